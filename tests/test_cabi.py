"""The C-ABI library loads on a CPU-only machine, exports every symbol include/gmj.h declares (and
nothing is missing from the ctypes binding), and fails loudly -- not silently -- without a GPU."""
import ctypes as C
import os
import re
import subprocess

import pytest
import torch

from gripper_mujoco_b200 import lib as gl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "gmj.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gmj_\w+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    so = gl.build()
    names = _declared()
    assert len(names) >= 19
    assert sorted(gl.SYMBOLS) == names, "lib.py SYMBOLS and include/gmj.h disagree"
    out = subprocess.check_output(["nm", "-D", "--defined-only", so], text=True)
    exported = set(re.findall(r" T (gmj_\w+)", out))
    assert set(names) <= exported, set(names) - exported
    L = gl.lib()
    for n in names:
        assert getattr(L, n) is not None


def test_struct_layouts_match_header():
    """ctypes mirrors have the sizes the C compiler gives the structs of include/gmj_types.h."""
    from gripper_mujoco_b200.model import GmjModel, GmjObject
    from gripper_mujoco_b200.settings import GmjSettings
    prog = ('#include <stdio.h>\n#include "gmj_types.h"\nint main(){printf("%zu %zu %zu %zu\\n", sizeof(GmjModel), '
            'sizeof(GmjObject), sizeof(GmjSettings), sizeof(GmjStats));return 0;}')
    exe = "/tmp/gmj_sizes"
    subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", exe], input=prog, text=True, check=True)
    sizes = [int(x) for x in subprocess.check_output([exe], text=True).split()]
    assert sizes == [C.sizeof(GmjModel), C.sizeof(GmjObject), C.sizeof(GmjSettings), C.sizeof(gl.GmjStats)]


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    from gripper_mujoco_b200.batch import GripperBatch
    from gripper_mujoco_b200.model import build_model
    from gripper_mujoco_b200.settings import Settings
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        GripperBatch(1)
    h = C.c_void_p()
    m, s = build_model().to_struct(), Settings().pack()
    rc = gl.lib().gmj_create(C.byref(m), C.byref(s), 1, 0, 0, 0, C.byref(h))
    assert rc != 0 and b"no CUDA device" in gl.lib().gmj_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "gripper_mujoco_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle|#include\s+\".*oracle", txt, flags=re.M), f
