"""Build differently compiled copies of libgripper_b200.so for A/B runs (GMJ_LIB selects one at load time).
usage: python profiles/build_variants.py name1:-DFOO=1,-DBAR=2 name2:... ; output build/variants/lib_<name>.so
Only the FP32 N=5 / N=8 instantiations differ between variants; everything is compiled anyway (parallel nvcc)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gripper_mujoco_b200 import lib as gl  # noqa: E402

os.makedirs(os.path.join(ROOT, "build", "variants"), exist_ok=True)
for spec in sys.argv[1:]:
    name, _, flags = spec.partition(":")
    out = os.path.join(ROOT, "build", "variants", f"lib_{name}.so")
    gl.build(force=True, out=out, extra=[f for f in flags.split(",") if f])
    print("built", out)
