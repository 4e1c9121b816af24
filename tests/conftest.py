import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def pytest_sessionfinish(session, exitstatus):
    """The FP32 / FP64 error maxima the parity tests saw (the bars in test_gpu_parity.TOL are set from this file)."""
    mod = sys.modules.get("test_gpu_parity")
    measured = getattr(mod, "MEASURED", None)
    if not measured:
        return
    import json
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "parity_measured.json"), "w") as f:
        json.dump({"columns": ["qpos", "qvel", "obs", "reward", "event_value"], "max_abs_error": measured}, f, indent=1, sort_keys=True)
