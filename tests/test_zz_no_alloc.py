"""gmj_probe_host_noalloc (include/gmj.h): the allocation-free variant of gmj_probe_host, whose device scratch is a
static array inside the library. Collected last on purpose. Through the C ABI: two
calls in a row equal the existing entry point's output exactly (same kernels, same values, FP64 copies), a stale
scratch from an earlier, larger call does not leak into a later one, and a handle beyond the static maximum is
rejected with a clear error."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _batch(n, N=5):
    from common import make_pair, objects
    from gripper_mujoco_b200.batch import GripperBatch
    mdl, st = make_pair(N, 4e-3, steps_per_action=20, noise=True, seed=5)
    gb = GripperBatch(n, mdl, st, precision="fp32")
    gb.set_objects(objects())
    return gb


def test_noalloc_probe_equals_probe_on_two_calls_in_a_row():
    from gripper_mujoco_b200.lib import GMJ_PROBE_LEN, GP
    big, gb = _batch(300), _batch(37)
    rng = np.random.default_rng(0)
    for b in (big, gb):
        b.reset(); b.spawn_with_retry(0.01, 0.5)
    big.step(rng.integers(0, big.n_actions - 1, big.n_envs).astype(np.int32))
    stale = big.probe_noalloc()                      # fills 300 rows of the shared scratch
    assert np.array_equal(stale, big.probe())
    out = np.full((gb.n_envs, GMJ_PROBE_LEN), np.nan)
    for call in range(2):                             # two calls in a row, state changing in between
        gb.step(rng.integers(0, gb.n_actions - 1, gb.n_envs).astype(np.int32))
        ref = gb.probe()
        got = gb.probe_noalloc(out)
        assert got is out and np.isfinite(got).all()
        assert np.array_equal(got, ref), f"call {call}: max diff {np.abs(got - ref).max()}"
        assert (got[:, GP["time"]] > 0).all() and got[:, GP["gauge0"]].std() >= 0
    # raw C-ABI call with a caller-owned buffer
    raw = np.zeros((gb.n_envs, GMJ_PROBE_LEN))
    assert gb.L.gmj_probe_host_noalloc(gb.h, raw.ctypes.data_as(C.c_void_p)) == 0
    assert np.array_equal(raw, gb.probe())
    big.close(); gb.close()


def test_noalloc_probe_rejects_sizes_beyond_the_static_scratch():
    from gripper_mujoco_b200.lib import GMJ_NOALLOC_MAX_ENVS, GMJ_PROBE_LEN, GmjError
    gb = _batch(GMJ_NOALLOC_MAX_ENVS + 1)
    with pytest.raises(GmjError, match="GMJ_NOALLOC_MAX_ENVS"):
        gb.probe_noalloc()
    assert gb.probe().shape == (GMJ_NOALLOC_MAX_ENVS + 1, GMJ_PROBE_LEN)     # the allocating entry point still serves it
    gb.close()
