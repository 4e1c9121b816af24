"""Parity at CONTRACT size (BASELINE.json configs[1], configs[2], configs[4]): the FP32 product build through the C ABI
against the FP64 oracle, env by env, on the same seeded inputs -- reset noise, MjEnv's spawn policy (search with the
env's own Philox draws, 3 tries, fallback pose), random discrete actions including the termination action.

What FP32 can and cannot match: a trajectory is identical in its discrete events until a contact appears or
disappears one substep earlier or later than in FP64; from there the env is a different (equally valid) sample of the
same chaotic contact. So the bars are (1) done / event rows identical in all but a stated fraction of envs, (2) state
and sensors within a tight bar for the envs whose event history still agrees and whose object was spawned clear of the
fingers and has not been touched yet, at the horizon SURVEY 8(c) names, (3) quantile bars over ALL envs, (4) episode statistics (return, success
rate, per-event frequency) within confidence intervals over >= 4096 paired episodes. Every bar is <= 3x the value
measured on B200 (printed by the test; table in DESIGN.md).

Knife edge excluded from (1): within_XY_distance compares the object distance with 10 mm, and the 2 mm spawn / base
grids put ~10 % of the objects EXACTLY on 10 mm; the reference's own float cast decides those, FP32 state storage
decides them differently. Envs whose oracle distance is within 1e-6 m of the threshold are not counted in that column.
"""
import numpy as np
import pytest

from common import make_pair, objects, oracle_rollout

pytestmark = pytest.mark.gpu

DISC = list(range(15)) + [15 + 2 * k for k in range(21)]
CONT = [16 + 2 * k for k in range(21)]
COL_WITHIN_XY, COL_XY_DIST = 14, 56          # binary event row / -distance last value (EventTrack::vectorise order)
SEARCH = (0.01, float(np.pi / 2), 3, 10, 5)  # rl/env/MjEnv.py:1212-1244: +-10 mm, +-90 deg, 3 tries, fallback 10 mm / 5 deg
TIMING = {5: (6.857e-3, 30), 8: (3.225e-3, 63), 24: (5.6e-4, None)}      # auto-calibrated timesteps of bench.py's configs


def grasp_script(rng, T, n_act):
    """close 35-60 steps, tilt 5-12, palm forward 0-10, terminate (lift); random actions after that."""
    a = rng.integers(0, n_act, T).astype(np.int32)
    k1, k2, k3 = rng.integers(35, 61), rng.integers(5, 13), rng.integers(0, 11)
    s = [0] * k1 + [2] * k2 + [4] * k3 + [n_act - 1]
    a[:len(s)] = s[:T]
    return a


def paired_rollout(N, n, T, sample, seed, bias=False, scripted_every=0, stop_at_done=False):
    """n envs on the GPU (all launched), `sample` of them replayed by the oracle. Returns per-step device and oracle
    arrays of the sampled envs, the spawn codes and the device stats."""
    from gripper_mujoco_b200.batch import GripperBatch
    dt, S = TIMING[N]
    mdl, st = make_pair(N, dt, steps_per_action=S, noise=True)
    gb = GripperBatch(n, mdl, st, precision="fp32"); gb.set_objects(objects())
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, len(objects()), n).astype(np.int32)
    acts = rng.integers(0, gb.n_actions, (T, n)).astype(np.int32)
    if bias:        # half of the non-terminating actions close the fingers / lower the palm: more grasps per episode
        b = (rng.random((T, n)) < 0.4) & (acts != gb.n_actions - 1)
        acts[b] = rng.choice([1, 3, 5, 13], size=int(b.sum()))
    if scripted_every:        # every k-th env runs a jittered grasp-and-lift script, so that lifts / successful grasps occur
        for k in range(0, n, scripted_every):
            acts[:, k] = grasp_script(rng, T, gb.n_actions)
    gb.reset(); gb.spawn_with_retry(SEARCH[0], SEARCH[1], None, idx, tries=SEARCH[2], fallback_xy_mm=SEARCH[3], fallback_rot_deg=SEARCH[4])
    spawned = gb.spawned.cpu().numpy()
    if sample < n:   # stratified: every (n / sample)-th env in object-index order, so every object family is covered
        sel = np.sort(np.argsort(idx, kind="stable")[np.linspace(0, n - 1, sample).astype(int)])
    else:
        sel = np.arange(n)
    ep = int(gb.probe_field("episode")[0])
    ora = oracle_rollout(N, st, sel, ep, idx[sel], np.zeros((len(sel), 3), np.float32), acts[:, sel], search=SEARCH, stop_at_done=stop_at_done)
    dev = {k: [] for k in ("obs", "reward", "done", "events", "qpos", "qvel")}
    for t in range(T):
        o, r, d, e = (v.cpu().numpy() for v in gb.step(acts[t]))
        q, v = gb.state()
        for k, a in zip(dev, (o, r, d.astype(bool), e, q, v)):
            dev[k].append(a[sel])
    dev = {k: np.stack(v) for k, v in dev.items()}
    stats = gb.stats()
    gb.close()
    return dev, ora, spawned[sel], mdl.B, stats


def event_mismatch(dev_e, ora_e):
    """[n] bool: any discrete event column differs (knife-edge envs excused in the within_XY_distance column)."""
    neq = dev_e[:, DISC] != ora_e[:, DISC]
    edge = np.abs(-ora_e[:, COL_XY_DIST] - 10e-3) < 1e-6
    neq[:, DISC.index(COL_WITHIN_XY)] &= ~edge
    return neq.any(1)


# bars: <= 3x the B200 measurement printed by the test (see the module docstring for what each one means)
CONTRACT = {
    # N: (n launched, T, envs compared, early steps, bars)        measured on B200 (round 2) in the comment beside each
    5: (4096, 10, 4096, 3, dict(
        done_mismatch=5e-4,           # 0: every done flag of 40960 env-steps identical
        event_mismatch_early=4e-3,    # 1.5e-3 of the envs have left the oracle's event history after 3 steps
        event_mismatch=2e-2,          # 6.8e-3 after 10 steps (object_contact / fingerN_force one step early or late)
        early_qpos_rel=1e-4,          # 2.1e-5 ... 5.4e-5 between builds (FP32 rounding only); SURVEY 8c's bar is 1e-4
        early_obs=1.5e-4,             # 9.6e-6 ... 4.6e-5 between builds; SURVEY 8c's bar is 1e-3
        early_reward=1.5e-6,          # 4.3e-7
        early_evval=6e-4,             # 2.0e-4 (forces in N)
        median_rel=6e-7, p90_rel=3.5e-6, p99_rel=1.6e-3, p99_obs=9e-4)),      # 1.9e-7, 1.2e-6, 5.4e-4, 2.9e-4: all envs, all steps
    8: (16384, 6, 512, 3, dict(
        done_mismatch=4e-3,           # 0 of 3072 (bar = 2 envs)
        event_mismatch_early=4e-3,    # 0
        event_mismatch=6e-3,          # 2.0e-3 (1 env)
        early_qpos_rel=2e-5,          # 6.5e-6
        early_obs=5e-6,               # 1.7e-6
        early_reward=2e-7,            # 6.0e-8
        early_evval=6e-4,             # 5.7e-5 ... 2.4e-4 between builds that differ in FP32 rounding only (forces in N)
        median_rel=5.5e-7, p90_rel=1.2e-6, p99_rel=4e-5, p99_obs=2.5e-5)),    # 1.8e-7, 4.1e-7, 1.3e-5, 5.6e-6 ... 8.4e-6
}


@pytest.mark.parametrize("N", [5, 8])
def test_contract_trajectories_fp32_vs_oracle(N):
    n, T, sample, early, bar = CONTRACT[N]
    dev, ora, spawned, B, stats = paired_rollout(N, n, T, sample, seed=7)
    assert stats["unstable_resets"] == 0 and stats["contact_overflow"] == 0
    assert (spawned >= 1).all() and (spawned == 1).mean() > 0.9
    agree = np.ones(len(spawned), bool)              # event history identical so far
    m = dict(done_mismatch=0.0, event_mismatch_early=0.0, event_mismatch=0.0, early_qpos_rel=0.0, early_obs=0.0,
             early_reward=0.0, early_evval=0.0, median_rel=0.0, p90_rel=0.0, p99_rel=0.0, p99_obs=0.0)
    cols = np.zeros(len(DISC), int); excused = 0
    touched = np.zeros(len(spawned), bool)
    for t in range(T):
        mis = event_mismatch(dev["events"][t], ora["events"][t])
        neq = dev["events"][t][:, DISC] != ora["events"][t][:, DISC]
        edge = np.abs(-ora["events"][t][:, COL_XY_DIST] - 10e-3) < 1e-6
        excused += int((neq[:, DISC.index(COL_WITHIN_XY)] & edge).sum())
        neq[:, DISC.index(COL_WITHIN_XY)] &= ~edge
        cols += (neq & agree[:, None]).sum(0)             # first divergence of an env: which column
        agree &= ~mis
        dq = np.abs(dev["qpos"][t] - ora["qpos"][t])
        rel = dq[:, :B].max(1) / np.maximum(np.abs(ora["qpos"][t][:, :B]).max(1), 1e-3)
        dobs = np.abs(dev["obs"][t] - ora["obs"][t]).max(1)
        m["done_mismatch"] = max(m["done_mismatch"], (dev["done"][t] != ora["done"][t]).mean())
        m["event_mismatch"] = max(m["event_mismatch"], 1 - agree.mean())
        m["median_rel"] = max(m["median_rel"], np.median(rel)); m["p90_rel"] = max(m["p90_rel"], np.quantile(rel, 0.9))
        m["p99_rel"] = max(m["p99_rel"], np.quantile(rel, 0.99)); m["p99_obs"] = max(m["p99_obs"], np.quantile(dobs, 0.99))
        # an env is "free" until a finger or the palm touches its object (object_contact row, or any of the finger / palm
        # force readings): from the first touch on FP32 rounding decides how the contact develops
        oe = ora["events"][t]
        touched |= (oe[:, 7] > 0) | (np.abs(oe[:, [16, 18, 46, 48, 50]]).max(1) > 0)
        if t < early:
            calm = agree & (spawned == 1) & ~touched
            m["free_fraction"] = float(calm.mean())
            m["event_mismatch_early"] = max(m["event_mismatch_early"], 1 - agree.mean())
            m["early_qpos_rel"] = max(m["early_qpos_rel"], rel[calm].max()); m["early_obs"] = max(m["early_obs"], dobs[calm].max())
            # reward / event values: an excused knife-edge env differs by the within_XY_distance reward, leave it out
            edge = np.abs(-ora["events"][t][:, COL_XY_DIST] - 10e-3) < 1e-6
            m["early_reward"] = max(m["early_reward"], np.abs(dev["reward"][t] - ora["reward"][t])[calm & ~edge].max())
            m["early_evval"] = max(m["early_evval"], np.abs(dev["events"][t][:, CONT] - ora["events"][t][:, CONT])[calm].max())
    print(f"\nMEASURED contract N={N} n={n} T={T} sample={sample}: " + " ".join(f"{k}={v:.3g}" for k, v in m.items())
          + f" | constraint solves with 3+ rounds {stats['relaxed_solves']}, at the round cap {stats['capped_solves']} of {stats['substeps']} substeps")
    assert stats["relaxed_solves"] <= 2e-3 * stats["substeps"]          # measured 6.7e-4 (N = 5), 4.7e-4 (N = 8) per substep
    assert stats["capped_solves"] <= 1e-5 * stats["substeps"]
    print(f"   first-divergence column counts {({DISC[k]: int(c) for k, c in enumerate(cols) if c})}, knife-edge within_XY rows excused {excused} of {T * len(spawned)}")
    assert m.pop("free_fraction") > 0.5
    for k, v in m.items():
        assert v <= bar[k], (k, v, bar[k])
    # the numbers above say nothing about gripper <-> gripper contacts (not simulated): measure how
    # many of the compared env-steps are in a pose where the reference would have made one
    from gripper_mujoco_b200 import selfcheck
    mdl = make_pair(N, TIMING[N][0], steps_per_action=TIMING[N][1], noise=True)[0]
    over = sum(int((selfcheck.self_overlap(mdl, dev["qpos"][t], depth=1e-4)[0] > 0).sum()) for t in range(0, T, 4))
    print(f"   gripper self-overlap (M9 state): {over} of {len(range(0, T, 4)) * len(spawned)} sampled env-steps")
    assert over <= 1e-3 * len(range(0, T, 4)) * len(spawned)


def episode_stats(tr, n_act, T):
    """Per-env episode summary of a [T, n, ...] rollout that runs on after done: return up to the first done,
    length, the binary event rows at the done step (any step for envs that never finish)."""
    done = tr["done"].astype(bool)
    first = np.where(done.any(0), done.argmax(0), T - 1)
    alive = np.arange(T)[:, None] <= first[None, :]
    ret = (tr["reward"] * alive).sum(0)
    rows = tr["events"][first, np.arange(done.shape[1])][:, :15] > 0
    fired = ((tr["events"][:, :, :15] > 0) & alive[:, :, None]).any(0)
    return ret, first + 1, rows, fired, done.any(0)


STAT = {5: (4096, 84, 4), 8: (4096, 84, 8), 24: (256, 8, 0)}        # N: (episodes, step cap, every k-th env scripted)


@pytest.mark.parametrize("N", [5, 8, 24])
def test_statistical_protocol_paired_episodes(N):
    """SURVEY 8(c) last row: return, success rate and per-event frequency of >= 4096 episodes (>= 256 at N = 24) on
    the GPU against the oracle's on the same draws. The episodes are PAIRED, so the bars are on paired differences:
    mean return within 3 standard errors of the paired difference (plus 0.1 % of the return's spread), every binary
    event's episode frequency and the success rate within 3 binomial standard errors of the DISAGREEING pairs."""
    from gripper_mujoco_b200.settings import BINARY_NAMES
    n, T, every = STAT[N]
    dev, ora, spawned, B, stats = paired_rollout(N, n, T, n, seed=100 + N, bias=True, scripted_every=every, stop_at_done=True)
    assert stats["unstable_resets"] == 0
    rg, lg, rowg, firedg, fing = episode_stats(dev, None, T)
    ro, lo, rowo, firedo, fino = episode_stats(ora, None, T)
    d = rg - ro
    se = d.std(ddof=1) / np.sqrt(n)
    # where the return differs: which event row differs first
    w = np.where(np.abs(d) > 1e-3)[0]
    firstcol = {}
    for k in w:
        ne = np.where((dev["events"][:, k][:, DISC] != ora["events"][:, k][:, DISC]).any(1))[0]
        if len(ne):
            c = np.where(dev["events"][ne[0], k][DISC] != ora["events"][ne[0], k][DISC])[0]
            key = (DISC[c[0]], int(np.sign(dev["events"][ne[0], k][DISC[c[0]]] - ora["events"][ne[0], k][DISC[c[0]]])))
        else:
            key = ("none", 0)
        firstcol[key] = firstcol.get(key, 0) + 1
    print(f"\n   envs with a different return: {len(w)}; first differing event column (col, sign gpu - oracle): {firstcol}; mean diff there {d[w].mean() if len(w) else 0:.3f}")
    print(f"\nMEASURED episodes N={N} n={n} T={T}: finished {fino.mean():.3f} mean len {lo.mean():.1f} | return oracle {ro.mean():.4f} "
          f"gpu {rg.mean():.4f} paired diff {d.mean():.2e} +- {se:.2e} (spread {ro.std():.3f}) | same length {np.mean(lg == lo):.4f} | "
          f"envs with identical return (1e-3) {np.mean(np.abs(d) < 1e-3):.4f}")
    assert abs(d.mean()) <= 3 * se + 1e-3 * ro.std(), (d.mean(), se)
    assert np.mean(lg == lo) >= 0.995
    succ = BINARY_NAMES.index("successful_grasp"); lifted = BINARY_NAMES.index("lifted")
    worst = 0.0
    for k, nm in enumerate(BINARY_NAMES):
        fg, fo = firedg[:, k].mean(), firedo[:, k].mean()
        disagree = (firedg[:, k] != firedo[:, k]).mean()
        bound = 3 * np.sqrt(max(disagree, 1.0 / n) / n)          # sd of a paired frequency difference
        worst = max(worst, abs(fg - fo) / bound)
        if k in (succ, lifted) or abs(fg - fo) > 0:
            print(f"   {nm:24s} oracle {fo:.4f} gpu {fg:.4f} disagreeing pairs {disagree:.4f}")
        assert abs(fg - fo) <= bound, (nm, fg, fo, disagree)
        assert disagree <= 0.05, (nm, disagree)
    if every:
        assert firedo[:, lifted].mean() > 0.02 and firedo[:, BINARY_NAMES.index("object_contact")].mean() > 0.1, "protocol needs grasps"
    if N <= 8:          # how many of these env-steps are in the gripper <-> gripper contact state the path does not simulate (M9)
        from gripper_mujoco_b200 import selfcheck
        mdl = make_pair(N, TIMING[N][0], steps_per_action=TIMING[N][1], noise=True)[0]
        ts = list(range(3, T, 8))
        over = live = 0
        for t in ts:                         # only while the episode runs (the device keeps stepping a finished env)
            run = t < lg
            over += int(((selfcheck.self_overlap(mdl, dev["qpos"][t], depth=1e-4)[0] > 0) & run).sum()); live += int(run.sum())
        print(f"   gripper self-overlap (M9 state): {over} of {live} sampled env-steps of running episodes")
        assert over <= 1e-2 * live              # measured on B200: see DESIGN.md section 8
