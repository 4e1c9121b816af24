#!/usr/bin/env python
"""bench.py -- gripper env action-steps/sec on B200 (BASELINE.json metric, configs[1] at N=1).

  python bench.py --gpus N --steps K --warmup W            our arm (torchrun launches N ranks)
  python bench.py --impl reference --gpus N ...            reference arm: CPU pool on the host cores

One "step" = one action-step of every env of the rank: set_action + S physics substeps + sensors + events +
observation + reward + done (gmj_step: set_action / collect pass, sort, combined substep launch), followed by the
masked auto-reset + object respawn of finished episodes (gmj_reset + gmj_spawn with MjEnv's retry policy), as
MjEnv.step / reset do.

Both arms run the SAME configuration: the auto-calibration (timestep search, sim steps per action, gauge / wrist
calibration -- MjClass::configure_settings) is done once per finger model by whichever arm runs first on the box and
shared through a file in the temp directory, so `config` (timestep, sim_steps_per_action ...) is identical.

The line also carries `extra.configs`: the other BASELINE.json shapes measured in the same invocation with the same
code path -- configs[2]/[3] per-GPU shape (16384 envs, 8-segment fingers, contact-heavy scripted grasp-and-lift
prefix; with --gpus 8 that is configs[3], gather included) and one configs[4] point (24-segment fingers).
"""
from __future__ import annotations

import argparse
import json
import os
import pickle
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "gripper env action-steps/sec"
UNIT = "action-steps/s"
EPISODE_CAP = 250            # rl/env/MjEnv.py:78 max_episode_steps
PREROLL = 30                 # untimed action steps before the warm-up: episodes spread over their phases (steady state)
# SURVEY 8d "contact-heavy" workload: scripted prefix of every episode, then random actions. Action indices under the
# header-default action table: 0 / 1 prismatic close / open, 2 / 3 tilt in / out, 4 / 5 palm forward / back,
# 10 / 11 base down / up, 14 termination.
CONTACT_PREFIX = [0] * 30 + [2] * 10 + [4] * 20 + [11] * 15


# ----------------------------------------------------------------------------- algorithmic work
def flops_per_action(N: int, S: float, nc: float, nb: int = 1) -> float:
    B, nv = nb + 7 + 3 * N, nb + 7 + 3 * N + 6
    P = nb * (nb + 1) // 2 + 3 * sum(nb + k for k in range(1, N + 3)) + (nb + 1)
    Q = sum(d * d for d in range(1, nb + 1)) + 3 * sum((nb + k) ** 2 for k in range(1, N + 3)) + (nb + 1) ** 2
    f_tree = 500 * (B + 1) + 14 * P + 2 * Q + 12 * nv
    f_col = 12 * (6 * N + 8) + 350 * nc
    R = 3 * nc                                   # limit rows not counted (conservative), welds = 0 rows
    dbar = nb + 2 + N / 2
    f_con = R * 4 * (P / 3 + nb + 30) + 2 * R * R * (dbar + 6) + 40 * R * R + 2 * R * nv
    return S * (f_tree + f_col + f_con) + 6000


def bytes_per_action(N: int, n_obs: int, nb: int = 1) -> float:
    B = nb + 7 + 3 * N
    nq, nv = B + 7, B + 6
    n_state = nq + 2 * nv + 10 + 28 + 7 * 12 + 5 + 3 * 36 + 30 + 24
    return 2 * 4 * n_state + 4 * n_obs + 10


# ----------------------------------------------------------------------------- clocks sampling
class ClockSampler(threading.Thread):
    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None

    def run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            names = {pynvml.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                     pynvml.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                     pynvml.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                     pynvml.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                     pynvml.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake"}
            while not self.stop_flag:
                self.samples.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
                time.sleep(0.05)
        except Exception as ex:  # noqa: BLE001
            self.reasons.add(f"nvml_unavailable:{type(ex).__name__}")

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# ----------------------------------------------------------------------------- shared calibration (same config in both arms)
PRESET = [None]          # --preset: a TrainingManager reward / sensor preset on top of the header defaults


def base_settings():
    from gripper_mujoco_b200.settings import Settings
    st = Settings()
    if PRESET[0]:
        # rl/TrainingManager.py configure_env: wipe -> apply_env_settings -> create_reward_function, with DISCRETE actions
        # (one action per step, the fused gmj_step; the continuous variant sets every action per step through
        # gmj_set_action and is covered by tests/test_presets.py)
        from gripper_mujoco_b200 import training_presets as tp

        class _Env:
            class params:
                max_episode_steps = EPISODE_CAP

            class mj:
                pass
        _Env.mj.set = st
        _Env.mj.set_sensor_noise_and_normalisation_to = staticmethod(lambda v: (st.set_use_normalisation(v), st.set_use_noise(v)))
        _Env.mj.yield_load = staticmethod(lambda: 3.458)
        cfg = tp.preset(PRESET[0])
        cfg["cpp"]["continous_actions"] = False
        cfg["cpp"]["use_termination_action"] = True
        tp.configure_env(_Env, cfg)
    st.debug = False
    st.random_seed = 1234
    st.set_use_noise(True)
    return st


def calib_path(N: int) -> str:
    tag = f"_{PRESET[0]}" if PRESET[0] else ""
    return os.path.join(tempfile.gettempdir(), f"gmj_bench_calibration_N{N}{tag}.pkl")


def shared_calibration(N: int, how: str, device: int = 0):
    """(settings, source, seconds): the calibrated Settings of the N-segment model. The first arm to run on a box does
    the auto-calibration (`how` = "cuda": MjClass over the CUDA path, "oracle": the CPU oracle) and leaves it in the temp
    directory; later arms and ranks load it, so every arm and rank runs the same timestep / sim steps / sensor scales."""
    path = calib_path(N)
    if os.path.exists(path):
        with open(path, "rb") as f:
            d = pickle.load(f)
        return d["settings"], d["source"] + " (shared file)", 0.0
    from gripper_mujoco_b200.model import GripperParams
    t0 = time.time()
    if how == "cuda":
        from gripper_mujoco_b200.mjclass import MjClass
        mj = MjClass(1, GripperParams(num_segments=N), settings=base_settings(), device=device, precision="fp32", calibrate=True)
        st = mj.set.copy()
        hst = getattr(mj, "highest_stable_timestep", None)
        mj.close()
        src = "auto-calibration over the CUDA path (FP32)"
    else:
        from oracle.calibrate import calibrated
        _, st, info = calibrated(GripperParams(num_segments=N), base_settings())
        hst = info.get("highest_stable_timestep")
        src = "auto-calibration over the CPU oracle (FP64)"
    dt = time.time() - t0
    tmp = path + f".{os.getpid()}.tmp"
    with open(tmp, "wb") as f:
        pickle.dump({"settings": st, "source": src, "highest_stable_timestep": hst}, f)
    os.replace(tmp, path)
    return st, src, dt


def workload_config(args, N, envs, st, workload):
    cfg = {(5, 4096): "configs[1]", (8, 16384): "configs[2]/[3] per-GPU shape", (24, 4096): "configs[4] sweep point",
           (24, 1024): "configs[4] sweep point"}.get((N, envs), "custom")
    acts = "random discrete actions" if workload == "random" else \
        "contact-heavy: scripted prefix per episode (close x30, tilt x10, palm x20, base-Z up x15) then random discrete actions"
    if PRESET[0]:
        acts += f", TrainingManager preset '{PRESET[0]}' (sensors, noise, reward function; discrete actions + termination action)"
    return {"workload": f"{cfg}: {envs} envs/GPU, {N}-segment fingers, set9_fullset-like 1750 objects, {acts}, "
                        f"auto-reset (cap {EPISODE_CAP}), spawn_into_scene +-10 mm / +-90 deg with MjEnv's retry policy",
            "envs_per_gpu": envs, "segments": N, "timestep": st.mujoco_timestep,
            "sim_steps_per_action": int(st.sim_steps_per_action), "noise": "on (Philox)",
            "l2": "flushed between timed steps (256 MiB write)"}


# ----------------------------------------------------------------------------- CPU arm
def run_cpu_pool(N, settings_struct, budget_s, cores=None):
    from oracle.pool import OraclePool
    pool = OraclePool(N, settings_struct, cores)
    pool.run(budget_s=0, max_steps=2)              # warm-up: build sets, first reset
    n, wall = pool.run(budget_s=budget_s)
    pool.close()
    return n, wall, pool.cores


CPU_KIND_NOTE = ("reference MjClass via bind.so is not buildable here (libmujoco, Armadillo absent): this arm times the CPU "
                 "oracle port, FP64, one env per process on every host core, in its timed-baseline mode (direct active-set "
                 "solve, tree-sparse skyline Cholesky; nv = B + 6, one live object). The reference publishes ~320 "
                 "action-steps/s = ~50 us per substep on one desktop thread at nv = 152 (src/test.cpp:181)")


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    N = args.segments
    st, src, _ = shared_calibration(N, "oracle")
    from oracle.pool import OraclePool
    pool = OraclePool(N, st.pack())
    cores = pool.cores
    per_step_budget = max(0.5, min(10.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        pool.run(budget_s=per_step_budget)
    tot_n, tot_t = 0, 0.0
    for _ in range(args.steps):
        n, wall = pool.run(budget_s=per_step_budget)
        tot_n += n; tot_t += wall
    pool.close()
    v = tot_n / tot_t
    S = int(st.sim_steps_per_action)
    us_sub = 1e6 * cores / (v * S * 34.0 / 30.0)          # ~34/30: the termination action's extra substeps
    sample = (f"{args.steps} x {per_step_budget:.1f}s slices of the same random-action rollout "
              f"(one oracle env per core, {tot_n} action-steps total; ~{us_sub:.0f} us per substep per core)")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(1, args.steps), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, N, args.envs, st, args.workload),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "calibration_source": src, "note": CPU_KIND_NOTE}
    print(json.dumps(line), flush=True)


_REAL_STDOUT = [None]


def print_line(line: dict) -> None:
    """The one JSON line of the contract, on the process's original stdout."""
    text = json.dumps(line) + "\n"
    if _REAL_STDOUT[0] is None:
        sys.stdout.write(text); sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT[0], text.encode())


# ----------------------------------------------------------------------------- our arm: one configuration
def run_config(args, N, envs, K, W, workload, rank, world, local, headline):
    """Measure one configuration on this rank's GPU; rank 0 returns the result dict (others None)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from gripper_mujoco_b200 import lib as gl
    from gripper_mujoco_b200.mjclass import MjClass
    from gripper_mujoco_b200.model import GripperParams

    if rank == 0:
        st, cal_src, t_cal = shared_calibration(N, "cuda", local)
    if world > 1:
        dist.barrier()
    if rank != 0:
        st, cal_src, t_cal = shared_calibration(N, "cuda", local)
    n = envs
    mj = MjClass(n, GripperParams(num_segments=N), settings=st.copy(), device=local, precision="fp32",
                 env_id_offset=rank * n, calibrate=False)
    st = mj.set
    dev = mj.device
    gb = mj.batch
    n_act = mj.get_n_actions()
    gen = torch.Generator(device="cpu"); gen.manual_seed(5678 + rank)
    total = PREROLL + 2 * (K + W)
    actions_host = torch.randint(0, n_act, (total, n), generator=gen, dtype=torch.int32).pin_memory()
    actions_dev = actions_host.to(dev)
    prefix = torch.tensor(CONTACT_PREFIX, dtype=torch.int32, device=dev)
    ep_steps = torch.zeros(n, dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    launches = [0]
    a_work = torch.empty(n, dtype=torch.int32, device=dev)

    def pick_actions(a_rand):
        """Random actions, or the scripted contact-heavy prefix while the episode is younger than the prefix."""
        if workload == "random":
            return a_rand
        k = ep_steps.clamp(max=len(CONTACT_PREFIX) - 1).to(torch.int64)
        torch.where(ep_steps < len(CONTACT_PREFIX), prefix[k], a_rand, out=a_work)
        return a_work

    def bookkeeping():
        """MjEnv.step's end of episode: done or 250 steps -> reset -> _spawn_object (3 tries + fallback) -> the first
        observation of the new episode (MjEnv.reset's return value) into the obs rows of those envs: three launches."""
        ep_steps.add_(1)
        m8 = gb.reset_done(gb.done, EPISODE_CAP)
        gb.spawn_with_retry(0.01, float(np.pi / 2), m8)
        gb.observe_masked(m8)
        ep_steps.mul_(1 - m8.to(torch.int32))
        launches[0] += 6       # ours: k_step (set_action pass), k_sort_scatter, k_step (combined), k_reset, k_spawn, k_observe_masked

    def env_step(a_dev):
        out = gb.step(pick_actions(a_dev))
        bookkeeping()
        return out[0], out[1], out[2]

    gb.spawn_with_retry(0.01, float(np.pi / 2))
    s0 = gb.stats()
    for k in range(PREROLL + W):                      # untimed: pre-roll to the steady-state episode mix, then warm-up
        env_step(actions_dev[k])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local) if headline else None
    if sampler:
        sampler.start()
    launches[0] = 0
    s1 = gb.stats()
    # ---- device-resident timing: K steps, each bracketed by events, L2 flushed between them
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True),
            torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for k in range(K):
        flush.fill_(k & 0xFF)
        a = pick_actions(actions_dev[PREROLL + W + k])
        evs[k][0].record()
        gb.step(a)
        evs[k][1].record()             # end of gmj_step (the dominant kernel + its set_action / sort passes)
        bookkeeping()
        evs[k][2].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    step_ms = sum(e[0].elapsed_time(e[2]) for e in evs)
    kern_ms = sum(e[0].elapsed_time(e[1]) for e in evs) / K
    s2 = gb.stats()
    gpu_launches = launches[0]
    t = torch.tensor([step_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = world * n * K / (total_ms * 1e-3)

    # ---- end to end through the public API with host buffers (pinned H2D actions, D2H results)
    obs_h = torch.empty((n, gb.n_obs), dtype=torch.float32).pin_memory()
    rew_h = torch.empty(n, dtype=torch.float32).pin_memory()
    done_h = torch.empty(n, dtype=torch.uint8).pin_memory()
    a_dev = torch.empty(n, dtype=torch.int32, device=dev)
    gatherer = None
    if world > 1:          # the path's only exchange: transitions of every shard to the learner rank (NCCL gather)
        from gripper_mujoco_b200.sharding import TransitionGather, shard_range
        gatherer = TransitionGather(shard_range(world * n, rank, world), gb.n_obs, dev, dst=0)
        for _ in range(3):                 # communicator set-up and first-use costs stay outside the timed region
            gatherer.gather(gb.obs, gb.reward, gb.done)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(K):
        a_dev.copy_(actions_host[PREROLL + K + W + k], non_blocking=True)
        obs, rew, done = env_step(a_dev)
        if gatherer is not None:
            gatherer.gather(obs, rew, done)
        obs_h.copy_(obs, non_blocking=True); rew_h.copy_(rew, non_blocking=True); done_h.copy_(done, non_blocking=True)
        torch.cuda.current_stream().synchronize()      # the learner consumes the transition on the host
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.barrier()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * n * K / (float(t.item()) * 1e-3)
    clocks = None
    if sampler:
        sampler.stop_flag = True; sampler.join(timeout=2)
        clocks = sampler.result()
    h2d = n * 4
    d2h = n * gb.n_obs * 4 + n * 4 + n
    n_obs = gb.n_obs
    mj.close()
    del flush
    torch.cuda.empty_cache()
    if rank != 0:
        return None

    sub = max(1, s2["substeps"] - s1["substeps"])
    nc = (s2["contacts"] - s1["contacts"]) / sub
    S_eff = sub / (n * K)                       # includes the 2S extra substeps of termination actions
    f_act = flops_per_action(N, S_eff, nc)
    b_act = bytes_per_action(N, n_obs)
    try:
        gl.build_fma_peak() if not os.path.exists(gl.FMA_SO) else None
        fp32_peak = gl.fma_peak_tflops(local)
        peak_src = "measured (csrc/fma_peak.cu, FFMA chains, this run)"
    except Exception:  # noqa: BLE001
        fp32_peak, peak_src = 74.4, "nominal 148 SM x 128 lanes x 2 x 1.965 GHz (microbenchmark failed)"
    hbm_peak = 6540.5
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            hbm_peak = float(json.load(f)["hbm_gbs"])
    except Exception:  # noqa: BLE001
        pass
    ach_tf = f_act * n / (kern_ms * 1e-3) * 1e-12
    ach_gb = b_act * n / (kern_ms * 1e-3) * 1e-9
    traffic, traffic_src = None, None
    try:
        # dram bytes of one launch come from a committed `ncu --set full` capture of the same launch shape, NOT from this run
        tfile = {(5, 4096): "r2_k_step_n5_traffic.json", (8, 16384): "r2_k_step_n8_traffic.json"}.get((N, envs))
        if tfile:
            with open(os.path.join(ROOT, "profiles", tfile)) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
            traffic_src = f"profiles/{tfile} (ncu capture of this launch shape, not measured in this run)"
    except Exception:  # noqa: BLE001
        pass
    roofline = {"bound": "fp32", "kernel": (f"k_step<float,{N}>" if N <= 8 else f"k_step_wide<float,{N}>"),
                "achieved": ach_tf, "peak": fp32_peak, "unit": "TFLOP/s", "frac": ach_tf / fp32_peak, "traffic": traffic,
                "traffic_source": traffic_src, "peak_source": peak_src,
                "kernel_ms": kern_ms, "flops_per_action_step": f_act, "mean_contacts_per_substep": nc,
                "substeps_per_action_step": S_eff,
                "hbm": {"achieved": ach_gb, "peak": hbm_peak, "unit": "GB/s", "frac": ach_gb / hbm_peak,
                        "bytes_per_action_step": b_act},
                "why": "neither HBM nor tensor cores bind this path (SURVEY 8d): FP32 pipe / latency"}
    return {"value": value, "ms_per_step": total_ms / K, "config": workload_config(args, N, envs, st, workload),
            "roofline": roofline, "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "clocks": clocks, "gpu_launches": gpu_launches, "steps": K, "warmup": W,
            "stats": {"unstable_resets": s2["unstable_resets"] - s0["unstable_resets"], "contact_overflow": s2["contact_overflow"],
                      "relaxed_solves_per_substep": (s2["relaxed_solves"] - s1["relaxed_solves"]) / sub,
                      "capped_solves": s2["capped_solves"] - s1["capped_solves"],
                      "calibration_s": t_cal, "calibration_source": cal_src, "preroll_steps": PREROLL,
                      "torch_bookkeeping_launches_per_step": 4},
            "settings": st}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--envs", type=int, default=4096, help="envs per GPU")
    ap.add_argument("--segments", type=int, default=5)
    ap.add_argument("--workload", default="random", choices=["random", "contact"],
                    help="random discrete actions, or the contact-heavy scripted prefix of SURVEY 8d")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the extra.configs sub-results")
    ap.add_argument("--preset", default=None, choices=["sensor_mixed_v1", "sparse", "termination_action_v1", "MAT_liftonly"],
                    help="run under a TrainingManager preset (rl/TrainingManager.py:804-1268) instead of the header defaults")
    args = ap.parse_args()
    PRESET[0] = args.preset
    if args.impl == "reference":
        return reference_arm(args)
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/tmp/nccl_%h_%p.log")   # keep stdout to the one JSON line
        # NCCL still writes its version banner to fd 1 when the communicator comes up: send everything that is not
        # the result line to stderr and keep the real stdout for print_line()
        sys.stdout.flush()
        _REAL_STDOUT[0] = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    K, W = args.steps, args.warmup
    head = run_config(args, args.segments, args.envs, K, W, args.workload, rank, world, local, headline=True)
    extras = []
    default_run = (args.segments, args.envs, args.workload, args.preset) == (5, 4096, "random", None)
    if not args.no_extras and default_run:
        # BASELINE.json configs[2] / [3]: 16384 envs per GPU, 8-segment fingers, contact-heavy grasp-and-lift episodes
        # (at --gpus 8 this IS configs[3]: 131072 envs, transitions gathered to rank 0 inside e2e); configs[4]: one point
        # of the 24-segment sweep; the same N = 8 shape under the headline's random-action workload beside it
        for N, envs, k, w, wl in ((8, 16384, 8, 3, "contact"), (8, 16384, 8, 3, "random"), (24, 4096, 3, 3, "random")):
            r = run_config(args, N, envs, k, w, wl, rank, world, local, headline=False)
            if r is not None:
                r.pop("settings"); r.pop("clocks")
                r.update({"metric": METRIC, "unit": UNIT, "n_gpus": world})
                extras.append(r)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    st = head.pop("settings")
    cpu = None
    if not args.no_cpu_baseline:
        ncpu, wall, cores = run_cpu_pool(args.segments, st.pack(), args.cpu_seconds)
        v = ncpu / wall
        us_sub = 1e6 * cores / (v * head["roofline"]["substeps_per_action_step"])
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{args.cpu_seconds:.0f} s of the same random-action rollout per core ({ncpu} action-steps, "
                         f"~{us_sub:.0f} us per substep per core), oracle FP64 in its timed-baseline mode, one env per process",
               "note": CPU_KIND_NOTE}
    line = {"metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": head["config"],
            "roofline": head["roofline"], "cpu_baseline": cpu, "e2e": head["e2e"],
            "clocks": head["clocks"], "gpu_launches": head["gpu_launches"], "stats": head["stats"],
            "extra": {"configs": extras}}
    print_line(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
