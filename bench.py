#!/usr/bin/env python
"""bench.py -- gripper env action-steps/sec on B200 (BASELINE.json metric, configs[1] at N=1).

  python bench.py --gpus N --steps K --warmup W            our arm (torchrun launches N ranks)
  python bench.py --impl reference --gpus N ...            reference arm: CPU pool on the host cores

One "step" = one action-step of every env of the rank: set_action + S physics substeps +
sensors + events + observation + reward + done (gmj_step: k_step set_action pass + k_step
physics pass), followed by the masked
auto-reset + object respawn of finished episodes (gmj_reset + gmj_spawn), as MjEnv.step/reset do.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "gripper env action-steps/sec"
UNIT = "action-steps/s"
EPISODE_CAP = 250            # rl/env/MjEnv.py:78 max_episode_steps


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


# ----------------------------------------------------------------------------- CPU arm
def calibrated_settings_for_cpu(N: int):
    """The CPU arm runs the reference's auto-calibration (timestep search, sim steps, sensor
    calibration) over the oracle itself, exactly like our arm does over the CUDA path."""
    from gripper_mujoco_b200.model import GripperParams
    from gripper_mujoco_b200.settings import Settings
    from oracle.calibrate import calibrated
    st = Settings()
    st.debug = False
    st.random_seed = 1234
    st.set_use_noise(True)
    return calibrated(GripperParams(num_segments=N), st)[1]


def run_cpu_pool(N, settings_struct, budget_s, cores=None):
    from oracle.pool import OraclePool
    pool = OraclePool(N, settings_struct, cores)
    pool.run(budget_s=0, max_steps=2)              # warm-up: build sets, first reset
    n, wall = pool.run(budget_s=budget_s)
    pool.close()
    return n, wall, pool.cores


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    st = calibrated_settings_for_cpu(args.segments)
    from oracle.pool import OraclePool
    pool = OraclePool(args.segments, st.pack())
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
    sample = (f"{args.steps} x {per_step_budget:.1f}s slices of the same random-action rollout "
              f"(one oracle env per core, {tot_n} action-steps total)")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(1, args.steps), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, st),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference MjClass via bind.so is not buildable here (libmujoco, Armadillo absent): "
                    "this arm times the CPU oracle port in a process pool on all host cores"}
    print(json.dumps(line), flush=True)


def workload_config(args, st):
    cfg = {(5, 4096): "configs[1]", (8, 16384): "configs[2]/[3] per-GPU shape", (24, 4096): "configs[4] sweep point"}.get(
        (args.segments, args.envs), "custom")
    return {"workload": f"{cfg}: {args.envs} envs/GPU, {args.segments}-segment fingers, set9_fullset-like 1750 objects, "
                        f"random discrete actions, auto-reset (cap {EPISODE_CAP})",
            "envs_per_gpu": args.envs, "segments": args.segments, "timestep": st.mujoco_timestep,
            "sim_steps_per_action": int(st.sim_steps_per_action), "noise": "on (Philox)",
            "l2": "flushed between timed steps (256 MiB write)"}


_REAL_STDOUT = [None]


def print_line(line: dict) -> None:
    """The one JSON line of the contract, on the process's original stdout."""
    text = json.dumps(line) + "\n"
    if _REAL_STDOUT[0] is None:
        sys.stdout.write(text); sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT[0], text.encode())


# ----------------------------------------------------------------------------- our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--envs", type=int, default=4096, help="envs per GPU")
    ap.add_argument("--segments", type=int, default=5)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-calibrate", action="store_true", help="skip the auto timestep / gauge calibration")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    args.warmup = max(args.warmup, 3)

    import numpy as np
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
    from gripper_mujoco_b200 import lib as gl
    from gripper_mujoco_b200.mjclass import MjClass
    from gripper_mujoco_b200.model import GripperParams
    from gripper_mujoco_b200.settings import Settings

    st = Settings()
    st.debug = False
    st.random_seed = 1234
    st.set_use_noise(True)
    n = args.envs
    t_cal = time.time()
    mj = MjClass(n, GripperParams(num_segments=args.segments), settings=st, device=local, precision="fp32",
                 env_id_offset=rank * n, calibrate=not args.no_calibrate)
    t_cal = time.time() - t_cal
    st = mj.set
    S = int(st.sim_steps_per_action)
    dev = mj.device
    gb = mj.batch
    n_act = mj.get_n_actions()
    K, W = args.steps, args.warmup
    gen = torch.Generator(device="cpu"); gen.manual_seed(5678 + rank)
    actions_host = torch.randint(0, n_act, (2 * (K + W), n), generator=gen, dtype=torch.int32).pin_memory()
    actions_dev = actions_host.to(dev)
    ep_steps = torch.zeros(n, dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    launches = [0]

    def env_step(a_dev):
        obs, rew, done, _ = gb.step(a_dev)
        ep_steps.add_(1)
        mask = (done != 0) | (ep_steps >= EPISODE_CAP)
        m8 = mask.to(torch.uint8)
        gb.reset(m8); gb.spawn_with_retry(0.01, float(np.pi / 2), m8)      # MjEnv.reset: reset -> spawn
        ep_steps.mul_((~mask).to(torch.int32))
        launches[0] += 5          # k_step (set_action pass), k_sort_scatter, k_step (combined), k_reset, k_spawn
        return obs, rew, done

    mj.batch.spawn_with_retry(0.01, float(np.pi / 2))
    s0 = gb.stats()
    # ---- device-resident timing: K steps, each bracketed by events, L2 flushed between them
    for k in range(W):
        env_step(actions_dev[k])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local); sampler.start()
    launches[0] = 0
    s1 = gb.stats()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True),
            torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for k in range(K):
        flush.fill_(k & 0xFF)
        evs[k][0].record()
        gb.step(actions_dev[W + k])
        evs[k][1].record()             # end of the dominant kernel (gmj_step)
        # the bookkeeping + masked reset / spawn of the step
        ep_steps.add_(1)
        mask = (gb.done != 0) | (ep_steps >= EPISODE_CAP)
        m8 = mask.to(torch.uint8)
        gb.reset(m8); gb.spawn_with_retry(0.01, float(np.pi / 2), m8)
        ep_steps.mul_((~mask).to(torch.int32))
        evs[k][2].record()
        launches[0] += 5          # k_step (set_action pass), k_sort_scatter, k_step (combined), k_reset, k_spawn
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
        a_dev.copy_(actions_host[K + W + k], non_blocking=True)
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
    sampler.stop_flag = True; sampler.join(timeout=2)
    h2d = n * 4
    d2h = n * gb.n_obs * 4 + n * 4 + n

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    sub = max(1, s2["substeps"] - s1["substeps"])
    nc = (s2["contacts"] - s1["contacts"]) / sub
    S_eff = sub / (n * K)                       # includes the 2S extra substeps of termination actions
    f_act = flops_per_action(args.segments, S_eff, nc)
    b_act = bytes_per_action(args.segments, gb.n_obs)
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
    traffic = None
    try:
        # ncu --set full captures exist for these launch shapes only (profiles/README.md)
        tfile = {(5, 4096): "k_step_traffic.json", (8, 16384): "r1_k_step_n8_traffic.json"}.get((args.segments, args.envs))
        if tfile:
            with open(os.path.join(ROOT, "profiles", tfile)) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
    except Exception:  # noqa: BLE001
        pass
    roofline = {"bound": "fp32", "kernel": (f"k_step<float,{args.segments}>" if args.segments <= 8 else f"k_step_wide<float,{args.segments}>"), "achieved": ach_tf, "peak": fp32_peak,
                "unit": "TFLOP/s", "frac": ach_tf / fp32_peak, "traffic": traffic, "peak_source": peak_src,
                "kernel_ms": kern_ms, "flops_per_action_step": f_act, "mean_contacts_per_substep": nc,
                "substeps_per_action_step": S_eff,
                "hbm": {"achieved": ach_gb, "peak": hbm_peak, "unit": "GB/s", "frac": ach_gb / hbm_peak,
                        "bytes_per_action_step": b_act},
                "why": "neither HBM nor tensor cores bind this path (SURVEY 8d): FP32 pipe / latency"}
    cpu = None
    if not args.no_cpu_baseline:
        ncpu, wall, cores = run_cpu_pool(args.segments, st.pack(), args.cpu_seconds)
        cpu = {"value": ncpu / wall, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{args.cpu_seconds:.0f} s of the same random-action rollout per core ({ncpu} action-steps), "
                         f"oracle FP64, one env per process"}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(args, st),
            "roofline": roofline, "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "clocks": sampler.result(), "gpu_launches": gpu_launches,
            "stats": {"unstable_resets": s2["unstable_resets"] - s0["unstable_resets"],
                      "contact_overflow": s2["contact_overflow"], "calibration_s": t_cal,
                      "highest_stable_timestep": getattr(mj, "highest_stable_timestep", None),
                      "torch_bookkeeping_launches_per_step": 6}}
    print_line(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
