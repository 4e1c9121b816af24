// gmj_step.cuh -- the step kernels and their launchers. Each (precision, N) instantiation is compiled in its own
// translation unit (gmj_inst.cu with -DINST_T / -DINST_N [/ -DINST_WIDE]) so the build runs in parallel.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <string>
#include <vector>
#include "../../include/gmj.h"
#include "gmj_wide.cuh"

// environments (warps) per block of the FP32 step kernel, one block per SM: as many as shared memory
// holds (N = 5: 12 x 16.1 KB, 14 measured no faster, N = 8: 12 x 18.8 KB of the 227 KB)
// envs are dealt to blocks in order of expected load so that the warps of a lockstep block finish each
// phase together (ncu: stall_barrier was the top stall with arrival order)
#ifndef GMJ_LIFT_WPB
#define GMJ_LIFT_WPB 12
#endif
#ifndef GMJ_LIFT_SMALL
#define GMJ_LIFT_SMALL 8            // lift-block warps when the launch is <= 3 rounds of full blocks (A/B: 6 and 4 slower)
#endif
#ifndef GMJ_SORT_DIV
#define GMJ_SORT_DIV 3
#endif
#define GMJ_SORT_ROWS ((GMJ_RMAX + GMJ_SORT_DIV) / GMJ_SORT_DIV)
#define GMJ_SORT_BUCKETS (3 * GMJ_SORT_ROWS)
// load key of an env: MPR passes of its last substep first (a pass costs the whole warp ~1.5 k instructions, whichever
// lanes run it), constraint rows second -- the warps of a lockstep block then agree phase by phase, not only in total
#ifndef GMJ_SORT_MPR
#define GMJ_SORT_MPR 1
#endif
template <typename E> __device__ __forceinline__ int sort_key_of(const E& e) {
  int rows = e.nrow_prev / GMJ_SORT_DIV; if (rows > GMJ_SORT_ROWS - 1) rows = GMJ_SORT_ROWS - 1;
  int mpr = GMJ_SORT_MPR ? e.mpr_prev : 0; if (mpr < 0) mpr = 0; if (mpr > 2) mpr = 2;
  return mpr * GMJ_SORT_ROWS + rows;
}
#ifndef GMJ_BLOCKS_PER_SM
#define GMJ_BLOCKS_PER_SM 1
#endif
#ifndef GMJ_WPB_F32_N5
#define GMJ_WPB_F32_N5 12
#endif
#ifndef GMJ_WPB_F32_N8
#define GMJ_WPB_F32_N8 12
#endif

#define GMJ_MAX_DEVICES 64
void gmj_set_error(const std::string& msg);      // gmj_api.cu (thread-local last error)
#define CK(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) { gmj_set_error(std::string(#x) + ": " + cudaGetErrorString(_e)); return 1; } } while (0)
#define FAIL(msg) do { gmj_set_error(msg); return 1; } while (0)

struct GmjHandle {
  int n_envs, device, precision, N, B;
  int wide;              // 1: B > 32, one env per 96-thread block (gmj_wide.cuh), state EnvState<T, 96>, model DevModel<T, 96>
  const void* model_s;   // the DevModelS<T> base of *model
  uint32_t env_id_offset;
  void* envs;            // EnvState<T>[n_envs]
  void* model;           // DevModel<T>
  DevSettings* settings; // device
  DevSettings hs;        // host copy
  GmjObject* objects; int n_objects;
  int* err_flag;
  int* env_list; int* env_count;   // envs that owe the 2S termination-lift substeps this step
  int* perm; int* hist;            // load-sorted order of the other envs; hist[0..16) counts, hist[16..32) cursors
  unsigned long long* stats_dev;
  double* scratch; size_t scratch_len;
  GmjModel hm;
  size_t smem_bytes;
  double* asin_tab;      // device table for the stepper (DevSettings::asin_tab)
  uint8_t* spawn_ok;     // per env: 1 when the last gmj_spawn placed its object (spawn_into_scene's return value)
  void* taps;            // T[n_envs][GMJ_MAXTAP][GMJ_TAP_HDR + 3 N]: sensor tap records of the running step (gmj_tail.cuh)
  int n_sm;              // SMs of the device
  int balance_sm;        // SM count when the combined launch balances envs per block over the SMs, 0 = full blocks
};

// ------------------------------------------------------------------------------------------ kernels
enum { STEP_SET_ACTION = 1, STEP_SUBSTEPS = 2, STEP_TAIL = 4, STEP_OUTPUT = 8, STEP_SENSORS = 16,
       STEP_COLLECT = 32,   // set_action only: append envs that owe extra substeps (termination lift) to env_list
       STEP_LIST = 64,      // run on the envs of env_list[0 .. *env_count) instead of all envs
       STEP_COMBINED = 128 }; // blocks [0, list_blocks) take env_list (owed lift substeps first), the rest all other envs

template <typename T> __device__ __forceinline__ void copy_words(void* dst, const void* src, int nbytes, int lane) {
  const uint32_t* s = (const uint32_t*)src; uint32_t* d = (uint32_t*)dst;
  for (int i = lane; i < nbytes / 4; i += 32) d[i] = s[i];
}

// One warp per environment, WPB environments per block. flags select the phases so the same kernel
// serves gmj_step (all), gmj_set_action, gmj_action_step, gmj_substeps and gmj_observe.
// The kernel is instruction-fetch bound (ncu: stall_no_inst dominates, the step is ~25 k SASS
// instructions of mostly straight-line code against a 32 KB L1.5 I-cache), so the warps of a block
// are kept in lockstep with a block barrier per substep: one fetch stream then feeds WPB warps.
template <typename T, int N> struct StepCfg {
  static constexpr int WPB = sizeof(T) == 4 ? (N <= 5 ? GMJ_WPB_F32_N5 : GMJ_WPB_F32_N8) : 4;
  static constexpr bool MODEL_SMEM = sizeof(T) == 4 && N > 5;        // block copy of DevModel in shared memory (k_step)
};

template <typename T, int N>
__global__ void __launch_bounds__(32 * StepCfg<T, N>::WPB, GMJ_BLOCKS_PER_SM)
k_step(EnvState<T>* __restrict__ envs, const DevModel<T>* __restrict__ Mp, const DevSettings* __restrict__ Sp, int n_envs,
       int flags, int nsub_override, const int32_t* __restrict__ actions, const float* __restrict__ frac,
       float* __restrict__ obs, float* __restrict__ reward, uint8_t* __restrict__ done, float* __restrict__ events,
       int* __restrict__ err_flag, int* __restrict__ env_list, int* __restrict__ env_count, int list_blocks,
       const int* __restrict__ perm, int* __restrict__ hist, int bal_sm, int lift_wpb, T* __restrict__ tap_buf) {
  constexpr int WPB = StepCfg<T, N>::WPB;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int s_nsub_max;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int bid = blockIdx.x;
  bool list_block = (flags & STEP_LIST) != 0;
  if (flags & STEP_COMBINED) { list_block = bid < list_blocks; if (!list_block) bid -= list_blocks; }
  int env = bid * WPB + warp;
  bool valid = env < n_envs;
  if (list_block) {                        // the collected envs, lift_wpb per block; surplus blocks leave at once
    // a lifting env runs 3S substeps back to back: few warps per block keep that critical path short
    // (a 4-warp block steps ~2.5x faster than a full one), the idle warps only pass the barriers
    const int LW = lift_wpb;
    const int cnt = *env_count;
    if (bid * LW >= cnt) return;
    env = bid * LW + warp;
    valid = warp < LW && env < cnt;
    env = valid ? env_list[env] : 0;
  } else if (flags & STEP_COMBINED) {      // everybody else, in the load-sorted order built by k_sort_scatter
    // envs per block: the fewest full rounds over the SMs, then an equal share per block, so the last round is
    // not a handful of full blocks on a few SMs (4096 envs / 12 = 2.3 rounds of 148 blocks). Lifting envs weigh 3.
    const int cnt = *env_count, n_rest = n_envs - cnt;
    int W = WPB;
    if (bal_sm > 0) {
      const int units = n_rest + 3 * cnt, rounds = (units + bal_sm * WPB - 1) / (bal_sm * WPB);
      W = (units + bal_sm * rounds - 1) / (bal_sm * rounds);
      W = W < 1 ? 1 : (W > WPB ? WPB : W);
    }
    if (bid * W >= n_rest) return;
    env = bid * W + warp;
    valid = warp < W && env < n_rest;
    env = valid ? perm[env] : 0;
  }
  WarpShared<T, N>& ws = *reinterpret_cast<WarpShared<T, N>*>(smem_raw + (size_t)warp * sizeof(WarpShared<T, N>));
  // N = 8 only: the model constants (73 per-lane fields + scalars, 9.9 KB in FP32), read by every lane in every substep,
  // get one copy per block in shared memory behind the per-warp records -- beside 203 KB of shared memory the L1 is
  // small (ncu r2: long_sb on the M.f[...] loads; 27.15 -> 26.79 ms at 16384 envs). At N = 5 the L1 holds them and the
  // copy was measured 4 % SLOWER, so it stays in global memory there.
  const DevModel<T>* Mq = Mp;
  if constexpr (StepCfg<T, N>::MODEL_SMEM) {
    DevModel<T>* const sM = reinterpret_cast<DevModel<T>*>(smem_raw + (size_t)WPB * sizeof(WarpShared<T, N>));
    for (int i = threadIdx.x; i < (int)(sizeof(DevModel<T>) / 4); i += blockDim.x)
      reinterpret_cast<unsigned*>(sM)[i] = reinterpret_cast<const unsigned*>(Mp)[i];
    Mq = sM;
  }
  const DevModel<T>& M = *Mq; const DevSettings& S = *Sp;
  if (threadIdx.x == 0) s_nsub_max = 0;
  if (valid) {
    copy_words<T>(&ws.e, &envs[env], sizeof(EnvState<T>), lane);
    if (lane == 0) { ws.ax.ncon = 0; ws.ax.nrow = 0; ws.ax.niter = 0; ws.ax.wrist_z = ws.e.wrist_z; }
    ws.qfk[lane] = ws.e.qpos[lane];
  }
  __syncthreads();
  EnvState<T>& e = ws.e;
  int extra = 0;
  if ((flags & STEP_SET_ACTION) && valid) {
    if (lane == 0) {
      extra = set_action<T>(e, S, M, actions[env], frac ? frac[env] : 0.0f);
      if (extra < 0) { extra = 0; if (err_flag) atomicExch(err_flag, 1); }
    }
    if ((flags & STEP_COLLECT) && lane == 0) {
      e.pending_extra = extra;
      const int key = sort_key_of(e);      // load key of the lockstep sort (k_sort_scatter fills perm / env_list)
      e.sort_key = key;
      if (extra > 0) { atomicAdd(env_count, 1); atomicAdd(&hist[2 * GMJ_SORT_BUCKETS + key], 1); }
      else atomicAdd(&hist[key], 1);
    }
    extra = __shfl_sync(GMJ_FULL, extra, 0);
    __syncwarp();
  }
  if (list_block && valid) {               // owed lift substeps run first, exactly as inside MjClass::set_action
    extra = (int)e.pending_extra;
    __syncwarp();
    if (lane == 0) e.pending_extra = 0;
  }
  if (flags & STEP_SUBSTEPS) {
    const int nsub = valid ? extra + (nsub_override >= 0 ? nsub_override : (int)S.s.general[GS_sim_steps_per_action]) : 0;
    if (lane == 0) atomicMax(&s_nsub_max, nsub);
    __syncthreads();
    const int nsub_max = s_nsub_max;
    const double dtd = S.s.general[GS_mujoco_timestep];
    const bool curve = S.s.general[GS_curve_validation] != 0;
    const T tipf = curve ? (T)S.s.general[GS_tip_force_applied] : T(0);
    T* const taps = tap_buf + (size_t)env * (GMJ_MAXTAP * (GMJ_TAP_HDR + 3 * N));
    int ntap = 0;
#ifdef GMJ_PROF
    const long long prof_loop0 = clock64();
#endif
    for (int i = 0; i < nsub_max; i++) {
      __syncthreads();                       // lockstep: shared instruction stream
      if (i < nsub) {
        if (curve && lane == 0) for (int id = 0; id < 6; id++) set_weld<T, N>(e, id, 1, ws.qfk);   // apply_segment_force re-locks every step
        __syncwarp();
        physics_substep<T, N, 32 * WPB>(ws, M, S, lane, dtd, tipf);
        phase_barrier<32 * WPB>(GMJ_NPHASE);
#ifdef GMJ_PROF
        long long prof_t0 = clock64();
#endif
        if (lane == 0) update_all<T, N>(e, ws.qfk, S.asin_tab);
        __syncwarp();
        if (flags & STEP_SENSORS) sensor_taps<T, N>(e, S, M, ws.ax, taps, ntap, lane);
        __syncwarp();
#ifdef GMJ_PROF
        if (lane == 0) e.prof[9] += (unsigned)(clock64() - prof_t0);
#endif
      } else { phase_idle<32 * WPB>(); phase_barrier<32 * WPB>(GMJ_NPHASE); }
    }
    if (valid && lane == 0) e.wrist_z = (float)ws.ax.wrist_z;
#ifdef GMJ_PROF
    if (valid && lane == 0) e.prof[11] += (unsigned)(clock64() - prof_loop0);
#endif
    // the taps of this launch, in time order, outside the lockstep loop (no other warp waits for the FP64 fits)
    if (valid && ntap > 0) { __syncwarp(); process_taps<T, N>(e, S, M, taps, ntap, lane); }
  }
  if (!valid) return;
  if ((flags & STEP_TAIL) && lane == 0) {
    sense_gripper_state<T>(e, S, M);
    update_env<T>(e, S, ws.ax);
    e.num_action_steps += 1;
  }
  if ((flags & STEP_OUTPUT) && lane == 0) {
    if (reward) { float r = calc_reward<T>(e, S); e.last_reward = r; reward[env] = r; }
    int d = calc_done<T>(e, S); e.last_done = d;
    if (done) done[env] = (uint8_t)d;
    if (obs) get_observation<T>(e, S, obs + (size_t)env * S.n_obs);
    if (events) event_state<T>(e, events + (size_t)env * GMJ_EVENT_STATE_LEN);
  }
  __syncwarp();
  copy_words<T>(&envs[env], &ws.e, sizeof(EnvState<T>), lane);
}

template <typename T, int N>
int launch_step_TN(GmjHandle* h, int flags, int nsub, const int32_t* actions, const float* frac, float* obs,
                          float* reward, uint8_t* done, float* events, cudaStream_t st) {
  const bool combined = (flags & STEP_COMBINED) != 0;
  constexpr int WPB = StepCfg<T, N>::WPB;
  size_t smem = sizeof(WarpShared<T, N>) * WPB + (StepCfg<T, N>::MODEL_SMEM ? sizeof(DevModel<T>) : 0);   // + the block's copy of the model
  // the opt-in to > 48 KB of dynamic shared memory is PER DEVICE: one flag per device (a handle on a second GPU of
  // the same process would otherwise launch without it and fail)
  static std::atomic<bool> attr_set[GMJ_MAX_DEVICES];
  if (h->device < 0 || h->device >= GMJ_MAX_DEVICES) FAIL("device index out of range");
  if (!attr_set[h->device].load(std::memory_order_acquire)) {
    CK(cudaFuncSetAttribute(k_step<T, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set[h->device].store(true, std::memory_order_release);
  }
  int blocks = (h->n_envs + WPB - 1) / WPB;
  // Few envs (<= 3 rounds of full blocks over the SMs): the step is as long as the chain of 3S substeps of the lifting
  // envs, and a block of 8 warps steps faster than a block of 12 (A/B on B200: +3 % at 4096 envs, -6 % at 16384).
  int LW = GMJ_LIFT_WPB < WPB ? GMJ_LIFT_WPB : WPB;
  if (h->n_sm > 0 && h->n_envs <= 3 * h->n_sm * WPB && LW > GMJ_LIFT_SMALL) LW = GMJ_LIFT_SMALL;
  if (flags & STEP_LIST) blocks = (h->n_envs + LW - 1) / LW;
  const int list_blocks = combined ? (h->n_envs + LW - 1) / LW : 0;   // worst case every env owes a lift; empty list blocks exit at once
  const int bal_sm = combined ? h->balance_sm : 0;
  if (bal_sm > 0) blocks = bal_sm * ((3 * h->n_envs + bal_sm * WPB - 1) / (bal_sm * WPB));   // bound (every env lifting); surplus blocks exit at once
  k_step<T, N><<<blocks + list_blocks, 32 * WPB, smem, st>>>((EnvState<T>*)h->envs, (const DevModel<T>*)h->model, h->settings,
      h->n_envs, flags, nsub, actions, frac, obs, reward, done, events, h->err_flag, h->env_list, h->env_count, list_blocks, h->perm, h->hist, bal_sm, LW, (T*)h->taps);
  CK(cudaGetLastError());
  h->smem_bytes = smem;
  return 0;
}

// ------------------------------------------------------------------------------------------ wide kernel
// 96 threads per environment (gmj_wide.cuh), WideCfg::EPB environments per block in lockstep (FP32: 4 at N = 24, one
// block per SM; FP64: 1). Same two-pass scheme as the lane-per-dof kernel: pass 1 = set_action + collect (envs that
// owe the termination action's 2S lift substeps -- quirk Q8, they run INSIDE MjClass::set_action -- go to env_list,
// the others are counting-sorted by load), pass 2 = combined launch, the lifting envs densely packed in the first
// blocks (a single lifter would hold its lockstep block for 3S substeps).
template <typename T, int N>
__global__ void __launch_bounds__(GMJ_WNT * WideCfg<T, N>::EPB, 1)
k_step_wide(EnvState<T, GMJ_WQC>* __restrict__ envs, const DevModel<T, GMJ_WNT>* __restrict__ Mp, const DevSettings* __restrict__ Sp,
            int n_envs, int flags, int nsub_override, const int32_t* __restrict__ actions, const float* __restrict__ frac,
            float* __restrict__ obs, float* __restrict__ reward, uint8_t* __restrict__ done, float* __restrict__ events,
            int* __restrict__ err_flag, int* __restrict__ env_list, int* __restrict__ env_count, int list_blocks,
            const int* __restrict__ perm, int* __restrict__ hist, T* __restrict__ tap_buf) {
  constexpr int EPB = WideCfg<T, N>::EPB;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int s_nsub_max;
  const int grp = threadIdx.x / GMJ_WNT, tid = threadIdx.x - grp * GMJ_WNT, lane = tid & 31, warp = tid >> 5;
  WideShared<T, N>& ws = *reinterpret_cast<WideShared<T, N>*>(smem_raw + (size_t)grp * sizeof(WideShared<T, N>));
  int bid = blockIdx.x;
  bool list_block = (flags & STEP_LIST) != 0;
  if (flags & STEP_COMBINED) { list_block = bid < list_blocks; if (!list_block) bid -= list_blocks; }
  int env = bid * EPB + grp;
  bool valid = env < n_envs;
  if (list_block) {
    const int cnt = *env_count;
    if (bid * EPB >= cnt) return;
    valid = env < cnt;
    env = valid ? env_list[env] : 0;
  } else if (flags & STEP_COMBINED) {
    const int n_rest = n_envs - *env_count;
    if (bid * EPB >= n_rest) return;
    valid = env < n_rest;
    env = valid ? perm[env] : 0;
  } else if (bid * EPB >= n_envs) return;
  const DevModel<T, GMJ_WNT>& M = *Mp; const DevSettings& S = *Sp;
  if (threadIdx.x == 0) s_nsub_max = 0;
  if (tid == 0) ws.grp = grp;
  if (valid) {
    const uint32_t* s = (const uint32_t*)&envs[env]; uint32_t* d = (uint32_t*)&ws.e;
    for (int i = tid; i < (int)(sizeof(EnvState<T, GMJ_WQC>) / 4); i += GMJ_WNT) d[i] = s[i];
  }
  __syncthreads();
  EnvState<T, GMJ_WQC>& e = ws.e;
  if (valid) {
    if (tid == 0) { ws.ax.ncon = 0; ws.ax.nrow = 0; ws.ax.niter = 0; ws.ax.wrist_z = e.wrist_z; ws.extra = 0; }
    for (int i = tid; i < GMJ_WQC; i += GMJ_WNT) ws.qfk[i] = e.qpos[i];
  }
  __syncthreads();
  if ((flags & STEP_SET_ACTION) && valid && tid == 0) {
    int extra = set_action<T>(e, S, M, actions[env], frac ? frac[env] : 0.0f);
    if (extra < 0) { extra = 0; if (err_flag) atomicExch(err_flag, 1); }
    ws.extra = extra;
    if (flags & STEP_COLLECT) {
      e.pending_extra = extra;
      const int key = sort_key_of(e);
      e.sort_key = key;
      if (extra > 0) { atomicAdd(env_count, 1); atomicAdd(&hist[2 * GMJ_SORT_BUCKETS + key], 1); }
      else atomicAdd(&hist[key], 1);
    }
  }
  if (list_block && valid && tid == 0) { ws.extra = (int)e.pending_extra; e.pending_extra = 0; }
  __syncthreads();
  if (flags & STEP_SUBSTEPS) {
    const int nsub = valid ? ws.extra + (nsub_override >= 0 ? nsub_override : (int)S.s.general[GS_sim_steps_per_action]) : 0;
    if (tid == 0) atomicMax(&s_nsub_max, nsub);
    __syncthreads();
    const int nsub_max = s_nsub_max;
    const double dtd = S.s.general[GS_mujoco_timestep];
    const bool curve = S.s.general[GS_curve_validation] != 0;
    const T tipf = curve ? (T)S.s.general[GS_tip_force_applied] : T(0);
    T* const taps = tap_buf + (size_t)env * (GMJ_MAXTAP * (GMJ_TAP_HDR + 3 * N));
    int ntap = 0;                            // (kept by warp 0 of the env)
    for (int i = 0; i < nsub_max; i++) {
      wide_phase<T, N>();                    // lockstep: one instruction stream for the EPB envs of the block
      if (i < nsub) {
        if (curve && tid == 0) for (int id = 0; id < 6; id++) set_weld<T, N>(e, id, 1, ws.qfk);   // apply_segment_force re-locks every step
        env_sync<T, N>(ws);
        physics_substep_wide<T, N>(ws, M, S, tid, dtd, tipf);
        env_sync<T, N>(ws);
        if (tid == 0) update_all<T, N>(e, ws.qfk, S.asin_tab);
        env_sync<T, N>(ws);
        if ((flags & STEP_SENSORS) && warp == 0) sensor_taps<T, N>(e, S, M, ws.ax, taps, ntap, lane);
        env_sync<T, N>(ws);
      } else wide_phase_idle<T, N>();
    }
    if (valid && tid == 0) e.wrist_z = (float)ws.ax.wrist_z;
    if (valid && warp == 0 && ntap > 0) { __syncwarp(); process_taps<T, N>(e, S, M, taps, ntap, lane); }
  }
  if (!valid) return;
  env_sync<T, N>(ws);
  if ((flags & STEP_TAIL) && tid == 0) {
    sense_gripper_state<T>(e, S, M);
    update_env<T>(e, S, ws.ax);
    e.num_action_steps += 1;
  }
  if ((flags & STEP_OUTPUT) && tid == 0) {
    if (reward) { float r = calc_reward<T>(e, S); e.last_reward = r; reward[env] = r; }
    int d = calc_done<T>(e, S); e.last_done = d;
    if (done) done[env] = (uint8_t)d;
    if (obs) get_observation<T>(e, S, obs + (size_t)env * S.n_obs);
    if (events) event_state<T>(e, events + (size_t)env * GMJ_EVENT_STATE_LEN);
  }
  env_sync<T, N>(ws);
  {
    uint32_t* d = (uint32_t*)&envs[env]; const uint32_t* s = (const uint32_t*)&ws.e;
    for (int i = tid; i < (int)(sizeof(EnvState<T, GMJ_WQC>) / 4); i += GMJ_WNT) d[i] = s[i];
  }
}

template <typename T, int N>
int launch_step_wide_TN(GmjHandle* h, int flags, int nsub, const int32_t* actions, const float* frac, float* obs,
                        float* reward, uint8_t* done, float* events, cudaStream_t st) {
  constexpr int EPB = WideCfg<T, N>::EPB;
  const bool combined = (flags & STEP_COMBINED) != 0;
  size_t smem = sizeof(WideShared<T, N>) * EPB;
  static std::atomic<bool> attr_set[GMJ_MAX_DEVICES];       // per device, see launch_step_TN
  if (h->device < 0 || h->device >= GMJ_MAX_DEVICES) FAIL("device index out of range");
  if (!attr_set[h->device].load(std::memory_order_acquire)) {
    CK(cudaFuncSetAttribute(k_step_wide<T, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set[h->device].store(true, std::memory_order_release);
  }
  const int blocks = (h->n_envs + EPB - 1) / EPB;
  const int list_blocks = combined ? blocks : 0;        // worst case every env owes a lift; empty list blocks exit at once
  k_step_wide<T, N><<<blocks + list_blocks, GMJ_WNT * EPB, smem, st>>>((EnvState<T, GMJ_WQC>*)h->envs, (const DevModel<T, GMJ_WNT>*)h->model, h->settings,
      h->n_envs, flags, nsub, actions, frac, obs, reward, done, events, h->err_flag, h->env_list, h->env_count, list_blocks, h->perm, h->hist, (T*)h->taps);
  CK(cudaGetLastError());
  h->smem_bytes = smem;
  return 0;
}
