// gmj_device.cuh -- device-side data layout and small math helpers for the batched gripper env.
//
// One environment per warp. Lane i owns gripper dof/body i in the reference order
// (include/gmj_types.h): lane 0 = base slide, finger f occupies lanes fb..fb+N+1 with
// fb = 1 + f*(N+2) (prismatic, revolute, N segment hinges), lane 3N+7 = palm. The free object
// (6 dofs) is replicated in every lane. Per-env state lives in one contiguous block of HBM
// (env-major), so a warp loads/stores its environment with coalesced accesses.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/gmj_types.h"

#define GMJ_FULL 0xffffffffu
#define GMJ_WIN 8            // ring depth kept per sensor stream (reference keeps 1000, reads <= 7)
#define GMJ_RMAX 48          // constraint rows per env per substep (== GMJ_MAXROW)
#define GMJ_NSTREAM_F 8      // normalised force streams: gauge1-3, axial1-3, palm, wristZ
#define GMJ_NSTREAM_S 7      // state streams: motor x,y,z, base x,y, base z, yaw
#define GMJ_NSTREAM_C 12     // cartesian contact streams (MAT)
#define GMJ_NSTREAM (GMJ_NSTREAM_F + GMJ_NSTREAM_S + GMJ_NSTREAM_C)

// luke::Gripper motor-space state (src/gripper.h)
struct DevGripper { double x, y, z, th; int sx, sy, sz, pad; };

// per-env persistent state (one block per env in HBM)
// QC = capacity of qpos/qvel: 40 for the lane-per-dof kernels (B <= 32), 96 for the wide kernels (B <= 88)
template <typename T, int QC = 40>
struct EnvState {
  static constexpr int OQ = QC - 8;   // first object entry
  T qpos[QC];                 // [0,B) gripper joints in reference dof order; [OQ,OQ+7) object pos(3)+quat(4)
  T qvel[QC];                 // [0,B) gripper; [OQ,OQ+6) object lin(3 world)+ang(3 body)
  T fprev[GMJ_RMAX];          // PGS warm start
  T weld_anchor[8];           // joint value at weld activation (7 used)
  T obj[20];                  // type,size3,round,quat4,mass,inertia3,friction,z_rest,bbox3
  double time, last_step_time;
  double last_read_time[4];   // bending, axial, palm, wristZ
  DevGripper tend, tnext;     // target_.end / target_.next
  double base_x, base_y, base_z, base_yaw;
  double start_z;
  double obj_xy[2];           // object x, y integrated in double beside the FP32 qpos (events compare it with thresholds the
                              // spawn grid puts objects exactly ON: within_XY_distance, src/mjclass.cpp:1237)
  float win[GMJ_NSTREAM][GMJ_WIN];
  int win_i[GMJ_NSTREAM];
  float si_latest[5];         // latest SI readings: gauge1-3, palm, wristZ
  float rand_mu[GSEN_COUNT][3];
  float ev_value[GBE_COUNT + GLE_COUNT];
  float ev_last[GBE_COUNT + GLE_COUNT];
  int ev_row[GBE_COUNT + GLE_COUNT];
  int ev_abs[GBE_COUNT + GLE_COUNT];
  float cumulative_reward, peak_lateral, last_reward, wrist_z;
  int num_action_steps, termination_sent, unstable, has_object;
  int weld_on, old_moving, nrow_prev, last_done;
  uint32_t rng_ctr, env_id, episode, obj_index;
  unsigned long long st_contacts, st_rows, st_iters, st_substeps;
  unsigned int st_unstable, st_overflow, pending_extra, sort_key;   // pending_extra: lift substeps owed by set_action
  unsigned int st_relaxed, st_capped;   // constraint solves accepted under a widened feasibility tolerance (3+ rounds) / at the round cap while infeasible
  int mpr_prev;                     // MPR passes (0-2: finger plates, hooks) the last substep ran: load key of the lockstep sort
#ifdef GMJ_PROF
  unsigned int prof[12];            // development build: cycles per phase (gmj_physics.cuh GMJ_PROF_*)
#endif
};

// model constants, lane-major [field][32] so a warp reads each field conflict-free
enum {
  MF_parent = 0, MF_jtype, MF_depth, MF_fb, MF_local, MF_group,
  MF_pos0, MF_pos1, MF_pos2,
  MF_R0, MF_R1, MF_R2, MF_R3, MF_R4, MF_R5, MF_R6, MF_R7, MF_R8,
  MF_ax0, MF_ax1, MF_ax2,
  MF_mass, MF_ip0, MF_ip1, MF_ip2,
  MF_ixx, MF_iyy, MF_izz, MF_ixy, MF_ixz, MF_iyz,
  MF_stiff, MF_damp, MF_lo, MF_hi, MF_qpos0, MF_kp, MF_kd, MF_tsel,
  // primary collision geom of the body (segment box / palm cylinder) and the hook plate
  MF_gtype, MF_gp0, MF_gp1, MF_gp2, MF_gs0, MF_gs1, MF_gs2, MF_gfric, MF_gcls,
  MF_gR0, MF_gR1, MF_gR2, MF_gR3, MF_gR4, MF_gR5, MF_gR6, MF_gR7, MF_gR8,
  MF_htype, MF_hp0, MF_hp1, MF_hp2, MF_hs0, MF_hs1, MF_hs2,
  MF_hR0, MF_hR1, MF_hR2, MF_hR3, MF_hR4, MF_hR5, MF_hR6, MF_hR7, MF_hR8,
  MF_COUNT
};

template <typename T>
struct DevModelS {               // scalars (all the tail logic needs)
  int N, B, NL, palm_lane;
  T gravity, solref_tc, solref_dr, solimp[5];
  T ground_friction, finger_length, segment_length, EI, clearance, zref;
  T hook_length, hook_angle_deg, finger_width;
  int tip_lane[3], finger_lane[3];
  T tip_mat[3][9];
  T qpos0[GMJ_MAXB];           // keyframe, reference dof order
};
// LN = 32: lane-per-dof kernels (lane == reference dof index); LN = 96: wide kernels, slot = 32 * finger + lane
template <typename T, int LN = 32>
struct DevModel : DevModelS<T> {
  T f[MF_COUNT][LN];
};

struct DevSettings {
  GmjSettings s;
  int n_actions, action_options[32];
  int prev_steps[GSEN_COUNT], readings_per_step[GSEN_COUNT];
  int n_obs;
  int obs_ok;   // 0: the sensor history asked for is deeper than the device ring (observation unavailable)
  double read_between[GSEN_COUNT];   // (double)(1.0f / (float)read_rate) per sensor: MjType::Sensor::ready_to_read's interval
  const double* asin_tab;   // asin(G_XY_STEP d / G_LEADSCREW), d in [-G_TH_TAB, G_TH_TAB] (stepper ticks, gmj_tail.cuh)
};

// ------------------------------------------------------------------ vec helpers
template <typename T> struct V3 { T x, y, z; };
template <typename T> __device__ __forceinline__ V3<T> mk(T x, T y, T z) { V3<T> r; r.x = x; r.y = y; r.z = z; return r; }
template <typename T> __device__ __forceinline__ V3<T> operator+(V3<T> a, V3<T> b) { return mk<T>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename T> __device__ __forceinline__ V3<T> operator-(V3<T> a, V3<T> b) { return mk<T>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename T> __device__ __forceinline__ V3<T> operator*(V3<T> a, T s) { return mk<T>(a.x * s, a.y * s, a.z * s); }
template <typename T> __device__ __forceinline__ T dot(V3<T> a, V3<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> __device__ __forceinline__ V3<T> cross(V3<T> a, V3<T> b) {
  return mk<T>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
template <typename T> __device__ __forceinline__ V3<T> shfl3(V3<T> v, int src) {
  return mk<T>(__shfl_sync(GMJ_FULL, v.x, src), __shfl_sync(GMJ_FULL, v.y, src), __shfl_sync(GMJ_FULL, v.z, src));
}
template <typename T> __device__ __forceinline__ V3<T> shflup3(V3<T> v, int d) {
  return mk<T>(__shfl_up_sync(GMJ_FULL, v.x, d), __shfl_up_sync(GMJ_FULL, v.y, d), __shfl_up_sync(GMJ_FULL, v.z, d));
}
template <typename T> __device__ __forceinline__ V3<T> shfldn3(V3<T> v, int d) {
  return mk<T>(__shfl_down_sync(GMJ_FULL, v.x, d), __shfl_down_sync(GMJ_FULL, v.y, d), __shfl_down_sync(GMJ_FULL, v.z, d));
}
// row-major 3x3
template <typename T> struct M3 { T m[9]; };
template <typename T> __device__ __forceinline__ V3<T> mul(const M3<T>& R, V3<T> v) {
  return mk<T>(R.m[0] * v.x + R.m[1] * v.y + R.m[2] * v.z, R.m[3] * v.x + R.m[4] * v.y + R.m[5] * v.z,
               R.m[6] * v.x + R.m[7] * v.y + R.m[8] * v.z);
}
template <typename T> __device__ __forceinline__ V3<T> mulT(const M3<T>& R, V3<T> v) {
  return mk<T>(R.m[0] * v.x + R.m[3] * v.y + R.m[6] * v.z, R.m[1] * v.x + R.m[4] * v.y + R.m[7] * v.z,
               R.m[2] * v.x + R.m[5] * v.y + R.m[8] * v.z);
}
template <typename T> __device__ __forceinline__ M3<T> mul(const M3<T>& A, const M3<T>& B) {
  M3<T> o;
#pragma unroll
  for (int r = 0; r < 3; r++)
#pragma unroll
    for (int c = 0; c < 3; c++) o.m[3 * r + c] = A.m[3 * r] * B.m[c] + A.m[3 * r + 1] * B.m[3 + c] + A.m[3 * r + 2] * B.m[6 + c];
  return o;
}
template <typename T> __device__ __forceinline__ M3<T> shflM(const M3<T>& A, int src) {
  M3<T> o;
#pragma unroll
  for (int k = 0; k < 9; k++) o.m[k] = __shfl_sync(GMJ_FULL, A.m[k], src);
  return o;
}
template <typename T> __device__ __forceinline__ M3<T> quat2mat(T w, T x, T y, T z) {
  M3<T> R;
  R.m[0] = 1 - 2 * (y * y + z * z); R.m[1] = 2 * (x * y - w * z); R.m[2] = 2 * (x * z + w * y);
  R.m[3] = 2 * (x * y + w * z); R.m[4] = 1 - 2 * (x * x + z * z); R.m[5] = 2 * (y * z - w * x);
  R.m[6] = 2 * (x * z - w * y); R.m[7] = 2 * (y * z + w * x); R.m[8] = 1 - 2 * (x * x + y * y);
  return R;
}
template <typename T> __device__ __forceinline__ void sincosT(T t, T* s, T* c);
// joint angles and per-substep rotation increments are well inside [-pi, pi]: MUFU sin/cos (abs error
// ~4e-7, the size of an FP32 ulp of the rotation-matrix entries) instead of the ~40-instruction libm path
template <> __device__ __forceinline__ void sincosT<float>(float t, float* s, float* c) { __sincosf(t, s, c); }
template <> __device__ __forceinline__ void sincosT<double>(double t, double* s, double* c) { sincos(t, s, c); }

// Double-precision libm calls expand to hundreds of SASS instructions each; the step kernel runs at
// the speed of instruction fetch, so every call site shares ONE out-of-line copy.
static __device__ __noinline__ double nd_asin(double x) { return asin(x); }
static __device__ __noinline__ double nd_sin(double x) { return sin(x); }
static __device__ __noinline__ double nd_cos(double x) { return cos(x); }
// round() (half away from zero), exact for |x| < 2^51: truncate, then step on the remainder. The libm
// expansion is ~70 instructions; this is called three times per stepper tick.
__device__ __forceinline__ double nd_round(double x) {
  double r = (double)__double2ll_rz(x), d = x - r;
  if (d >= 0.5) r += 1.0; else if (d <= -0.5) r -= 1.0;
  return r;
}
static __device__ __noinline__ double nd_log(double x) { return log(x); }
template <typename T> __device__ __forceinline__ T sqrtT(T x);
template <> __device__ __forceinline__ float sqrtT<float>(float x) {
  // MUFU.RSQ + one Newton step (<= 1 ulp), no out-of-line slow path: sqrtf() is ~12 instructions per site
  float r = rsqrtf(x), s0 = x * r;
  s0 = fmaf(fmaf(-s0, s0, x), 0.5f * r, s0);
  return x > 0.0f ? s0 : 0.0f;
}
template <> __device__ __forceinline__ double sqrtT<double>(double x) { return sqrt(x); }
template <typename T> __device__ __forceinline__ T rsqrtT(T x);
template <> __device__ __forceinline__ float rsqrtT<float>(float x) { float r = rsqrtf(x); return r * fmaf(fmaf(-x * r, r, 1.0f), 0.5f, 1.0f); }
template <> __device__ __forceinline__ double rsqrtT<double>(double x) { return 1.0 / sqrt(x); }
// 1 / x without the IEEE division's special-case subroutine: ncu r1 showed ptxas' slow path entered ~17 times per substep
// by a third of the lanes (idle lanes, tiny pivots) and the whole warp waits for it. MUFU.RCP + one Newton step is
// within an ulp for normal operands; FP64 keeps the exact division (the FP64 build is the parity check of the algorithm).
template <typename T> __device__ __forceinline__ T rcpT(T x);
template <> __device__ __forceinline__ float rcpT<float>(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return fmaf(r, fmaf(-x, r, 1.0f), r);
}
template <> __device__ __forceinline__ double rcpT<double>(double x) { return 1.0 / x; }
template <typename T> __device__ __forceinline__ T absT(T x) { return x < 0 ? -x : x; }
template <typename T> __device__ __forceinline__ T maxT(T a, T b) { return a > b ? a : b; }
template <typename T> __device__ __forceinline__ T minT(T a, T b) { return a < b ? a : b; }
template <typename T> __device__ __forceinline__ T powT(T a, T b);
template <> __device__ __forceinline__ float powT<float>(float a, float b) { return powf(a, b); }
template <> __device__ __forceinline__ double powT<double>(double a, double b) { return pow(a, b); }

template <typename T> __device__ __forceinline__ M3<T> rot_axis(V3<T> a, T t) {
  T s, c; sincosT<T>(t, &s, &c);
  T v = 1 - c;
  M3<T> R;
  R.m[0] = c + a.x * a.x * v; R.m[1] = a.x * a.y * v - a.z * s; R.m[2] = a.x * a.z * v + a.y * s;
  R.m[3] = a.y * a.x * v + a.z * s; R.m[4] = c + a.y * a.y * v; R.m[5] = a.y * a.z * v - a.x * s;
  R.m[6] = a.z * a.x * v - a.y * s; R.m[7] = a.z * a.y * v + a.x * s; R.m[8] = c + a.z * a.z * v;
  return R;
}
template <typename T> __device__ __forceinline__ V3<T> normalized(V3<T> a) {
  T l = sqrtT<T>(dot(a, a));
  return l > T(1e-30) ? a * rcpT<T>(l) : a;      // (a denormal length flushes to zero in rcp.approx.ftz)
}
template <typename T> __device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(GMJ_FULL, v, o);
  return v;
}
template <typename T> __device__ __forceinline__ T warp_max(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { T u = __shfl_xor_sync(GMJ_FULL, v, o); v = v > u ? v : u; }
  return v;
}

// Philox4x32-10, first output word; identical to the oracle's generator (oracle/env.c)
__device__ __forceinline__ uint32_t philox_first(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; r++) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return c0;
}
