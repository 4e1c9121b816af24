// gmj_tail.cuh -- per-env scalar logic around the physics substeps (executed by lane 0 on the
// env state staged in shared memory): motor-space kinematics, stepper + weld toggling, 10 Hz sensor
// taps, sliding windows, events, reward, done and observation assembly.
// Reference lines each block replaces are cited inline (paths relative to the reference tree).
#pragma once
#include "gmj_device.cuh"

// contact record kept in shared memory for the force tally
template <typename T>
struct DevContact {
  T pos[3], frame[9], dist, mu, force[3];
  int body1, body2, cls1, cls2, row;
};

template <typename T>
struct WarpAux {                 // per-warp scratch written by the physics each substep
  T fingerR[3][9], palmR[9];     // xmat of finger_1..3 and palm (local force frames)
  T wrist_z;
  int ncon, nrow, niter;
  DevContact<T> con[GMJ_MAXCON];
};

// ------------------------------------------------------------------ luke::Gripper (src/gripper.h, src/gripper.cpp)
#define G_LEADSCREW 35e-3
#define G_XY_MIN 49e-3
#define G_XY_MAX 134e-3
#define G_Z_MAX 165e-3
#define G_LIMIT_TOL 1e-4
#define G_XY_STEP (4.0 / (1.0 * 400 * 1e3))
#define G_Z_STEP (4.8768 / (1 * 400 * 1e3))
#define G_TH_TAB 2400                    // |sx - sy| <= 2250 at the +-40 degree limit
#define G_XY_HOME (G_XY_MAX - 1.0 * (4 * 1e-3 / 1.0))
#define G_Z_HOME (0.0 + 1.0 * (4.8768 * 1e-3 / 1))

__device__ __forceinline__ double g_calc_th(double x, double y) { return nd_asin((y - x) / G_LEADSCREW); }
__device__ __forceinline__ double g_calc_y(double x, double th) { return x + G_LEADSCREW * nd_sin(th); }

// tab != nullptr: x and y were just set from the integer step counts (g_step_to) whose difference is d; when no
// limit moved them the angle comes from the table asin(G_XY_STEP d / G_LEADSCREW) -- the same value to the last
// bit or two (x and y are rounded images of the lattice) without ~100 FP64 instructions on one lane.
__device__ inline int g_update_xy(DevGripper& g, const double* tab = nullptr, int d = 0) {   // gripper.cpp:6-90
  int wl = 1;
  const double to_rad = 3.14159265358979323846 / 180.0;
  const double th_min = -40 * to_rad, th_max = 40 * to_rad;
  if (g.x > G_XY_MAX + G_LIMIT_TOL) { g.x = G_XY_MAX; wl = 0; }
  if (g.x < G_XY_MIN - G_LIMIT_TOL) { g.x = G_XY_MIN; wl = 0; }
  if (g.y > G_XY_MAX + G_LIMIT_TOL) { g.y = G_XY_MAX; wl = 0; }
  if (g.y < G_XY_MIN - G_LIMIT_TOL) { g.y = G_XY_MIN; wl = 0; }
  double nth = (tab && wl && d >= -G_TH_TAB && d <= G_TH_TAB) ? __ldg(tab + d + G_TH_TAB) : g_calc_th(g.x, g.y);
  if (nth > th_max) { nth = th_max; g.y = g_calc_y(g.x, th_max); wl = 0; }
  if (nth < th_min) { nth = th_min; g.y = g_calc_y(g.x, th_min); wl = 0; }
  g.th = nth;
  const double hyp = sqrt(235e-3 * 235e-3 + 35.0e-3 * 35.0e-3), rest = atan(35.0e-3 / 235e-3);
  // fingertip_radius_min = -1 (disabled): the asin argument is below -1 for every reachable x, the
  // reference gets NaN and the comparison is false; the call is only made when it can matter
  const double targ = (-1.0 - g.x) / hyp;
  if (targ >= -1.0 && targ <= 1.0) {
    double th_lim = nd_asin(targ) + rest;
    if (g.th < th_lim) { g.y = g_calc_y(g.x, th_lim); g.th = th_lim; wl = 0; }
  }
  g.sx = (int)nd_round((G_XY_MAX - g.x) / G_XY_STEP);
  g.sy = (int)nd_round((G_XY_MAX - g.y) / G_XY_STEP);
  return wl;
}
__device__ inline int g_update_z(DevGripper& g) {                     // gripper.cpp:92-118
  int wl = 1;
  if (g.z > G_Z_MAX + G_LIMIT_TOL) { g.z = G_Z_MAX; wl = 0; }
  if (g.z < 0.0 - G_LIMIT_TOL) { g.z = 0.0; wl = 0; }
  g.sz = (int)nd_round(g.z / G_Z_STEP);
  return wl;
}
__device__ inline int g_set_xyz_m(DevGripper& g, double x, double y, double z) {
  g.x = x; g.y = y; g.z = z;
  int a = g_update_xy(g), b = g_update_z(g);
  return a && b;
}
__device__ inline int g_set_xyz_m_rad(DevGripper& g, double x, double th, double z) {
  g.x = x; g.y = g_calc_y(g.x, th);
  int in_lim = g_update_xy(g);
  g.z = z;
  return g_update_z(g) ? in_lim : 0;
}
__device__ inline void g_reset(DevGripper& g) { g_set_xyz_m(g, G_XY_HOME, G_XY_HOME, G_Z_HOME); }
__device__ inline void g_step_to(DevGripper& g, const DevGripper& t, int num, const double* tab = nullptr) { // gripper.cpp:158-228
  int d[3] = { t.sx - g.sx, t.sy - g.sy, t.sz - g.sz };
#pragma unroll
  for (int k = 0; k < 3; k++) { if (d[k] < -num) d[k] = -num; if (d[k] > num) d[k] = num; }
  g.x = G_XY_MAX - G_XY_STEP * (g.sx + d[0]); g.y = G_XY_MAX - G_XY_STEP * (g.sy + d[1]);
  g_update_xy(g, tab, (g.sx + d[0]) - (g.sy + d[1]));
  g.z = G_Z_STEP * (g.sz + d[2]);
  g_update_z(g);
}

// ------------------------------------------------------------------ windows (src/slidingwindow.h:9-55)
enum { ST_gauge = 0, ST_axial = 3, ST_palm = 6, ST_wristz = 7, ST_motor = 8, ST_basexy = 11, ST_basez = 13,
       ST_yaw = 14, ST_cart = 15 };
template <typename T, int QC> __device__ __forceinline__ void win_add(EnvState<T, QC>& e, int s, float v) {
  int i = e.win_i[s] + 1;
  if (i > GMJ_WIN - 1) i = 0;
  e.win_i[s] = i; e.win[s][i] = v;
}
template <typename T, int QC> __device__ __forceinline__ float win_read(const EnvState<T, QC>& e, int s, int n) {
  int idx = e.win_i[s] - n;
  while (idx < 0) idx += GMJ_WIN;
  return e.win[s][idx];
}

// samplers MjType::Sensor::*_sample (src/mjclass.h:278-453)
template <typename T, int QC>
__device__ inline int sample_stream(const EnvState<T, QC>& e, int st, int mode, int p, int r, float* out) {
  int Tn = 1 + r * p;
  if (mode == GSM_raw) {
    int n = Tn - 1;
    for (int j = n - 1, k = 0; j >= 0; j--, k++) out[k] = win_read(e, st, j);
    return n;
  }
  out[0] = win_read(e, st, Tn - 1);
  for (int i = 0; i < p; i++) {
    int first = Tn - 1 - i * r;
    out[2 * i + 2] = win_read(e, st, first - r);
    float a = out[2 * i], b = out[2 * i + 2], v = 0.0f;
    if (mode == GSM_change) v = b - a;
    else if (mode == GSM_average) {
      v = 0;
      for (int j = 0; j < r + 1; j++) v += win_read(e, st, first - j);
      v /= r + 1;
    } else if (mode == GSM_median) {
      float vals[GMJ_WIN]; int n = r + 1;
      for (int j = 0; j < n; j++) vals[j] = win_read(e, st, first - j);
      for (int x = 1; x < n; x++) { float t = vals[x]; int y = x - 1; while (y >= 0 && vals[y] > t) { vals[y + 1] = vals[y]; y--; } vals[y + 1] = t; }
      int h = n / 2; v = vals[h];
      if (!(n & 1)) v = (float)((vals[h - 1] + v) / 2.0);
    } else if (mode == GSM_sign) {
      float c = b - a; v = (c > 1e-6f) ? 1.0f : ((c < -1e-6f) ? -1.0f : 0.0f);
    } else if (mode == GSM_scaled_change) {
      v = (b - a) / 0.07f; if (v > 1.0) v = 1.0f; else if (v < -1.0) v = -1.0f;
    } else if (mode == GSM_scaled_change_sq) {
      const float m2 = 0.10f; const float fac = (float)(1.0 / (m2 * m2));
      float c = fabsf(b - a); v = (b - a) * c * fac;
      if (v > 1.0) v = 1.0f; else if (v < -1.0) v = -1.0f;
    }
    out[2 * i + 1] = v;
  }
  return 2 * p + 1;
}

// ------------------------------------------------------------------ sensors
template <typename T, int QC> __device__ __noinline__ float dev_uniform(EnvState<T, QC>& e, const DevSettings& S) {
  uint64_t seed = (uint64_t)S.s.general[GS_random_seed];
  uint32_t x = philox_first(e.rng_ctr++, e.env_id, 0, 0, (uint32_t)seed, (uint32_t)(seed >> 32));
  return (float)(x >> 8) * (1.0f / 16777216.0f);
}
__device__ inline float apply_normalisation(const double* sn, float value) {       // src/mjclass.h:155-174
  if (!sn[GSF_use_normalisation]) return value;
  float normalise = (float)sn[GSF_normalise];
  if (normalise <= 0) return value < 0 ? -1.0f : 1.0f;
  else if (value > normalise) return 1.0f;
  if (value < -normalise) return -1.0f;
  return value / normalise;
}
template <typename T, int QC>
__device__ inline float apply_noise(EnvState<T, QC>& e, const DevSettings& S, int sensor, float value, int i) { // src/mjclass.h:176-223
  const double* sn = S.s.sensor[sensor];
  if (!sn[GSF_use_noise] || !S.s.general[GS_use_noise_seeded]) return value;
  const float eps = 1.1920929e-07f;
  float mu = e.rand_mu[sensor][i - 1];
  float noise_std = (float)sn[GSF_noise_std], noise_mag = (float)sn[GSF_noise_mag];
  if (noise_std < eps) {
    float noise = mu + noise_mag * (2 * dev_uniform(e, S) - 1);
    value += noise;
  } else {
    float u1, u2;
    do { u1 = dev_uniform(e, S); } while (u1 <= eps);
    u2 = dev_uniform(e, S);
    float mag = (float)(noise_std * sqrt(-2.0 * nd_log((double)u1)));
    float z0 = (float)(mag * nd_cos((double)(6.2831853071795864769f * u2))) + mu;
    value += z0;
  }
  if (value > 1) value = 1; else if (value < -1) value = -1;
  return value;
}
template <typename T, int QC> __device__ inline int ready_to_read(EnvState<T, QC>& e, const DevSettings& S, int sensor, int slot) { // src/mjclass.h:225-241
  const double between = S.read_between[sensor];       // = (double)(1 / (float)read_rate), formed once on the host
  if (e.time > e.last_read_time[slot] + between) { e.last_read_time[slot] = e.time; return 1; }
  return 0;
}

// luke::read_armadillo_gauge (src/myfunctions.cpp:2699-2795): cubic least squares through the
// N+1 joint positions of one finger evaluated at 50 mm, in mm. FP64 (arma::polyfit is double).
template <typename T, int N>
__device__ inline float read_gauge_q(const T* qf, double L) {          // qf: the N segment joint angles of one finger
  double X[N + 1][4], y[N + 1], fx[N + 1];
  double cum = 0; fx[0] = 0; y[0] = 0;
  for (int i = 0; i < N; i++) {
    cum += (double)qf[i];
    fx[i + 1] = fx[i] + L * nd_cos(cum);
    y[i + 1] = y[i] + L * nd_sin(cum);
  }
  const int n = N + 1;
  for (int i = 0; i < n; i++) { X[i][0] = fx[i] * fx[i] * fx[i]; X[i][1] = fx[i] * fx[i]; X[i][2] = fx[i]; X[i][3] = 1; }
  for (int k = 0; k < 4; k++) {           // Householder QR
    double nrm = 0;
    for (int i = k; i < n; i++) nrm += X[i][k] * X[i][k];
    nrm = sqrt(nrm);
    if (nrm == 0) continue;
    double alpha = X[k][k] > 0 ? -nrm : nrm;
    double v[N + 1], vn = 0;
    for (int i = k; i < n; i++) v[i] = X[i][k];
    v[k] -= alpha;
    for (int i = k; i < n; i++) vn += v[i] * v[i];
    if (vn == 0) continue;
    for (int c = k; c < 4; c++) {
      double d = 0;
      for (int i = k; i < n; i++) d += v[i] * X[i][c];
      d = 2 * d / vn;
      for (int i = k; i < n; i++) X[i][c] -= d * v[i];
    }
    double d = 0;
    for (int i = k; i < n; i++) d += v[i] * y[i];
    d = 2 * d / vn;
    for (int i = k; i < n; i++) y[i] -= d * v[i];
  }
  double coeff[4];
  for (int k = 3; k >= 0; k--) {
    double s = y[k];
    for (int c = k + 1; c < 4; c++) s -= X[k][c] * coeff[c];
    coeff[k] = (X[k][k] != 0) ? s / X[k][k] : 0;
  }
  float yv = 0.0f;
  const double xp = 50e-3;
  yv = (float)(yv + coeff[0] * (xp * xp * xp));
  yv = (float)(yv + coeff[1] * (xp * xp));
  yv = (float)(yv + coeff[2] * xp);
  yv = (float)(yv + coeff[3] * 1.0);
  return yv * 1000;
}
template <typename T, int N, int QC>
__device__ inline float read_gauge(const EnvState<T, QC>& e, const DevModelS<T>& M, int finger) {
  return read_gauge_q<T, N>(&e.qpos[3 + finger * (N + 2)], (double)M.segment_length);
}

// ObjectHandler::extract_forces_faster (src/objecthandler.cpp:737-992) with geom classes
struct DevForces {
  double obj_f[3][3], obj_palm[3], obj_gnd[3], all_f[3][3], all_palm[3], gnd_f[3][3];
  double obj_f_loc[3][3], obj_palm_loc[3], all_f_loc[3][3], all_palm_loc[3], gnd_f_loc[3][3];
};
template <typename T, int QC>
__device__ inline void extract_forces(const EnvState<T, QC>& e, const WarpAux<T>& ax, DevForces& F) {
  double* z = (double*)&F;
  for (int k = 0; k < (int)(sizeof(DevForces) / sizeof(double)); k++) z[k] = 0;
  if (!e.has_object) return;
  for (int c = 0; c < ax.ncon; c++) {
    const DevContact<T>& cn = ax.con[c];
    int a = cn.cls1, b = cn.cls2;
    double g[3] = { 0, 0, 0 };
    for (int k = 0; k < 3; k++) for (int d = 0; d < 3; d++) g[d] += (double)cn.force[k] * (double)cn.frame[3 * k + d];
    int with_obj = (a == GMJ_CLS_OBJECT || b == GMJ_CLS_OBJECT);
    int with_gnd = (a == GMJ_CLS_GROUND || b == GMJ_CLS_GROUND);
    int with_palm = (a == GMJ_CLS_PALM || b == GMJ_CLS_PALM);
    for (int f = 0; f < 3; f++) {
      int with_f = (a == GMJ_CLS_F1 + f || b == GMJ_CLS_F1 + f);
      for (int d = 0; d < 3; d++) {
        if (with_obj && with_f) F.obj_f[f][d] += g[d];
        if (with_f) { F.all_f[f][d] += g[d]; if (with_gnd) F.gnd_f[f][d] += g[d]; }
      }
    }
    for (int d = 0; d < 3; d++) {
      if (with_obj && with_palm) F.obj_palm[d] += g[d];
      if (with_obj && with_gnd) F.obj_gnd[d] += g[d];
      if (with_palm) F.all_palm[d] += g[d];
    }
  }
  for (int f = 0; f < 3; f++)
    for (int c = 0; c < 3; c++) {
      double so = 0, sa = 0, sg = 0;
      for (int r = 0; r < 3; r++) { double R = (double)ax.fingerR[f][3 * r + c]; so += R * F.obj_f[f][r]; sa += R * F.all_f[f][r]; sg += R * F.gnd_f[f][r]; }
      F.obj_f_loc[f][c] = so; F.all_f_loc[f][c] = sa; F.gnd_f_loc[f][c] = sg;
    }
  for (int c = 0; c < 3; c++) {
    double so = 0, sa = 0;
    for (int r = 0; r < 3; r++) { double R = (double)ax.palmR[3 * r + c]; so += R * F.obj_palm[r]; sa += R * F.all_palm[r]; }
    F.obj_palm_loc[c] = so; F.all_palm_loc[c] = sa;
  }
}

// MjClass::monitor_sensors (src/mjclass.cpp:741-898), split in two.
//  sensor_taps()   -- every substep, warp-parallel and cheap: the four 10 Hz timers (lanes 0-3, Sensor::ready_to_read)
//                     and, when one fires, a RECORD of the raw inputs of that tap (segment joint angles, finger-axial
//                     and palm force sums, wrist force) appended to the env's tap buffer in HBM.
//  process_taps()  -- after the substep loop: the records in time order through the reference's per-tap pipeline
//                     (cubic gauge fit in FP64, SI copy, normalisation, noise draws, window adds), so results and RNG
//                     draw order are exactly those of doing it inside the loop.
// Why: the FP64 fits + noise (libm log / cos) cost ~60 k cycles on three lanes; inside the lockstep loop every warp of
// the block waited for whichever env had a tap that substep (r2 per-env cycle counters: 6.3 k cycles of tail work and
// ~11 k cycles of waiting per substep).
// The same fit for the three fingers at once, one lane per ROW of the least-squares system (lane f (N + 1) + i = point i
// of finger f): the (N + 1) x 4 matrix lives in registers, every sum runs over the lanes of the finger in row order by
// shuffle, so the arithmetic (operands and order) is exactly read_gauge_q's and the result is bit-identical. The serial
// version kept X / v / y in local memory (FP64, 3 lanes): ncu r2, 5 % of the stall samples of the step at N = 8.
// q3: the 3 N segment angles of a tap record. Returns the fit of finger f on every lane of its group (lanes >= 3 (N + 1): 0).
template <typename T, int N>
__device__ inline float gauge_fit_warp(const T* q3, double L, int lane) {
  constexpr int n = N + 1;
  static_assert(3 * n <= 32, "one lane per row of the three fits");
  const int f = lane / n, i = lane - f * n, base = f * n;
  const bool on = f < 3;
  const T* qf = q3 + (on ? f : 0) * N;
  // point i (i >= 1) sits at the end of segment i - 1: cumulative angle q_0 + ... + q_{i-1}, summed in order
  double cum = 0;
  for (int j = 0; j < N; j++) if (on && j < i) cum += (double)qf[j];
  const double cs = (on && i >= 1) ? L * nd_cos(cum) : 0.0, sn = (on && i >= 1) ? L * nd_sin(cum) : 0.0;
  double fx = 0, y = 0;
  for (int j = 1; j < n; j++) {                       // fx[i] = fx[i - 1] + L cos(cum_i): running sums in index order
    const double cj = __shfl_sync(GMJ_FULL, cs, on ? base + j : lane), sj = __shfl_sync(GMJ_FULL, sn, on ? base + j : lane);
    if (j <= i) { fx += cj; y += sj; }
  }
  double X[4] = { fx * fx * fx, fx * fx, fx, 1.0 };
  auto ordered_sum = [&](double term, int k) {        // sum over rows i >= k of the finger, in row order
    double acc = 0;
    for (int j = 0; j < n; j++) { const double t = __shfl_sync(GMJ_FULL, term, on ? base + j : lane); if (j >= k) acc += t; }
    return acc;
  };
#pragma unroll
  for (int k = 0; k < 4; k++) {                       // Householder QR, column k
    double nrm = sqrt(ordered_sum(X[k] * X[k], k));
    const double xkk = __shfl_sync(GMJ_FULL, X[k], on ? base + k : lane);
    const double alpha = xkk > 0 ? -nrm : nrm;
    double v = (i >= k) ? X[k] : 0.0;
    if (i == k) v -= alpha;
    const double vn = ordered_sum(v * v, k);
    const bool skip = nrm == 0 || vn == 0;            // uniform inside a finger's group
#pragma unroll
    for (int c = k; c < 4; c++) {
      double d = ordered_sum(v * X[c], k);
      d = 2 * d / vn;
      if (!skip && i >= k) X[c] -= d * v;
    }
    double d = ordered_sum(v * y, k);
    d = 2 * d / vn;
    if (!skip && i >= k) y -= d * v;
  }
  // back substitution on the 4 x 4 triangle held by rows 0-3
  double coeff[4];
#pragma unroll
  for (int k = 3; k >= 0; k--) {
    double s = __shfl_sync(GMJ_FULL, y, on ? base + k : lane);
#pragma unroll
    for (int c = k + 1; c < 4; c++) s -= __shfl_sync(GMJ_FULL, X[c], on ? base + k : lane) * coeff[c];
    const double xkk = __shfl_sync(GMJ_FULL, X[k], on ? base + k : lane);
    coeff[k] = (xkk != 0) ? s / xkk : 0;
  }
  float yv = 0.0f;
  const double xp = 50e-3;
  yv = (float)(yv + coeff[0] * (xp * xp * xp));
  yv = (float)(yv + coeff[1] * (xp * xp));
  yv = (float)(yv + coeff[2] * xp);
  yv = (float)(yv + coeff[3] * 1.0);
  return on ? yv * 1000 : 0.0f;
}
#define GMJ_MAXTAP 8
#define GMJ_TAP_HDR 6                 // mask, wrist, palm, axial[3]; then 3 N joint angles
template <typename T, int N, int QC>
__device__ inline void process_taps(EnvState<T, QC>& e, const DevSettings& S, const DevModelS<T>& M, const T* taps, int ntap, int lane) {
  const double* g = S.s.general;
  constexpr int STRIDE = GMJ_TAP_HDR + 3 * N;
  for (int r = 0; r < ntap; r++) {
    const T* rec = taps + (size_t)r * STRIDE;
    const int mask = (int)rec[0];
    float gfit0 = 0.0f, gfit1 = 0.0f, gfit2 = 0.0f;
    if (mask & 1) {                                     // warp-uniform
      if constexpr (3 * (N + 1) <= 32) {
        const float gfit = gauge_fit_warp<T, N>(rec + GMJ_TAP_HDR, (double)M.segment_length, lane);
        gfit0 = __shfl_sync(GMJ_FULL, gfit, 0); gfit1 = __shfl_sync(GMJ_FULL, gfit, N + 1); gfit2 = __shfl_sync(GMJ_FULL, gfit, 2 * (N + 1));
      } else {
        float gfit = 0.0f;
        if (lane < 3) gfit = read_gauge_q<T, N>(rec + GMJ_TAP_HDR + lane * N, (double)M.segment_length);
        gfit0 = __shfl_sync(GMJ_FULL, gfit, 0); gfit1 = __shfl_sync(GMJ_FULL, gfit, 1); gfit2 = __shfl_sync(GMJ_FULL, gfit, 2);
      }
    }
    if (lane != 0) continue;
    if (mask & 1) {
      for (int f = 0; f < 3; f++) {
        float gv = f == 0 ? gfit0 : (f == 1 ? gfit1 : gfit2);
        e.si_latest[f] = (float)((double)gv * g[GS_sim_gauge_raw_to_N_factor]);   // gfloat * double member (src/mjclass.cpp:755-757)
        gv = apply_normalisation(S.s.sensor[GSEN_bending_gauge], gv);
        gv = apply_noise(e, S, GSEN_bending_gauge, gv, f + 1);
        win_add(e, ST_gauge + f, gv);
      }
    }
    if (mask & 2) {
      for (int f = 0; f < 3; f++) {
        float av = (float)rec[3 + f];
        av = apply_normalisation(S.s.sensor[GSEN_axial_gauge], av);
        av = apply_noise(e, S, GSEN_axial_gauge, av, f + 1);
        win_add(e, ST_axial + f, av);
      }
    }
    if (mask & 4) {
      float pv = (float)rec[2];
      pv = (float)(pv * g[GS_palm_scale_factor]);
      e.si_latest[3] = pv;
      pv = apply_normalisation(S.s.sensor[GSEN_palm], pv);
      pv = apply_noise(e, S, GSEN_palm, pv, 1);
      win_add(e, ST_palm, pv);
    }
    if (mask & 8) {
      float z = (float)rec[1];
      z -= (float)S.s.sensor[GSEN_wrist_Z][GSF_raw_value_offset];
      e.si_latest[4] = z;
      z = apply_normalisation(S.s.sensor[GSEN_wrist_Z], z);
      z = apply_noise(e, S, GSEN_wrist_Z, z, 1);
      win_add(e, ST_wristz, z);
    }
  }
  __syncwarp();
}
// ntap: records already in the buffer (kept by the caller, uniform over the warp). A full buffer is processed on the
// spot (settings with a very high read rate), so any setting gives the reference's results.
template <typename T, int N, int QC>
__device__ __forceinline__ void sensor_taps(EnvState<T, QC>& e, const DevSettings& S, const DevModelS<T>& M, const WarpAux<T>& ax,
                                            T* taps, int& ntap, int lane) {
  int due = 0;
  if (lane < 4) {
    const int sen = lane == 0 ? GSEN_bending_gauge : (lane == 1 ? GSEN_axial_gauge : (lane == 2 ? GSEN_palm : GSEN_wrist_Z));
    if (e.time > e.last_read_time[lane] + S.read_between[sen]) { e.last_read_time[lane] = e.time; due = 1; }   // src/mjclass.h:225-241
  }
  const unsigned m = __ballot_sync(GMJ_FULL, due) & 0xfu;
  if (!m) return;
  if (ntap == GMJ_MAXTAP) { __syncwarp(); process_taps<T, N>(e, S, M, taps, ntap, lane); ntap = 0; }
  constexpr int STRIDE = GMJ_TAP_HDR + 3 * N;
  T* rec = taps + (size_t)ntap * STRIDE;
  if (m & 1) for (int i = lane; i < 3 * N; i += 32) { const int f = i / N; rec[GMJ_TAP_HDR + i] = e.qpos[3 + f * (N + 2) + (i - f * N)]; }
  T ax3[3] = { 0, 0, 0 }, palm = 0;
  if ((m & 6) && e.has_object) {
    // ObjectHandler::extract_forces_faster restricted to what the taps read: local x of the force sums on each finger
    // and on the palm over ALL contacts (src/objecthandler.cpp:737-992); lane c takes contact c
    T gx = 0, gy = 0, gz = 0; int ca = 0, cb = 0;
    if (lane < ax.ncon) {
      const DevContact<T>& cn = ax.con[lane];
      ca = cn.cls1; cb = cn.cls2;
#pragma unroll
      for (int k = 0; k < 3; k++) { gx += cn.force[k] * cn.frame[3 * k]; gy += cn.force[k] * cn.frame[3 * k + 1]; gz += cn.force[k] * cn.frame[3 * k + 2]; }
    }
#pragma unroll
    for (int f = 0; f < 4; f++) {
      const int cls = f < 3 ? GMJ_CLS_F1 + f : GMJ_CLS_PALM;
      const bool w = (ca == cls || cb == cls);
      const T* R = f < 3 ? ax.fingerR[f] : ax.palmR;
      // local x component: column 0 of the body's rotation matrix dotted with the world force
      const T v = warp_sum(w ? (R[0] * gx + R[3] * gy + R[6] * gz) : T(0));
      if (f < 3) ax3[f] = v; else palm = v;
    }
  }
  if (lane == 0) { rec[0] = (T)(int)m; rec[1] = ax.wrist_z; rec[2] = palm; rec[3] = ax3[0]; rec[4] = ax3[1]; rec[5] = ax3[2]; }
  ntap++;
  __syncwarp();
}

// ------------------------------------------------------------------ stepper + welds
// Each weld joins two bodies connected by one 1-dof joint, so its anchor (relative pose at
// activation, luke::set_constraint src/myfunctions.cpp:1177-1279) reduces to the joint value at
// the pose of the last kinematics pass (`qfk`). Bits of weld_on: 0-2 pris, 3-5 rev, 6 palm.
template <typename T, int N, int QC>
__device__ inline void set_weld(EnvState<T, QC>& e, int id, int on, const T* qfk) {
  if (on) {
    int j = (id < 3) ? (1 + id * (N + 2)) : ((id < 6) ? (2 + (id - 3) * (N + 2)) : (3 * N + 7));
    e.weld_anchor[id] = qfk[j];
    e.weld_on |= (1 << id);
  } else e.weld_on &= ~(1 << id);
}

// luke::update_all (src/myfunctions.cpp:2110-2284) + apply_antiroll (src/objecthandler.cpp:1034-1059)
template <typename T, int N, int QC>
__device__ inline void update_all(EnvState<T, QC>& e, const T* qfk, const double* asin_tab) {
  const double time_per_step = 10 / 2000.0;
  if (e.time > e.last_step_time + time_per_step) {
    // old_moving bit 3 ("steady"): the previous tick moved no motor, the stepper state was already the
    // canonical image of its step counts and the target has not changed since (set_action clears the
    // bit), so this tick would recompute exactly the same values: only the tick clock advances.
    // bit 4: tnext is the output of g_step_to (canonical).
    if (e.old_moving & 8) e.last_step_time = e.time;
    else {
      int nx = e.tend.sx != e.tnext.sx;
      int nz = e.tend.sz != e.tnext.sz;
      // Target::revolute_moving compares calc_th of both targets; .th holds exactly that (g_update_xy), or the
      // limit angle where y was clamped to it (equal within an ulp, the 0.5 degree test is 30 steps wide)
      double a = (180.0 / 3.14159265358979323846) * e.tend.th;
      double b = (180.0 / 3.14159265358979323846) * e.tnext.th;
      int ny = !(fabs(a - b) < 5e-1);
      int ox = e.old_moving & 1, oy = (e.old_moving >> 1) & 1, oz = (e.old_moving >> 2) & 1;
      if (nx != ox) for (int i = 0; i < 3; i++) set_weld<T, N>(e, i, !nx, qfk);
      if (ny != oy) { if (ny) for (int i = 3; i < 6; i++) set_weld<T, N>(e, i, 0, qfk); }
      if (nz != oz) set_weld<T, N>(e, 6, !nz, qfk);
      const int nomove = e.tend.sx == e.tnext.sx && e.tend.sy == e.tnext.sy && e.tend.sz == e.tnext.sz;
      const int was_canon = (e.old_moving >> 4) & 1;
      e.old_moving = nx | (ny << 1) | (nz << 2) | ((nomove && was_canon) ? 8 : 0) | 16;
      e.last_step_time = e.time;
      g_step_to(e.tnext, e.tend, 10, asin_tab);
    }
  }
  if (e.has_object) {
    T vx = e.qvel[QC - 8], vy = e.qvel[QC - 7], vz = e.qvel[QC - 6];
    if ((double)vx * vx + (double)vy * vy + (double)vz * vz < 1e-12)          // |v| < 1e-6 without the FP64 square root
      for (int k = 0; k < 6; k++) e.qvel[QC - 8 + k] = 0;
  }
}

// ------------------------------------------------------------------ set_action (src/mjclass.cpp:1528-1630)
template <typename T, int QC>
__device__ inline int clamp_base(EnvState<T, QC>& e, const double* g) {
  int wl = 1;
  if (e.base_x > g[GS_base_x_max]) { e.base_x = g[GS_base_x_max]; wl = 0; }
  if (e.base_x < g[GS_base_x_min]) { e.base_x = g[GS_base_x_min]; wl = 0; }
  if (e.base_y > g[GS_base_y_max]) { e.base_y = g[GS_base_y_max]; wl = 0; }
  if (e.base_y < g[GS_base_y_min]) { e.base_y = g[GS_base_y_min]; wl = 0; }
  if (e.base_z > g[GS_base_z_max]) { e.base_z = g[GS_base_z_max]; wl = 0; }
  if (e.base_z < g[GS_base_z_min]) { e.base_z = g[GS_base_z_min]; wl = 0; }
  return wl;
}
template <typename T, int QC>
__device__ inline int call_action(EnvState<T, QC>& e, const DevSettings& S, int a, double v) { // src/mjclass.h:546-571
  v *= (int)S.s.action[a][GAF_sign];
  DevGripper& t = e.tend;
  const double* g = S.s.general;
  switch (a) {
    case GACT_gripper_X: return g_set_xyz_m(t, t.x + v, t.y, t.z);
    case GACT_gripper_Y: return g_set_xyz_m(t, t.x, t.y + v, t.z);
    case GACT_gripper_Z: return g_set_xyz_m(t, t.x, t.y, t.z + v);
    case GACT_gripper_prismatic_X: return g_set_xyz_m_rad(t, t.x + v, t.th, t.z);
    case GACT_gripper_revolute_Y: return g_set_xyz_m_rad(t, t.x, t.th + v, t.z);
    case GACT_base_X: e.base_x += v; return clamp_base(e, g);
    case GACT_base_Y: e.base_y += v; return clamp_base(e, g);
    case GACT_base_Z: e.base_z += v; return clamp_base(e, g);
    default: {
      int wl = 1;
      if (a == GACT_base_yaw) e.base_yaw += v;
      if (e.base_yaw > g[GS_base_yaw_max]) { e.base_yaw = g[GS_base_yaw_max]; wl = 0; }
      if (e.base_yaw < g[GS_base_yaw_min]) { e.base_yaw = g[GS_base_yaw_min]; wl = 0; }
      return wl;
    }
  }
}
// returns number of EXTRA substeps to burn before the action (termination lift, quirk Q8), or -1
// for an out-of-range action
template <typename T, int QC>
__device__ inline int set_action(EnvState<T, QC>& e, const DevSettings& S, const DevModelS<T>& M, int action, float fraction) {
  int wl = 1, extra = 0;
  e.termination_sent = 0;
  if (action < 0 || action >= S.n_actions) return -1;
  e.old_moving &= ~8;                       // the target may change: the stepper tick is no longer a no-op
  int code = S.action_options[action];
  const double* g = S.s.general;
  if (code == 3 * GACT_COUNT) {
    float value = g[GS_continous_actions] ? fraction : 1.0f;
    if (value > (float)g[GS_termination_threshold]) {
      e.termination_sent = 1;
      if (g[GS_lift_on_termination]) {
        e.base_z = -g[GS_base_z_max];
        if (e.base_z > g[GS_base_z_max]) e.base_z = g[GS_base_z_max];
        if (e.base_z < g[GS_base_z_min]) e.base_z = g[GS_base_z_min];
        extra = 2 * (int)g[GS_sim_steps_per_action];
      }
    }
  } else {
    int a = code / 3, kind = code % 3;
    double v = S.s.action[a][GAF_value];
    if (kind == 0) wl = call_action(e, S, a, v);
    else if (kind == 1) wl = call_action(e, S, a, -1 * v);
    else {
      // no clamp here: MjClass::set_action(int, float) takes the fraction as given, only
      // set_continous_action clamps to [-1, 1] first (src/mjclass.cpp:1517-1526; host side: mjclass.py)
      wl = call_action(e, S, a, v * fraction);
      e.ev_value[GBE_COUNT + GLE_action_penalty_lin] += fabsf(fraction);
      e.ev_value[GBE_COUNT + GLE_action_penalty_sq] += fraction * fraction;
    }
  }
  {  // luke::get_fingertip_z_height (src/myfunctions.cpp:3608-3620)
    float straight = (float)(-g[GS_base_z_min] - e.base_z);
    float tip_lift = (float)((double)M.finger_length * (1 - nd_cos(g_calc_th(e.tend.x, e.tend.y))));
    float above = straight + tip_lift;
    float h = (float)((double)above + g[GS_base_z_min]);
    if (h < g[GS_fingertip_min_mm] * 1e-3) wl = 0;
  }
  if (!wl) e.ev_value[GBE_exceed_limits] = 1.0f;
  return extra;
}

// ------------------------------------------------------------------ sense_gripper_state (src/mjclass.cpp:900-964)
__device__ inline float normalise_between(float val, float mn, float mx) {  // src/mjclass.cpp:4896-4904
  if (val > mx) return 1.0f; else if (val < mn) return -1.0f;
  return 2 * (val - mn) / (mx - mn) - 1;
}
template <typename T, int QC>
__device__ inline void sense_gripper_state(EnvState<T, QC>& e, const DevSettings& S, const DevModelS<T>& M) {
  const double* g = S.s.general;
  float sx = normalise_between((float)e.tend.x, (float)49e-3, (float)134e-3);
  float sy = normalise_between((float)e.tend.y, (float)49e-3, (float)134e-3);
  float sz = normalise_between((float)e.tend.z, (float)0e-3, (float)165e-3);
  float bx = normalise_between((float)e.base_x, (float)g[GS_base_x_min], (float)g[GS_base_x_max]);
  float by = normalise_between((float)e.base_y, (float)g[GS_base_y_min], (float)g[GS_base_y_max]);
  float bz = normalise_between((float)e.base_z, (float)g[GS_base_z_min], (float)g[GS_base_z_max]);
  float byaw = normalise_between((float)e.base_yaw, (float)g[GS_base_yaw_min], (float)g[GS_base_yaw_max]);
  sx = apply_noise(e, S, GSEN_motor_state, sx, 1);
  sy = apply_noise(e, S, GSEN_motor_state, sy, 2);
  sz = apply_noise(e, S, GSEN_motor_state, sz, 3);
  bx = apply_noise(e, S, GSEN_base_XY, bx, 1);
  by = apply_noise(e, S, GSEN_base_XY, by, 2);
  bz = apply_noise(e, S, GSEN_base_Z, bz, 1);
  byaw = apply_noise(e, S, GSEN_base_yaw, byaw, 1);
  win_add(e, ST_motor + 0, sx); win_add(e, ST_motor + 1, sy); win_add(e, ST_motor + 2, sz);
  win_add(e, ST_basexy + 0, bx); win_add(e, ST_basexy + 1, by);
  win_add(e, ST_basez, bz); win_add(e, ST_yaw, byaw);
  // MAT cartesian contact points: luke::get_fingerend_and_palm_xyz (src/myfunctions.cpp:3622-3688)
  double Fg[3] = { e.si_latest[0], e.si_latest[1], e.si_latest[2] };
  double palmF = e.si_latest[3];
  double fing_x = e.tend.x, fing_th = g_calc_th(e.tend.x, e.tend.y), palm_z = e.tend.z;
  double L = (double)M.finger_length;
  double untilted = (double)M.clearance - e.base_z;
  double tilted = untilted + L * (1 - nd_cos(fing_th));
  const double ft = 0.2, PI_23 = 3.14159265358979323846 * (2.0 / 3.0);
  for (int i = 0; i < 3; i++) {
    double ang = i * PI_23;
    double tilt_x = fing_x - L * nd_sin(fing_th);
    double defl = Fg[i] * (L * L * L) / (3 * (double)M.EI);
    double fx = tilt_x + defl * nd_cos(fing_th);
    double px = -fx * nd_sin(ang + e.base_yaw) + e.base_x;
    double py = -fx * nd_cos(ang + e.base_yaw) + e.base_y;
    double pz = tilted - defl * nd_sin(fing_th);
    int on = fabs(Fg[i]) > ft;
    win_add(e, ST_cart + 3 * i + 0, on ? (float)px : 0.0f);
    win_add(e, ST_cart + 3 * i + 1, on ? (float)py : 0.0f);
    win_add(e, ST_cart + 3 * i + 2, on ? (float)pz : 0.0f);
  }
  int on = palmF > ft;
  win_add(e, ST_cart + 9, on ? (float)e.base_x : 0.0f);
  win_add(e, ST_cart + 10, on ? (float)e.base_y : 0.0f);
  win_add(e, ST_cart + 11, on ? (float)(untilted + 165e-3 - palm_z) : 0.0f);
}

// ------------------------------------------------------------------ events, reward, done, observation
#define EVB(k) (k)
#define EVL(k) (GBE_COUNT + (k))
__device__ inline float linear_reward(float val, float mn, float mx, float overshoot) { // src/mjclass.cpp:4866-4894
  if (val < mn) return 0.0f;
  if (val > mx) {
    if (overshoot < mx) return 1.0f;
    if (val > overshoot) return 0.0f;
    mn = 0; mx = overshoot - mx; val = overshoot - val;
  }
  return (val - mn) / (mx - mn);
}
template <typename T, int QC>
__device__ inline void update_events(EnvState<T, QC>& e, const DevSettings& S) {     // src/mjclass.cpp:5437-5469
  for (int k = 0; k < GBE_COUNT; k++) {
    int v = e.ev_value[k] != 0;
    e.ev_row[k] = e.ev_row[k] * v + v;
    e.ev_abs[k] += v;
    e.ev_last[k] = (float)v;
    e.ev_value[k] = 0;
  }
  for (int k = 0; k < GLE_COUNT; k++) {
    float mn = (float)S.s.linear[k][GLF_min], ov = (float)S.s.linear[k][GLF_overshoot];
    float val = e.ev_value[EVL(k)];
    int active = (val > mn && (val < ov || ov < 0)) ? 1 : 0;
    e.ev_row[EVL(k)] = e.ev_row[EVL(k)] * active + active;
    e.ev_abs[EVL(k)] += active;
    e.ev_last[EVL(k)] = val;
    e.ev_value[EVL(k)] = 0.0f;
  }
}
__device__ inline float mag3d(const double* v) { return (float)sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }

// MjClass::update_env (src/mjclass.cpp:966-1346), single live object
template <typename T, int QC>
__device__ inline void update_env(EnvState<T, QC>& e, const DevSettings& S, const WarpAux<T>& ax) {
  const double* g = S.s.general;
  const float ftol = 1e-5f;
  float* V = e.ev_value;
  if (!e.has_object) { V[GBE_step_num] = 1; update_events(e, S); return; }
  DevForces F;
  extract_forces(e, ax, F);
  // x, y: the double-integrated copy (FP32 storage of a 2 mm grid pose would decide the 10 mm threshold by rounding)
  double qx = sizeof(T) == 8 ? (double)e.qpos[QC - 8] : e.obj_xy[0], qy = sizeof(T) == 8 ? (double)e.qpos[QC - 7] : e.obj_xy[1], qz = (double)e.qpos[QC - 6];
  double rx = qx - e.base_x, ry = qy - e.base_y;
  float distance = (float)sqrt(rx * rx + ry * ry);
  float f1 = mag3d(F.obj_f_loc[0]), f2 = mag3d(F.obj_f_loc[1]), f3 = mag3d(F.obj_f_loc[2]);
  float pm = mag3d(F.obj_palm_loc), gm = mag3d(F.obj_gnd);
  float palm_axial = (float)F.obj_palm_loc[0];
  float lift_height = (float)(qz - e.start_z);
  float avg_finger = (float)(0.33333 * (f1 + f2 + f3));
  float peak_lat = (float)(-1 * fmin(F.obj_f_loc[0][1], fmin(F.obj_f_loc[1][1], F.obj_f_loc[2][1])));
  float ov_avg = 0, ov_palm = 0, ov_lift = 0;
  if (avg_finger > ov_avg) ov_avg = avg_finger;
  if (palm_axial > ov_palm) ov_palm = palm_axial;
  if (peak_lat > e.peak_lateral) e.peak_lateral = peak_lat;
  if (lift_height > ov_lift) ov_lift = lift_height;
  float gripper_z_height = (float)(-1 * e.base_z);
  float peak_axial = (float)(-1 * fmin(F.gnd_f_loc[0][0], fmin(F.gnd_f_loc[1][0], F.gnd_f_loc[2][0])));
  float g1 = e.si_latest[0], g2 = e.si_latest[1], g3 = e.si_latest[2];
  float last_palm = e.si_latest[3], last_wrist = e.si_latest[4];
  float max_gauge = fmaxf(fmaxf(g1, g2), g3);
  float avg_gauge = (float)((1.0 / 3.0) * (g1 + g2 + g3));
  V[GBE_step_num] = 1;
  int lifted = 0, oob = 0, l2h = 0, th = 0, stable = 0;
  if (gm < ftol && 0.0f < ftol) { V[GBE_lifted] = 1; lifted = 1; }          // quirk Q4
  double od = g[GS_oob_distance];
  if (qx > od || qx < -od || qy > od || qy < -od) { V[GBE_oob] = 1; oob = 1; }
  if (ov_lift > g[GS_lift_height] - ftol && lifted && !oob) { V[GBE_lifted_to_height] = 1; l2h = 1; }
  if (l2h && gripper_z_height > g[GS_gripper_target_height] - ftol) { V[GBE_target_height] = 1; th = 1; }
  if (f1 > ftol || f2 > ftol || f3 > ftol || pm > ftol) V[GBE_object_contact] = 1;
  if (f1 > g[GS_stable_finger_force] && f2 > g[GS_stable_finger_force] && f3 > g[GS_stable_finger_force] &&
      f1 < g[GS_stable_finger_force_lim] && f2 < g[GS_stable_finger_force_lim] && f3 < g[GS_stable_finger_force_lim] &&
      pm > g[GS_stable_palm_force] && pm < g[GS_stable_palm_force_lim] && V[GBE_lifted] != 0) {
    V[GBE_object_stable] = 1; stable = 1;
  }
  if (stable && th) V[GBE_stable_height] = 1;
  if (e.termination_sent) {
    if (S.s.binary[GBE_lifted_termination][GBF_done]) {
      if (l2h) V[GBE_lifted_termination] = 1; else V[GBE_failed_termination] = 1;
    } else {
      if (V[GBE_stable_height] != 0) V[GBE_stable_termination] = 1; else V[GBE_failed_termination] = 1;
    }
  }
  // The 2 mm spawn and base grids put every tenth object EXACTLY on the 10 mm threshold, where the reference's float
  // cast of the distance decides "within" (0.01f < 0.01). An FP32 resting object carries ~1e-9 m of solver noise, so the
  // FP32 build takes ties within 1e-7 m the way the reference takes the exact tie.
  if ((double)distance < g[GS_XY_distance_threshold] + (sizeof(T) == 8 ? 0.0 : 1e-7)) V[GBE_within_XY_distance] = 1;
  {  // dropped (src/mjclass.cpp:1249-1255, quirk Q7)
    int dr = e.ev_row[GBE_dropped], lv = V[GBE_lifted] != 0, lr = e.ev_row[GBE_lifted];
    int v = ((!dr) * (!lv) * lr) ? 1 : (lv ? 0 : (dr ? dr + 1 : 0));
    V[GBE_dropped] = (v != 0) ? 1.0f : 0.0f;
  }
  V[EVL(GLE_exceed_axial)] = peak_axial;
  V[EVL(GLE_exceed_lateral)] = e.peak_lateral;
  V[EVL(GLE_palm_force)] = ov_palm * (V[GBE_lifted] != 0);
  V[EVL(GLE_exceed_palm)] = ov_palm;
  V[EVL(GLE_finger_force)] = ov_avg;
  V[EVL(GLE_finger1_force)] = f1; V[EVL(GLE_finger2_force)] = f2; V[EVL(GLE_finger3_force)] = f3;
  V[EVL(GLE_ground_force)] = gm;
  V[EVL(GLE_good_bend_sensor)] = avg_gauge;
  V[EVL(GLE_exceed_bend_sensor)] = max_gauge; V[EVL(GLE_dangerous_bend_sensor)] = max_gauge;
  V[EVL(GLE_good_palm_sensor)] = last_palm; V[EVL(GLE_exceed_palm_sensor)] = last_palm;
  V[EVL(GLE_dangerous_palm_sensor)] = last_palm;
  V[EVL(GLE_exceed_wrist_sensor)] = last_wrist; V[EVL(GLE_dangerous_wrist_sensor)] = last_wrist;
  float denom = (float)(S.n_actions - (int)g[GS_use_termination_action]);
  V[EVL(GLE_action_penalty_lin)] /= denom;
  V[EVL(GLE_action_penalty_sq)] /= denom;
  V[EVL(GLE_object_XY_distance)] = (float)(-(double)distance);
  for (int k = 0; k < GBE_COUNT; k++)
    if (V[k] != 0 && (float)S.s.binary[k][GBF_reward] >= (1.0 - 1e-5) && S.s.binary[k][GBF_done] &&
        e.ev_row[k] + 1 >= (int)S.s.binary[k][GBF_trigger])
      V[GBE_successful_grasp] = 1;
  update_events(e, S);
}

// calc_rewards + MjClass::reward (src/mjclass.cpp:5471-5528, 3000-3049)
template <typename T, int QC>
__device__ inline float calc_reward(EnvState<T, QC>& e, const DevSettings& S) {
  float reward = 0;
  for (int k = 0; k < GBE_COUNT; k++)
    if (e.ev_row[k] >= (int)S.s.binary[k][GBF_trigger]) reward += (float)S.s.binary[k][GBF_reward];
  for (int k = 0; k < GLE_COUNT; k++)
    if (e.ev_row[EVL(k)] >= (int)S.s.linear[k][GLF_trigger]) {
      float fr = linear_reward(e.ev_last[EVL(k)], (float)S.s.linear[k][GLF_min], (float)S.s.linear[k][GLF_max],
                               (float)S.s.linear[k][GLF_overshoot]);
      float sc = (float)S.s.linear[k][GLF_reward] * fr;
      reward += sc;
    }
  float lo = (float)S.s.general[GS_reward_cap_lower_bound], hi = (float)S.s.general[GS_reward_cap_upper_bound];
  e.cumulative_reward += reward;
  if (e.cumulative_reward < lo && S.s.general[GS_cap_reward]) { reward += lo - e.cumulative_reward; e.cumulative_reward = lo; }
  if (e.cumulative_reward > hi && S.s.general[GS_cap_reward]) { reward += hi - e.cumulative_reward; e.cumulative_reward = hi; }
  return reward;
}
// MjClass::is_done (src/mjclass.cpp:1632-1698)
template <typename T, int QC>
__device__ inline int calc_done(const EnvState<T, QC>& e, const DevSettings& S) {
  const float ftol = 1e-5f;
  if (e.unstable) return 1;
  for (int k = 0; k < GBE_COUNT; k++) { int d = (int)S.s.binary[k][GBF_done]; if (d && e.ev_row[k] >= d) return 1; }
  for (int k = 0; k < GLE_COUNT; k++) { int d = (int)S.s.linear[k][GLF_done]; if (d && e.ev_row[EVL(k)] >= d) return 1; }
  if (S.s.general[GS_cap_reward] && S.s.general[GS_quit_if_cap_exceeded]) {
    if (e.cumulative_reward - ftol < (float)S.s.general[GS_reward_cap_lower_bound]) return 1;
    if (e.cumulative_reward + ftol > (float)S.s.general[GS_reward_cap_upper_bound]) return 1;
  }
  return 0;
}
// EventTrack::vectorise (src/mjclass.cpp:5594-5615)
template <typename T, int QC>
__device__ inline void event_state(const EnvState<T, QC>& e, float* out) {
  int n = 0;
  for (int k = 0; k < GBE_COUNT; k++) out[n++] = (float)e.ev_row[k];
  for (int k = 0; k < GLE_COUNT; k++) { out[n++] = (float)e.ev_row[EVL(k)]; out[n++] = e.ev_last[EVL(k)]; }
}
// MjClass::get_observation (src/mjclass.cpp:1700-1959)
template <typename T, int QC>
__device__ inline int get_observation(const EnvState<T, QC>& e, const DevSettings& S, float* obs) {
  const GmjSettings& s = S.s;
  int n = 0;
  int smode = (int)s.general[GS_sensor_sample_mode], tmode = (int)s.general[GS_state_sample_mode];
  if (tmode == GSM_scaled_change_sq) { smode = GSM_scaled_change_sq; tmode = GSM_sign; }   // quirk Q2
#define SAMP(SEN, MODE, ST) n += sample_stream(e, ST, MODE, S.prev_steps[SEN], S.readings_per_step[SEN], obs + n)
  if (s.sensor[GSEN_bending_gauge][GSF_in_use]) for (int k = 0; k < 3; k++) SAMP(GSEN_bending_gauge, smode, ST_gauge + k);
  if (s.sensor[GSEN_axial_gauge][GSF_in_use]) for (int k = 0; k < 3; k++) SAMP(GSEN_axial_gauge, smode, ST_axial + k);
  if (s.sensor[GSEN_palm][GSF_in_use]) SAMP(GSEN_palm, smode, ST_palm);
  if (s.sensor[GSEN_wrist_XY][GSF_in_use]) {     // wrist X/Y are never written (commented out in the reference): zeros
    int len = (smode == GSM_raw) ? S.prev_steps[GSEN_wrist_XY] * S.readings_per_step[GSEN_wrist_XY] : 2 * S.prev_steps[GSEN_wrist_XY] + 1;
    for (int k = 0; k < 2 * len; k++) obs[n++] = 0.0f;
  }
  if (s.sensor[GSEN_wrist_Z][GSF_in_use]) SAMP(GSEN_wrist_XY, smode, ST_wristz);  // quirk Q2: XY sensor object
  if (s.sensor[GSEN_motor_state][GSF_in_use]) for (int k = 0; k < 3; k++) SAMP(GSEN_motor_state, tmode, ST_motor + k);
  if (s.sensor[GSEN_base_XY][GSF_in_use]) for (int k = 0; k < 2; k++) SAMP(GSEN_base_XY, tmode, ST_basexy + k);
  if (s.sensor[GSEN_base_Z][GSF_in_use]) SAMP(GSEN_base_Z, tmode, ST_basez);
  if (s.sensor[GSEN_base_yaw][GSF_in_use]) SAMP(GSEN_base_yaw, tmode, ST_yaw);
  if (s.sensor[GSEN_cartesian][GSF_in_use]) for (int k = 0; k < 12; k++) SAMP(GSEN_cartesian, GSM_change, ST_cart + k);
#undef SAMP
  return n;
}
