// gmj_api.cu -- kernels + C ABI (include/gmj.h) of libgripper_b200.so.
// Build: see gripper_mujoco_b200/lib.py (one translation unit per step-kernel instantiation, linked into one .so)
#include "gmj_step.cuh"

static thread_local std::string g_err;
void gmj_set_error(const std::string& msg) { g_err = msg; }

// Every entry point makes the handle's device current for its own calls and puts the CALLER'S device back on return
// (torch and other users of the runtime API rely on the current device staying where they left it).
struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; if (prev != dev) cudaSetDevice(dev); else prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// counting-sort scatter by load key, heaviest first: envs that owe no lift substeps into perm, the lifting envs
// into env_list. hist = [counts | cursors | lifter counts | lifter cursors], GMJ_SORT_BUCKETS each.
template <typename T, int QC>
__global__ void k_sort_scatter(const EnvState<T, QC>* __restrict__ envs, int n_envs, int* __restrict__ hist,
                               int* __restrict__ perm, int* __restrict__ env_list, int sort_lift) {
  int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= n_envs) return;
  const bool lift = envs[env].pending_extra > 0;
  const int* cnt = hist + (lift ? 2 * GMJ_SORT_BUCKETS : 0);
  int* cursor = hist + (lift ? 3 * GMJ_SORT_BUCKETS : GMJ_SORT_BUCKETS);
  // small launches keep the lifting envs in arrival order: the step is as long as the slowest lift block, and a block
  // of equally heavy envs is slower than one heavy env among light ones (measured: +5 % step time at 4096 envs sorted,
  // -3.5 % at 16384)
  const int key = (lift && !sort_lift) ? GMJ_SORT_BUCKETS - 1 : (int)envs[env].sort_key;
  int off = 0;
  for (int k = GMJ_SORT_BUCKETS - 1; k > key; k--) off += cnt[k];
  (lift ? env_list : perm)[off + atomicAdd(&cursor[key], 1)] = env;
}

// MjClass::reset + luke::reset (src/mjclass.cpp:434-486, src/myfunctions.cpp:544-570): one thread per env
template <typename T, int QC>
__global__ void k_reset(EnvState<T, QC>* envs, const DevModelS<T>* Mp, const DevSettings* Sp, int n_envs,
                        const uint8_t* mask, uint32_t env_id_offset, int step_cap, uint8_t* mask_out) {
  constexpr int OQ = QC - 8;
  int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= n_envs) return;
  EnvState<T, QC>& e = envs[env];
  // gmj_reset_done: the env resets when its done flag is set OR its episode reached step_cap action steps (MjEnv.step's
  // truncation, rl/env/MjEnv.py:616-637); the resolved mask goes out for the spawn that follows
  bool go = !mask || mask[env];
  if (step_cap > 0) go = (mask && mask[env]) || e.num_action_steps >= step_cap;
  if (mask_out) mask_out[env] = go ? 1 : 0;
  if (!go) return;
  const DevModelS<T>& M = *Mp; const DevSettings& S = *Sp;
  const int N = M.N, NL = N + 2, B = M.B;
  uint32_t episode = e.episode + 1;
  unsigned long long c0 = e.st_contacts, c1 = e.st_rows, c2 = e.st_iters, c3 = e.st_substeps; unsigned int c4 = e.st_unstable + (e.unstable ? 1 : 0), c5 = e.st_overflow, c6 = e.st_relaxed, c7 = e.st_capped;
  memset(&e, 0, sizeof(EnvState<T, QC>));
  e.st_contacts = c0; e.st_rows = c1; e.st_iters = c2; e.st_substeps = c3; e.st_unstable = c4; e.st_overflow = c5; e.st_relaxed = c6; e.st_capped = c7;
  e.episode = episode; e.env_id = env_id_offset + env; e.rng_ctr = episode << 12;
  g_reset(e.tend); g_reset(e.tnext);
  for (int i = 0; i < B; i++) e.qpos[i] = M.qpos0[i];
  // parked object (dead collision group): not simulated
  e.obj_xy[0] = 0; e.obj_xy[1] = 2;
  e.qpos[OQ] = 0; e.qpos[OQ + 1] = 2; e.qpos[OQ + 2] = 1; e.qpos[OQ + 3] = 0; e.qpos[OQ + 4] = 0; e.qpos[OQ + 5] = 0; e.qpos[OQ + 6] = 1;
  // weld anchors from the keyframe pose (xpos/xmat are stale when set_all_constraints runs), then
  // the stored equilibrium of base + motor joints (calibrate_reset)
  for (int f = 0; f < 3; f++) { e.weld_anchor[f] = e.qpos[1 + f * NL]; e.weld_anchor[3 + f] = e.qpos[2 + f * NL]; }
  e.weld_anchor[6] = e.qpos[B - 1];
  e.weld_on = 0x7f; e.old_moving = 0;
  const double* g = S.s.general;
  e.qpos[0] = (T)g[GS_reset_qpos_base];
  for (int f = 0; f < 3; f++) { e.qpos[1 + f * NL] = (T)g[GS_reset_qpos_g0 + 2 * f]; e.qpos[2 + f * NL] = (T)g[GS_reset_qpos_g1 + 2 * f]; }
  e.qpos[B - 1] = (T)g[GS_reset_qpos_g6];
  for (int s = 0; s < GMJ_NSTREAM; s++) e.win_i[s] = -1;
  // Settings::apply_noise_params (src/mjclass.cpp:5293-5346)
  for (int k = 0; k < GSEN_COUNT; k++) {
    float mu = (float)(S.s.sensor[k][GSF_noise_overriden] ? S.s.sensor[k][GSF_noise_mu] : g[GS_sensor_noise_mu]);
    for (int j = 0; j < 3; j++) e.rand_mu[k][j] = mu * (2 * dev_uniform(e, S) - 1);
  }
  const int state_sensors[5] = { GSEN_motor_state, GSEN_base_XY, GSEN_base_Z, GSEN_base_yaw, GSEN_cartesian };
  for (int qn = 0; qn < 5; qn++) {
    int k = state_sensors[qn];
    float mu = (float)(S.s.sensor[k][GSF_noise_overriden] ? S.s.sensor[k][GSF_noise_mu] : g[GS_state_noise_mu]);
    for (int j = 0; j < 3; j++) e.rand_mu[k][j] = mu * (2 * dev_uniform(e, S) - 1);
  }
  // random_base_Z_movement (src/mjclass.cpp:1423-1434)
  double size = g[GS_base_position_noise];
  double zmove = (g[GS_use_noise_seeded] && size > 0) ? (-size + 2 * size * (double)dev_uniform(e, S)) : 0.0;
  float zf = (float)zmove;
  e.base_z = zf;
  if (e.base_z > g[GS_base_z_max]) e.base_z = g[GS_base_z_max];
  if (e.base_z < g[GS_base_z_min]) e.base_z = g[GS_base_z_min];
  e.qpos[0] = (T)(e.base_z + g[GS_reset_qpos_base]);
}

// luke::Box2d::overlapsWith (src/customtypes.h:86-140)
__device__ inline void box2d_init(double* b, double cx, double cy, double w, double h, double theta) {
  double hw = w / 2.0, hh = h / 2.0;
  b[0] = cx - hw; b[4] = cy - hh; b[1] = cx + hw; b[5] = cy - hh; b[2] = cx + hw; b[6] = cy + hh; b[3] = cx - hw; b[7] = cy + hh;
  double mx = (b[0] + b[1] + b[2] + b[3]) / 4.0, my = (b[4] + b[5] + b[6] + b[7]) / 4.0;
  for (int i = 0; i < 4; i++) {
    double nx = mx + (b[i] - mx) * cos(theta) - (b[4 + i] - my) * sin(theta);
    double ny = my + (b[i] - mx) * sin(theta) + (b[4 + i] - my) * cos(theta);
    b[i] = nx; b[4 + i] = ny;
  }
}
__device__ inline int box2d_overlap(const double* a, const double* o, double min_distance) {
  int contains_other = 1;
  for (int i = 0; i < 4; i++) {
    int j = (i + 1) % 4;
    double ex = a[j] - a[i], ey = a[4 + j] - a[4 + i];
    double px = -ey, py = ex, len = sqrt(px * px + py * py);
    px /= len; py /= len;
    double min1 = a[0] * px + a[4] * py, max1 = min1, min2 = o[0] * px + o[4] * py, max2 = min2;
    for (int k = 1; k < 4; k++) {
      double p1 = a[k] * px + a[4 + k] * py, p2 = o[k] * px + o[4 + k] * py;
      if (p1 < min1) min1 = p1; if (p1 > max1) max1 = p1; if (p2 < min2) min2 = p2; if (p2 > max2) max2 = p2;
    }
    if (max1 + min_distance < min2 || max2 + min_distance < min1) return 0;
    if (max1 < min2 || max2 < min1) contains_other = 0;
  }
  return contains_other;
}

// MjClass::spawn_object / spawn_into_scene (src/mjclass.cpp:2352-2420,2475-2654)
// Given pose (xyr != NULL): spawn_object, placed unconditionally. Otherwise the reference's search: candidate grid of
// (x, y) in 2 mm steps within +-xy_range and of rotations in pi/30 steps within +-rot_range, BOTH lists shuffled, walked
// pairwise for max(n_xy, n_rot) tries; the first pose that is inside the outer bounds and whose rotated footprint
// clears the three fingertip boxes (taken at reset: target at home, base at the origin, src/mjclass.h:893-904) by 1 mm
// spawns; when none does the env stays without object and ok[env] = 0 -- the caller retries as MjEnv does
// (rl/env/MjEnv.py:1216-1244). std::shuffle on the process-global engine becomes a Fisher-Yates walk on the counter RNG
// (oracle/env.c or_spawn_into_scene draws the same numbers).
#define GMJ_SPAWN_MAXXY 1024
#define GMJ_SPAWN_MAXROT 128
// One WARP per env: the Fisher-Yates draws (counter RNG, draw k = philox(ctr + k)) are formed by all lanes, the swaps
// applied in order by lane 0, and the candidate walk tests 32 poses at a time (first passing index wins, so the result
// equals the serial walk). One thread per env made the unlucky env (large object, 3 x 121 FP64 SAT tests) a ~1 ms
// serial tail on every bench step.
#define GMJ_SPAWN_WPB 4
__device__ __forceinline__ float spawn_draw(uint32_t ctr, uint32_t env_id, uint64_t seed) {
  uint32_t x = philox_first(ctr, env_id, 0, 0, (uint32_t)seed, (uint32_t)(seed >> 32));
  return (float)(x >> 8) * (1.0f / 16777216.0f);
}
__device__ inline void fy_shuffle_warp(uint32_t& ctr, uint32_t env_id, uint64_t seed, unsigned short* a, unsigned short* js, int n, int lane) {
  for (int i = lane; i < n; i += 32) a[i] = (unsigned short)i;
  for (int k = lane; k < n - 1; k += 32) {              // draw k belongs to i = n - 1 - k
    const int i = n - 1 - k;
    int j = (int)(spawn_draw(ctr + k, env_id, seed) * (float)(i + 1)); if (j > i) j = i;
    js[i] = (unsigned short)j;
  }
  ctr += n - 1;
  __syncwarp();
  if (lane == 0) for (int i = n - 1; i >= 1; i--) { const int j = js[i]; unsigned short t = a[i]; a[i] = a[j]; a[j] = t; }
  __syncwarp();
}
template <typename T, int QC>
__global__ void __launch_bounds__(32 * GMJ_SPAWN_WPB)
k_spawn(EnvState<T, QC>* envs, const DevModelS<T>* Mp, const DevSettings* Sp, const GmjObject* objects, int n_objects,
        int n_envs, const int32_t* idx, const float* xyr, const uint8_t* mask, double xy_range, double rot_range,
        int tries_max, int fb_xy_mm, int fb_rot_deg, uint8_t* ok_out, int* err_flag) {
  __shared__ unsigned short s_pxy[GMJ_SPAWN_WPB][GMJ_SPAWN_MAXXY], s_js[GMJ_SPAWN_WPB][GMJ_SPAWN_MAXXY], s_pr[GMJ_SPAWN_WPB][GMJ_SPAWN_MAXROT];
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int env = blockIdx.x * GMJ_SPAWN_WPB + wib;
  if (env >= n_envs) return;
  if (mask && !mask[env]) return;
  constexpr int OQ = QC - 8;
  EnvState<T, QC>& e = envs[env];
  const DevSettings& S = *Sp; const DevModelS<T>& M = *Mp;
  const uint64_t seed = (uint64_t)S.s.general[GS_random_seed];
  const uint32_t env_id = e.env_id;
  uint32_t ctr = e.rng_ctr;                      // every lane keeps the same counter; lane 0 stores it at the end
  int oi = 0; double x = 0, y = 0, rot = 0;
  int found = 0;
  if (idx) {
    oi = idx[env];
    if (oi < 0 || oi >= n_objects) { if (lane == 0) { ok_out[env] = 0; if (err_flag) atomicExch(err_flag, 2); } return; }
  }
  if (idx && xyr) { x = xyr[3 * env]; y = xyr[3 * env + 1]; rot = xyr[3 * env + 2]; found = 1; }
  else {
    if (!idx) { oi = (int)(spawn_draw(ctr++, env_id, seed) * n_objects); if (oi >= n_objects) oi = n_objects - 1; }
    const GmjObject& ob = objects[oi];
    const double inc = 2e-3, rinc = 3.14159265358979323846 / 30.0, gap = 1e-3;
    const int num_x = (int)((2 * xy_range) / inc) + 1, num_y = num_x, num_r = (int)((2 * rot_range) / rinc) + 1;
    const int nxy = num_x * num_y;
    if (nxy > GMJ_SPAWN_MAXXY || num_r > GMJ_SPAWN_MAXROT) {
      if (lane == 0) { ok_out[env] = 0; e.rng_ctr = ctr; if (err_flag) atomicExch(err_flag, 3); }
      return;
    }
    double tips[3][8];
    {
      DevGripper home; g_reset(home);
      const double hook_th = (double)M.hook_angle_deg * (3.14159265358979323846 / 180.000);
      const double hookx = 0.5 * (double)M.hook_length * sin(hook_th);
      for (int i = 0; i < 3; i++) {
        double ang = i * (3.14159265358979323846 * (2.0 / 3.0));
        double px = -(home.x - hookx) * sin(ang) + 0.0, py = -(home.x - hookx) * cos(ang) + 0.0;
        box2d_init(tips[i], px, py, (double)M.finger_width, (double)M.hook_length, -ang);
      }
    }
    unsigned short* pxy = s_pxy[wib]; unsigned short* pr = s_pr[wib]; unsigned short* js = s_js[wib];
    const int tries = nxy > num_r ? nxy : num_r;
    // MjEnv._spawn_object: up to tries_max calls of spawn_into_scene (each with fresh shuffles), rl/env/MjEnv.py:1216-1227
    for (int attempt = 0; attempt < tries_max && !found; attempt++) {
      if (nxy > 1) fy_shuffle_warp(ctr, env_id, seed, pxy, js, nxy, lane); else { if (lane == 0) pxy[0] = 0; __syncwarp(); }
      if (num_r > 1) fy_shuffle_warp(ctr, env_id, seed, pr, js, num_r, lane); else { if (lane == 0) pr[0] = 0; __syncwarp(); }
      for (int base = 0; base < tries && !found; base += 32) {
        const int i = base + lane;
        int ok = 0;
        if (i < tries) {
          const int k = pxy[i % nxy], ix = k / num_y, iy = k % num_y, ir = pr[i % num_r];
          x = num_x > 1 ? -xy_range + ix * inc + 0.0 : 0.0;
          y = num_y > 1 ? -xy_range + iy * inc + 0.0 : 0.0;
          rot = num_r > 1 ? -rot_range + ir * rinc + 0.0 : 0.0;
          double ours[8];
          box2d_init(ours, x, y, ob.bbox[0], ob.bbox[1], rot);
          ok = 1;
          for (int c = 0; c < 4; c++) if (ours[c] < -100 || ours[c] > 100 || ours[4 + c] < -100 || ours[4 + c] > 100) ok = 0;
          for (int t = 0; t < 3 && ok; t++) if (box2d_overlap(ours, tips[t], gap)) ok = 0;
        }
        const unsigned hit = __ballot_sync(0xffffffffu, ok);
        if (hit) {
          const int src = __ffs(hit) - 1;
          x = __shfl_sync(0xffffffffu, x, src); y = __shfl_sync(0xffffffffu, y, src); rot = __shfl_sync(0xffffffffu, rot, src);
          found = 1;
        }
      }
      __syncwarp();
    }
    if (!found) {
      if (fb_xy_mm < 0) { if (lane == 0) { ok_out[env] = 0; e.rng_ctr = ctr; } return; }
      // "old method" of MjEnv._spawn_object (rl/env/MjEnv.py:1229-1244): a pose on the millimetre grid with one of the
      // base angles 0 / 60 / 120 degrees +- fb_rot_deg, placed without a collision test
      int xm = (int)(spawn_draw(ctr++, env_id, seed) * (float)(2 * fb_xy_mm + 1)); if (xm > 2 * fb_xy_mm) xm = 2 * fb_xy_mm;
      int ym = (int)(spawn_draw(ctr++, env_id, seed) * (float)(2 * fb_xy_mm + 1)); if (ym > 2 * fb_xy_mm) ym = 2 * fb_xy_mm;
      int an = (int)(spawn_draw(ctr++, env_id, seed) * (float)(2 * fb_rot_deg + 1)); if (an > 2 * fb_rot_deg) an = 2 * fb_rot_deg;
      int bi = (int)(spawn_draw(ctr++, env_id, seed) * 3.0f); if (bi > 2) bi = 2;
      x = (xm - fb_xy_mm) * 1e-3; y = (ym - fb_xy_mm) * 1e-3;
      rot = (60 * bi + an - fb_rot_deg) * (3.14159265358979323846 / 180.0);
      found = 2;
    }
  }
  if (lane != 0) return;
  ok_out[env] = (uint8_t)found;
  e.rng_ctr = ctr;
  const GmjObject& ob = objects[oi];
  e.obj[0] = (T)ob.type; e.obj[1] = (T)ob.size[0]; e.obj[2] = (T)ob.size[1]; e.obj[3] = (T)ob.size[2]; e.obj[4] = (T)ob.round;
  e.obj[5] = (T)ob.quat[0]; e.obj[6] = (T)ob.quat[1]; e.obj[7] = (T)ob.quat[2]; e.obj[8] = (T)ob.quat[3];
  e.obj[9] = (T)ob.mass; e.obj[10] = (T)ob.inertia[0]; e.obj[11] = (T)ob.inertia[1]; e.obj[12] = (T)ob.inertia[2];
  e.obj[13] = (T)ob.friction; e.obj[14] = (T)ob.z_rest;
  e.obj_index = oi; e.has_object = 1;
  double z = ob.z_rest + 1e-6;
  e.qpos[OQ] = (T)x; e.qpos[OQ + 1] = (T)y; e.qpos[OQ + 2] = (T)z;
  e.obj_xy[0] = x; e.obj_xy[1] = y;
  e.qpos[OQ + 3] = (T)sin(-rot / 2.0); e.qpos[OQ + 4] = 0; e.qpos[OQ + 5] = 0; e.qpos[OQ + 6] = (T)cos(-rot / 2.0);
  for (int k = 0; k < 6; k++) e.qvel[OQ + k] = 0;
  e.start_z = (double)(T)z;
  e.nrow_prev = 0; e.mpr_prev = 0;
}

template <typename T, int QC>
__global__ void k_dump(const EnvState<T, QC>* envs, int n_envs, int B, double* qpos, double* qvel, double* probe) {
  constexpr int OQ = QC - 8;
  int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= n_envs) return;
  const EnvState<T, QC>& e = envs[env];
  if (qpos) { for (int i = 0; i < B; i++) qpos[(size_t)env * (B + 7) + i] = (double)e.qpos[i]; for (int k = 0; k < 7; k++) qpos[(size_t)env * (B + 7) + B + k] = (double)e.qpos[OQ + k]; }
  if (qvel) { for (int i = 0; i < B; i++) qvel[(size_t)env * (B + 6) + i] = (double)e.qvel[i]; for (int k = 0; k < 6; k++) qvel[(size_t)env * (B + 6) + B + k] = (double)e.qvel[OQ + k]; }
  if (probe) {
    double* p = probe + (size_t)env * GMJ_PROBE_LEN;
    for (int k = 0; k < GMJ_PROBE_LEN; k++) p[k] = 0;
    p[GP_time] = e.time; p[GP_wrist_z] = e.wrist_z;
    for (int k = 0; k < 5; k++) p[GP_si_gauge0 + k] = e.si_latest[k];
    p[GP_unstable] = e.unstable; p[GP_weld_on] = e.weld_on; p[GP_base_z] = e.base_z;
    p[GP_tend_x] = e.tend.x; p[GP_tend_y] = e.tend.y; p[GP_tend_z] = e.tend.z; p[GP_tend_th] = e.tend.th;
    p[GP_tnext_x] = e.tnext.x; p[GP_tnext_y] = e.tnext.y; p[GP_tnext_z] = e.tnext.z; p[GP_tnext_th] = e.tnext.th;
    p[GP_cumulative_reward] = e.cumulative_reward; p[GP_has_object] = e.has_object; p[GP_obj_index] = e.obj_index;
    p[GP_episode] = e.episode; p[GP_last_reward] = e.last_reward; p[GP_last_done] = e.last_done;
    p[GP_nrow] = e.nrow_prev; p[GP_last_nrow] = e.nrow_prev;
    p[GP_base_x] = e.base_x; p[GP_base_y] = e.base_y; p[GP_base_yaw] = e.base_yaw;
    for (int k = 0; k < 3; k++) p[GP_read_motor_x + k] = win_read(e, ST_motor + k, 0);
    p[GP_read_base_x] = win_read(e, ST_basexy, 0); p[GP_read_base_y] = win_read(e, ST_basexy + 1, 0);
    p[GP_read_base_z] = win_read(e, ST_basez, 0);
    p[GP_niter] = e.st_substeps ? (double)e.st_iters / (double)e.st_substeps : 0.0;   // mean LCP pivoting rounds per substep so far
    p[GP_ncon] = e.st_substeps ? (double)e.st_contacts / (double)e.st_substeps : 0.0;
  }
}

template <typename T, int N, int QC>
__global__ void k_gauge(const EnvState<T, QC>* envs, const DevModelS<T>* Mp, int n_envs, double* probe) {
  int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= n_envs) return;
  for (int f = 0; f < 3; f++) probe[(size_t)env * GMJ_PROBE_LEN + GP_gauge0 + f] = (double)read_gauge<T, N>(envs[env], *Mp, f);
}

template <typename T, int QC>
__global__ void k_set_welds(EnvState<T, QC>* envs, int n_envs, int B, int N, int mask_on) {
  int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= n_envs) return;
  EnvState<T, QC>& e = envs[env];
  for (int id = 0; id < 7; id++) {
    if ((mask_on >> id) & 1) {
      int j = (id < 3) ? (1 + id * (N + 2)) : ((id < 6) ? (2 + (id - 3) * (N + 2)) : (B - 1));
      e.weld_anchor[id] = e.qpos[j];
    }
  }
  e.weld_on = mask_on & 0x7f;
}

template <typename T, int QC>
__global__ void k_stats(const EnvState<T, QC>* envs, int n_envs, unsigned long long* out) {
  int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= n_envs) return;
  const EnvState<T, QC>& e = envs[env];
  atomicAdd(&out[0], e.st_substeps); atomicAdd(&out[1], (unsigned long long)(e.st_unstable + (e.unstable ? 1 : 0)));
  atomicAdd(&out[2], e.st_contacts); atomicAdd(&out[3], e.st_rows); atomicAdd(&out[4], e.st_iters);
  atomicAdd(&out[5], (unsigned long long)e.st_overflow); atomicAdd(&out[6], (unsigned long long)e.st_relaxed); atomicAdd(&out[7], (unsigned long long)e.st_capped);
}

// MjClass::get_observation of the masked envs straight from the resident state (no stepping, no side effects)
template <typename T, int QC>
__global__ void k_observe_masked(const EnvState<T, QC>* envs, const DevSettings* Sp, int n_envs, const uint8_t* mask, float* obs) {
  int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= n_envs || !mask[env]) return;
  get_observation<T>(envs[env], *Sp, obs + (size_t)env * Sp->n_obs);
}

// EventTrack of every env (MjType::EventTrack, src/mjclass.h:575-647): row / abs counters and last values
template <typename T, int QC>
__global__ void k_event_track(const EnvState<T, QC>* envs, int n_envs, int32_t* row_abs, float* last, float* cum_reward) {
  int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= n_envs) return;
  const EnvState<T, QC>& e = envs[env];
  constexpr int NE = GBE_COUNT + GLE_COUNT;
  for (int k = 0; k < NE; k++) {
    row_abs[((size_t)env * NE + k) * 2] = e.ev_row[k]; row_abs[((size_t)env * NE + k) * 2 + 1] = e.ev_abs[k];
    last[(size_t)env * NE + k] = e.ev_last[k];
  }
  cum_reward[env] = e.cumulative_reward;
}

#ifdef GMJ_PROF
template <typename T, int QC>
__global__ void k_prof(EnvState<T, QC>* envs, int n_envs, unsigned int* out, int clear) {
  int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= n_envs) return;
  for (int k = 0; k < 12; k++) { out[(size_t)env * 12 + k] = envs[env].prof[k]; if (clear) envs[env].prof[k] = 0; }
}
#endif

// ------------------------------------------------------------------------------------------ host side

static void quat2mat_h(const double* q, double* R) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  R[0] = 1 - 2 * (y * y + z * z); R[1] = 2 * (x * y - w * z); R[2] = 2 * (x * z + w * y);
  R[3] = 2 * (x * y + w * z); R[4] = 1 - 2 * (x * x + z * z); R[5] = 2 * (y * z - w * x);
  R[6] = 2 * (x * z - w * y); R[7] = 2 * (y * z + w * x); R[8] = 1 - 2 * (x * x + y * y);
}

// Model constants in slot order. LN = 32: slot == body index (lane-per-dof kernels). LN = 96 (wide kernels):
// finger f, chain position l -> slot 32 f + l; palm -> slot 30; base -> slot 31 of EVERY warp (replicas).
template <typename T, int LN>
static int build_model(const GmjModel* m, DevModel<T, LN>* d) {
  memset((void*)d, 0, sizeof(*d));
  const int N = m->N, NL = N + 2, B = m->B;
  const bool wide = LN != 32;
  if (m->nb != 1) FAIL("only nb = 1 (world_to_base) is supported");
  if (B != 3 * N + 8) FAIL("model does not have B = 3N + 8 gripper bodies");
  if (!wide && B > 32) FAIL("lane-per-dof layout needs B = 3N + 8 <= 32 (N <= 8)");
  if (wide && (NL > 30 || B > GMJ_MAXB)) FAIL("wide layout needs N + 2 <= 30 chain lanes per finger (N <= 28) and B <= 88");
  auto slots = [&](int body, int* out) -> int {
    if (!wide) { out[0] = body; return 1; }
    if (body == 0) { out[0] = 31; out[1] = 63; out[2] = 95; return 3; }
    if (body == B - 1) { out[0] = 30; return 1; }
    int f = (body - 1) / NL; out[0] = 32 * f + (body - 1) - f * NL; return 1;
  };
  d->N = N; d->B = B; d->NL = NL; d->palm_lane = wide ? 30 : B - 1;
  d->zref = (T)0.2;
  for (int i = 0; i < LN; i++) { d->f[MF_gtype][i] = -1; d->f[MF_htype][i] = -1; d->f[MF_mass][i] = 0; d->f[MF_R0][i] = d->f[MF_R4][i] = d->f[MF_R8][i] = 1; }
  for (int i = 0; i < B; i++) {
    double R[9]; quat2mat_h(m->quat[i], R);
    // structural checks for the lane layout
    int exp_parent = (i == 0) ? -1 : ((i == B - 1) ? 0 : (((i - 1) % NL == 0) ? 0 : i - 1));
    if (m->parent[i] != exp_parent) FAIL("model tree does not match the lane layout");
    int exp_hinge = (i != 0 && i != B - 1 && ((i - 1) % NL) > 0);
    if (m->jtype[i] != exp_hinge) FAIL("model joint types do not match the lane layout");
    // the lane-per-dof kernel walks the segments after the first as a planar chain (gmj_physics.cuh, kinematics): their
    // joints must be hinges about the local y axis, offset along local x, with no fixed rotation between segments
    if (!wide && i != 0 && i != B - 1 && ((i - 1) % NL) >= 2) {
      const bool first = ((i - 1) % NL) == 2;
      const bool ax_ok = fabs(m->axis[i][0]) < 1e-12 && fabs(m->axis[i][1] - 1.0) < 1e-12 && fabs(m->axis[i][2]) < 1e-12;
      bool rest_ok = true;
      if (!first) {
        rest_ok = fabs(m->pos[i][1]) < 1e-12 && fabs(m->pos[i][2]) < 1e-12;
        for (int k = 0; k < 9; k++) rest_ok = rest_ok && fabs(R[k] - ((k % 4 == 0) ? 1.0 : 0.0)) < 1e-12;
      }
      if (!ax_ok || !rest_ok) FAIL("finger segments are not a planar chain (hinges about local y, offsets along local x, no fixed rotations)");
    }
    d->qpos0[i] = (T)m->qpos0[i];
    int sl[3]; const int ns = slots(i, sl);
    for (int r = 0; r < ns; r++) {
      const int s = sl[r];
      d->f[MF_parent][s] = (T)m->parent[i]; d->f[MF_jtype][s] = (T)m->jtype[i]; d->f[MF_group][s] = (T)m->group[i];
      for (int k = 0; k < 3; k++) { d->f[MF_pos0 + k][s] = (T)m->pos[i][k]; d->f[MF_ax0 + k][s] = (T)m->axis[i][k]; d->f[MF_ip0 + k][s] = (T)m->ipos[i][k]; }
      if (m->parent[i] < 0) d->f[MF_pos2][s] = (T)(m->pos[i][2] - (double)d->zref);
      for (int k = 0; k < 9; k++) d->f[MF_R0 + k][s] = (T)R[k];
      d->f[MF_mass][s] = (T)m->mass[i];
      for (int k = 0; k < 6; k++) d->f[MF_ixx + k][s] = (T)m->inertia[i][k];
      d->f[MF_stiff][s] = (T)m->stiffness[i]; d->f[MF_damp][s] = (T)m->damping[i];
      d->f[MF_lo][s] = (T)m->range_lo[i]; d->f[MF_hi][s] = (T)m->range_hi[i]; d->f[MF_qpos0][s] = (T)m->qpos0[i];
    }
  }
  // PD control (src/myfunctions.cpp:1965-2057): target selector 1 base z, 2 x, 3 th, 4 z
  auto set_pd = [&](int body, int tsel, double kp, double kd) {
    int sl[3]; const int ns = slots(body, sl);
    for (int r = 0; r < ns; r++) { d->f[MF_tsel][sl[r]] = (T)tsel; d->f[MF_kp][sl[r]] = (T)kp; d->f[MF_kd][sl[r]] = (T)kd; }
  };
  set_pd(0, 1, m->base_kp, m->base_kd);
  for (int f = 0; f < 3; f++) { set_pd(1 + f * NL, 2, m->kp[0], m->kd[0]); set_pd(2 + f * NL, 3, m->kp[1], m->kd[1]); }
  set_pd(B - 1, 4, m->kp[2], m->kd[2]);
  // geoms: first geom of a body is its primary, a second one (the hook plate) goes to the h* slot
  for (int g = 0; g < m->ngeom; g++) {
    if (m->g_body[g] <= 0 || m->g_body[g] >= B) FAIL("collision geom on the base or outside the gripper");
    int sl[3]; slots(m->g_body[g], sl);
    const int b = sl[0];
    double R[9]; quat2mat_h(m->g_quat[g], R);
    int base = (d->f[MF_gtype][b] < 0) ? MF_gtype : ((d->f[MF_htype][b] < 0) ? MF_htype : -1);
    if (base < 0) FAIL("more than two collision geoms on one body");
    if (base == MF_gtype) {
      d->f[MF_gtype][b] = (T)m->g_type[g]; d->f[MF_gfric][b] = (T)m->g_friction[g]; d->f[MF_gcls][b] = (T)m->g_class[g];
      for (int k = 0; k < 3; k++) { d->f[MF_gp0 + k][b] = (T)m->g_pos[g][k]; d->f[MF_gs0 + k][b] = (T)m->g_size[g][k]; }
      for (int k = 0; k < 9; k++) d->f[MF_gR0 + k][b] = (T)R[k];
    } else {
      d->f[MF_htype][b] = (T)m->g_type[g];
      for (int k = 0; k < 3; k++) { d->f[MF_hp0 + k][b] = (T)m->g_pos[g][k]; d->f[MF_hs0 + k][b] = (T)m->g_size[g][k]; }
      for (int k = 0; k < 9; k++) d->f[MF_hR0 + k][b] = (T)R[k];
    }
  }
  d->gravity = (T)m->gravity; d->solref_tc = (T)m->solref[0]; d->solref_dr = (T)m->solref[1];
  for (int k = 0; k < 5; k++) d->solimp[k] = (T)m->solimp[k];
  d->ground_friction = (T)m->ground_friction; d->finger_length = (T)m->finger_length; d->segment_length = (T)m->segment_length;
  d->EI = (T)m->EI; d->clearance = (T)m->fingertip_clearance;
  d->hook_length = (T)m->hook_length; d->hook_angle_deg = (T)m->hook_angle_deg; d->finger_width = (T)m->finger_width;
  for (int f = 0; f < 3; f++) {
    int sl[3]; slots(m->tip_body[f], sl); d->tip_lane[f] = sl[0];
    slots(m->finger_body[f], sl); d->finger_lane[f] = sl[0];
    for (int k = 0; k < 9; k++) d->tip_mat[f][k] = (T)m->tip_mat[f][k];
  }
  return 0;
}

// MjClass::configure_settings action table + Settings::update_sensor_settings
// (src/mjclass.cpp:108-141, 5236-5265, quirk Q11)
static int build_settings(const GmjSettings* s, DevSettings* d) {
  memset(d, 0, sizeof(*d));
  d->s = *s;
  int n = 0, cont = (int)s->general[GS_continous_actions];
  for (int a = 0; a < GACT_COUNT; a++) {
    if (!s->action[a][GAF_in_use]) continue;
    if (cont) d->action_options[n++] = 3 * a + 2;
    else { d->action_options[n++] = 3 * a; d->action_options[n++] = 3 * a + 1; }
  }
  if (s->general[GS_use_termination_action]) d->action_options[n++] = 3 * GACT_COUNT;
  d->n_actions = n;
  double tps = s->general[GS_mujoco_timestep] * (int)s->general[GS_sim_steps_per_action];
  int deep = 0;
  for (int k = 0; k < GSEN_COUNT; k++) {
    int is_state = (k == GSEN_motor_state || k == GSEN_base_XY || k == GSEN_base_Z || k == GSEN_base_yaw || k == GSEN_cartesian);
    if (is_state) { d->prev_steps[k] = (int)s->general[GS_state_n_prev_steps]; d->readings_per_step[k] = 1; }
    else {
      d->prev_steps[k] = (int)s->general[GS_sensor_n_prev_steps];
      float t = (float)tps, rr = (float)s->sensor[k][GSF_read_rate];
      double rs = t * rr;
      d->readings_per_step[k] = (int)floor(rs);
    }
    if (1 + d->prev_steps[k] * d->readings_per_step[k] > GMJ_WIN) deep = 1;
  }
  // observation length (src/mjclass.cpp:1700-1959)
  int smode = (int)s->general[GS_sensor_sample_mode], tmode = (int)s->general[GS_state_sample_mode];
  if (tmode == GSM_scaled_change_sq) { smode = GSM_scaled_change_sq; tmode = GSM_sign; }
  auto len = [&](int sen, int mode) { return mode == GSM_raw ? d->prev_steps[sen] * d->readings_per_step[sen] : 2 * d->prev_steps[sen] + 1; };
  int no = 0;
  if (s->sensor[GSEN_bending_gauge][GSF_in_use]) no += 3 * len(GSEN_bending_gauge, smode);
  if (s->sensor[GSEN_axial_gauge][GSF_in_use]) no += 3 * len(GSEN_axial_gauge, smode);
  if (s->sensor[GSEN_palm][GSF_in_use]) no += len(GSEN_palm, smode);
  if (s->sensor[GSEN_wrist_XY][GSF_in_use]) no += 2 * len(GSEN_wrist_XY, smode);
  if (s->sensor[GSEN_wrist_Z][GSF_in_use]) no += len(GSEN_wrist_XY, smode);
  if (s->sensor[GSEN_motor_state][GSF_in_use]) no += 3 * len(GSEN_motor_state, tmode);
  if (s->sensor[GSEN_base_XY][GSF_in_use]) no += 2 * len(GSEN_base_XY, tmode);
  if (s->sensor[GSEN_base_Z][GSF_in_use]) no += len(GSEN_base_Z, tmode);
  if (s->sensor[GSEN_base_yaw][GSF_in_use]) no += len(GSEN_base_yaw, tmode);
  if (s->sensor[GSEN_cartesian][GSF_in_use]) no += 12 * len(GSEN_cartesian, GSM_change);
  // the reference keeps 1000 readings per sensor; the device ring keeps GMJ_WIN. Settings that need
  // more (e.g. the timestep search runs with a stale sim_steps_per_action) can still be stepped, but
  // asking for the observation is an error (launch_step)
  d->obs_ok = !(deep || no > GMJ_MAX_OBS);
  d->n_obs = no;
  for (int k = 0; k < GSEN_COUNT; k++) { const float rr = (float)s->sensor[k][GSF_read_rate]; d->read_between[k] = (double)(1 / rr); }
  return 0;
}

// ------------------------------------------------------------------------------------------ C ABI
// step-kernel instantiations live in their own translation units (gmj_inst.cu)
#define GMJ_DECL_NARROW(T, N) extern template int launch_step_TN<T, N>(GmjHandle*, int, int, const int32_t*, const float*, float*, float*, uint8_t*, float*, cudaStream_t)
#define GMJ_DECL_WIDE(T, N) extern template int launch_step_wide_TN<T, N>(GmjHandle*, int, int, const int32_t*, const float*, float*, float*, uint8_t*, float*, cudaStream_t)
GMJ_DECL_NARROW(float, 5); GMJ_DECL_NARROW(float, 8); GMJ_DECL_NARROW(double, 5); GMJ_DECL_NARROW(double, 8);
GMJ_DECL_WIDE(float, 16); GMJ_DECL_WIDE(float, 24); GMJ_DECL_WIDE(double, 16); GMJ_DECL_WIDE(double, 24);

// run `...` with T = the handle's physics type and QC = its state capacity
#define GMJ_FOR_STATE(h, ...) do { \
    if (!(h)->wide) { if ((h)->precision == 0) { using T = float; constexpr int QC = 40; __VA_ARGS__; } else { using T = double; constexpr int QC = 40; __VA_ARGS__; } } \
    else { if ((h)->precision == 0) { using T = float; constexpr int QC = GMJ_WQC; __VA_ARGS__; } else { using T = double; constexpr int QC = GMJ_WQC; __VA_ARGS__; } } \
  } while (0)

template <typename T>
static int launch_step_T(GmjHandle* h, int flags, int nsub, const int32_t* actions, const float* frac, float* obs,
                         float* reward, uint8_t* done, float* events, cudaStream_t st) {
  switch (h->N) {
    case 5: return launch_step_TN<T, 5>(h, flags, nsub, actions, frac, obs, reward, done, events, st);
    case 8: return launch_step_TN<T, 8>(h, flags, nsub, actions, frac, obs, reward, done, events, st);
    case 16: return launch_step_wide_TN<T, 16>(h, flags, nsub, actions, frac, obs, reward, done, events, st);
    case 24: return launch_step_wide_TN<T, 24>(h, flags, nsub, actions, frac, obs, reward, done, events, st);
    default: FAIL("this build instantiates the step kernel for N = 5, 8 (one env per warp) and N = 16, 24 (one env per block) segments");
  }
}
static int launch_step(GmjHandle* h, int flags, int nsub, const int32_t* actions, const float* frac, float* obs,
                       float* reward, uint8_t* done, float* events, void* stream) {
  if (obs && !h->hs.obs_ok) FAIL("observation unavailable: sensor history deeper than the device ring (GMJ_WIN) or longer than GMJ_MAX_OBS");
  DeviceGuard dg_(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (h->precision == 0) return launch_step_T<float>(h, flags, nsub, actions, frac, obs, reward, done, events, st);
  return launch_step_T<double>(h, flags, nsub, actions, frac, obs, reward, done, events, st);
}
static size_t env_bytes(const GmjHandle* h) { size_t n = 0; GMJ_FOR_STATE(h, n = sizeof(EnvState<T, QC>)); return n; }

template <typename T, int LN>
static int upload_model(GmjHandle* h, const GmjModel* model) {
  DevModel<T, LN>* dm = new DevModel<T, LN>();
  if (build_model<T, LN>(model, dm)) { delete dm; return 1; }
  DevModel<T, LN>* dev = nullptr;
  cudaError_t e = cudaMalloc(&dev, sizeof(*dm));
  if (e == cudaSuccess) e = cudaMemcpy(dev, dm, sizeof(*dm), cudaMemcpyHostToDevice);
  delete dm;
  if (e != cudaSuccess) { gmj_set_error(std::string("model upload: ") + cudaGetErrorString(e)); return 1; }
  h->model = dev; h->model_s = static_cast<const DevModelS<T>*>(dev);
  return 0;
}

extern "C" {

const char* gmj_last_error(void) { return g_err.c_str(); }

int gmj_set_settings(GmjHandle* h, const GmjSettings* settings) {
  if (!h || !settings) FAIL("null argument");
  DeviceGuard dg_(h->device);
  CK(cudaDeviceSynchronize());        // host-synchronising entry point: ordered after everything queued on any stream of the device
  if (build_settings(settings, &h->hs)) return 1;
  { const char* t = getenv("GMJ_ASIN_TAB"); h->hs.asin_tab = (t && atoi(t) == 0) ? nullptr : h->asin_tab; }   // development switch
  CK(cudaMemcpy(h->settings, &h->hs, sizeof(DevSettings), cudaMemcpyHostToDevice));
  return 0;
}

static int create_impl(GmjHandle* h, const GmjModel* model, const GmjSettings* settings, int n_envs, int device, int precision) {
  {
    int nsm = 0;
    CK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, device));
    // development switch GMJ_BALANCE=1: equal env share per block over the SMs instead of full blocks. Measured on
    // B200 (profiles/README.md): no gain on the bench workload, 7 % slower on the lift-free profile workload -> off.
    const char* bal = getenv("GMJ_BALANCE");
    h->balance_sm = (bal && atoi(bal) == 1) ? nsm : 0;
    h->n_sm = nsm;
  }
  int rc;
  if (!h->wide) rc = precision == 0 ? upload_model<float, 32>(h, model) : upload_model<double, 32>(h, model);
  else rc = precision == 0 ? upload_model<float, GMJ_WNT>(h, model) : upload_model<double, GMJ_WNT>(h, model);
  if (rc) return 1;
  CK(cudaMalloc(&h->envs, env_bytes(h) * (size_t)n_envs));
  CK(cudaMemset(h->envs, 0, env_bytes(h) * (size_t)n_envs));
  CK(cudaMalloc(&h->settings, sizeof(DevSettings)));
  {                                    // finger angle of every leadscrew step difference (g_update_xy_steps, gmj_tail.cuh)
    std::vector<double> tab(2 * G_TH_TAB + 1);
    for (int d = -G_TH_TAB; d <= G_TH_TAB; d++) tab[d + G_TH_TAB] = asin((G_XY_STEP * d) / G_LEADSCREW);
    CK(cudaMalloc(&h->asin_tab, tab.size() * sizeof(double)));
    CK(cudaMemcpy(h->asin_tab, tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice));
  }
  CK(cudaMalloc(&h->taps, (size_t)n_envs * GMJ_MAXTAP * (GMJ_TAP_HDR + 3 * model->N) * (precision == 0 ? sizeof(float) : sizeof(double))));
  CK(cudaMalloc(&h->err_flag, sizeof(int))); CK(cudaMemset(h->err_flag, 0, sizeof(int)));
  CK(cudaMalloc(&h->stats_dev, 8 * sizeof(unsigned long long)));
  CK(cudaMalloc(&h->env_list, sizeof(int) * (size_t)n_envs)); CK(cudaMalloc(&h->env_count, sizeof(int)));
  CK(cudaMemset(h->env_count, 0, sizeof(int)));
  CK(cudaMalloc(&h->perm, sizeof(int) * (size_t)n_envs)); CK(cudaMalloc(&h->hist, sizeof(int) * 4 * GMJ_SORT_BUCKETS));
  CK(cudaMalloc(&h->spawn_ok, (size_t)n_envs)); CK(cudaMemset(h->spawn_ok, 0, (size_t)n_envs));
  if (gmj_set_settings(h, settings)) return 1;
  return gmj_reset(h, nullptr, nullptr);
}

int gmj_create(const GmjModel* model, const GmjSettings* settings, int n_envs, int device, int precision,
               uint32_t env_id_offset, GmjHandle** out) {
  if (!model || !settings || !out || n_envs <= 0) FAIL("bad argument to gmj_create");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) FAIL("no CUDA device: libgripper_b200 has no CPU fallback");
  if (device < 0 || device >= ndev) FAIL("gmj_create: no such CUDA device");
  DeviceGuard dg_(device);
  GmjHandle* h = new GmjHandle();
  memset((void*)h, 0, sizeof(*h));
  h->n_envs = n_envs; h->device = device; h->precision = precision; h->N = model->N; h->B = model->B; h->hm = *model;
  h->env_id_offset = env_id_offset;
  h->wide = model->B > 32;
  if (create_impl(h, model, settings, n_envs, device, precision)) {
    const std::string keep = g_err;          // every allocation made so far goes back; the first error is the one reported
    gmj_destroy(h);
    g_err = keep;
    return 1;
  }
  *out = h;
  return 0;
}

int gmj_destroy(GmjHandle* h) {
  if (!h) return 0;
  DeviceGuard dg_(h->device);
  cudaDeviceSynchronize();
  cudaFree(h->spawn_ok);
  cudaFree(h->envs); cudaFree(h->model); cudaFree(h->settings); cudaFree(h->err_flag); cudaFree(h->stats_dev); cudaFree(h->env_list); cudaFree(h->env_count); cudaFree(h->perm); cudaFree(h->hist); cudaFree(h->asin_tab); cudaFree(h->taps);
  if (h->objects) cudaFree(h->objects);
  if (h->scratch) cudaFree(h->scratch);
  delete h;
  return 0;
}

int gmj_set_objects(GmjHandle* h, const GmjObject* objects_host, int n_objects) {
  if (!h || !objects_host || n_objects <= 0) FAIL("bad argument to gmj_set_objects");
  DeviceGuard dg_(h->device);
  if (h->objects) CK(cudaFree(h->objects));
  CK(cudaMalloc(&h->objects, sizeof(GmjObject) * (size_t)n_objects));
  CK(cudaMemcpy(h->objects, objects_host, sizeof(GmjObject) * (size_t)n_objects, cudaMemcpyHostToDevice));
  h->n_objects = n_objects;
  return 0;
}

int gmj_reset(GmjHandle* h, const uint8_t* mask_dev, void* stream) {
  if (!h) FAIL("null handle");
  DeviceGuard dg_(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  int tb = 128, nb = (h->n_envs + tb - 1) / tb;
  GMJ_FOR_STATE(h, k_reset<T, QC><<<nb, tb, 0, st>>>((EnvState<T, QC>*)h->envs, (const DevModelS<T>*)h->model_s, h->settings, h->n_envs, mask_dev, h->env_id_offset, 0, nullptr));
  CK(cudaGetLastError());
  return 0;
}

int gmj_reset_done(GmjHandle* h, const uint8_t* done_dev, int step_cap, uint8_t* mask_out_dev, void* stream) {
  if (!h || !done_dev) FAIL("gmj_reset_done: null handle or done flags");
  if (step_cap <= 0) FAIL("gmj_reset_done: step_cap must be positive");
  DeviceGuard dg_(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  int tb = 128, nb = (h->n_envs + tb - 1) / tb;
  GMJ_FOR_STATE(h, k_reset<T, QC><<<nb, tb, 0, st>>>((EnvState<T, QC>*)h->envs, (const DevModelS<T>*)h->model_s, h->settings, h->n_envs, done_dev, h->env_id_offset, step_cap, mask_out_dev));
  CK(cudaGetLastError());
  return 0;
}

int gmj_spawn(GmjHandle* h, const int32_t* idx_dev, const float* xyr_dev, const uint8_t* mask_dev, double xy_range,
              double rot_range, int tries, int fallback_xy_mm, int fallback_rot_deg, uint8_t* ok_dev, void* stream) {
  if (!h) FAIL("null handle");
  if (!h->objects) FAIL("gmj_spawn before gmj_set_objects");
  if (!idx_dev && xyr_dev) FAIL("gmj_spawn: xyr_dev given without idx_dev");
  if (xy_range < 0 || rot_range < 0) FAIL("gmj_spawn: negative range");
  if (tries < 1) tries = 1;
  if (fallback_rot_deg < 0) fallback_rot_deg = 0;
  DeviceGuard dg_(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  int tb = 32 * GMJ_SPAWN_WPB, nb = (h->n_envs + GMJ_SPAWN_WPB - 1) / GMJ_SPAWN_WPB;
  uint8_t* ok = ok_dev ? ok_dev : h->spawn_ok;
  GMJ_FOR_STATE(h, k_spawn<T, QC><<<nb, tb, 0, st>>>((EnvState<T, QC>*)h->envs, (const DevModelS<T>*)h->model_s, h->settings, h->objects, h->n_objects, h->n_envs, idx_dev, xyr_dev, mask_dev, xy_range, rot_range, tries, fallback_xy_mm, fallback_rot_deg, ok, h->err_flag));
  CK(cudaGetLastError());
  return 0;
}

// Pass 1: set_action for every env; envs that owe the termination action's 2 x sim_steps_per_action
// lift substeps (quirk Q8: they run INSIDE MjClass::set_action) are recorded in env_list / pending_extra, the
// others are counting-sorted by expected load.
static int launch_collect(GmjHandle* h, const int32_t* actions_dev, const float* frac_dev, void* stream) {
  DeviceGuard dg_(h->device);
  CK(cudaMemsetAsync(h->env_count, 0, sizeof(int), (cudaStream_t)stream));
  CK(cudaMemsetAsync(h->hist, 0, sizeof(int) * 4 * GMJ_SORT_BUCKETS, (cudaStream_t)stream));
  if (launch_step(h, STEP_SET_ACTION | STEP_COLLECT, 0, actions_dev, frac_dev, nullptr, nullptr, nullptr, nullptr, stream)) return 1;
  int tb = 256, nb = (h->n_envs + tb - 1) / tb;
  const int sort_lift = h->n_sm > 0 && h->n_envs > 3 * h->n_sm * 12;
  GMJ_FOR_STATE(h, k_sort_scatter<T, QC><<<nb, tb, 0, (cudaStream_t)stream>>>((const EnvState<T, QC>*)h->envs, h->n_envs, h->hist, h->perm, h->env_list, sort_lift));
  CK(cudaGetLastError());
  return 0;
}
// Fused step: pass 1 + ONE combined launch: the first blocks take the (few) lifting envs, densely packed, and run
// lift + action substeps back to back; the other blocks take everybody else (in a lockstep block a single lifting
// env would otherwise hold the whole block for 3S substeps). Same scheme for the lane-per-dof and the wide kernels.
int gmj_step(GmjHandle* h, const int32_t* actions_dev, const float* frac_dev, float* obs_dev, float* reward_dev,
             uint8_t* done_dev, float* events_dev, void* stream) {
  if (!h || !actions_dev) FAIL("gmj_step: null handle or actions");
  if (launch_collect(h, actions_dev, frac_dev, stream)) return 1;
  return launch_step(h, STEP_SUBSTEPS | STEP_SENSORS | STEP_TAIL | STEP_OUTPUT | STEP_COMBINED, -1, nullptr, nullptr,
                     obs_dev, reward_dev, done_dev, events_dev, stream);
}
int gmj_set_action(GmjHandle* h, const int32_t* actions_dev, const float* frac_dev, void* stream) {
  if (!h || !actions_dev) FAIL("gmj_set_action: null handle or actions");
  if (launch_collect(h, actions_dev, frac_dev, stream)) return 1;
  return launch_step(h, STEP_SUBSTEPS | STEP_SENSORS | STEP_LIST, 0, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, stream);
}
int gmj_action_step(GmjHandle* h, void* stream) {
  if (!h) FAIL("null handle");
  return launch_step(h, STEP_SUBSTEPS | STEP_SENSORS | STEP_TAIL, -1, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, stream);
}
int gmj_substeps(GmjHandle* h, int n, int sensors, void* stream) {
  if (!h || n < 0) FAIL("bad argument to gmj_substeps");
  return launch_step(h, STEP_SUBSTEPS | (sensors ? STEP_SENSORS : 0), n, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, stream);
}
int gmj_observe(GmjHandle* h, float* obs_dev, float* reward_dev, uint8_t* done_dev, float* events_dev, void* stream) {
  if (!h) FAIL("null handle");
  return launch_step(h, STEP_OUTPUT, 0, nullptr, nullptr, obs_dev, reward_dev, done_dev, events_dev, stream);
}

int gmj_observe_masked(GmjHandle* h, const uint8_t* mask_dev, float* obs_dev, void* stream) {
  if (!h || !mask_dev || !obs_dev) FAIL("bad argument to gmj_observe_masked");
  if (!h->hs.obs_ok) FAIL("observation unavailable: sensor history deeper than the device ring (GMJ_WIN) or longer than GMJ_MAX_OBS");
  DeviceGuard dg_(h->device);
  int tb = 64, nb = (h->n_envs + tb - 1) / tb;
  GMJ_FOR_STATE(h, k_observe_masked<T, QC><<<nb, tb, 0, (cudaStream_t)stream>>>((const EnvState<T, QC>*)h->envs, h->settings, h->n_envs, mask_dev, obs_dev));
  CK(cudaGetLastError());
  return 0;
}

int gmj_n_obs(GmjHandle* h) { return h ? h->hs.n_obs : -1; }
int gmj_n_actions(GmjHandle* h) { return h ? h->hs.n_actions : -1; }
int gmj_n_envs(GmjHandle* h) { return h ? h->n_envs : -1; }

static int ensure_scratch(GmjHandle* h, size_t n) {
  if (h->scratch_len >= n) return 0;
  if (h->scratch) CK(cudaFree(h->scratch));
  CK(cudaMalloc(&h->scratch, n * sizeof(double)));
  h->scratch_len = n;
  return 0;
}

int gmj_get_state_host(GmjHandle* h, double* qpos_host, double* qvel_host) {
  if (!h) FAIL("null handle");
  DeviceGuard dg_(h->device);
  CK(cudaDeviceSynchronize());        // host-synchronising entry point: ordered after everything queued on any stream of the device
  size_t nq = (size_t)h->n_envs * (h->B + 7), nv = (size_t)h->n_envs * (h->B + 6);
  if (ensure_scratch(h, nq + nv)) return 1;
  int tb = 128, nb = (h->n_envs + tb - 1) / tb;
  GMJ_FOR_STATE(h, k_dump<T, QC><<<nb, tb>>>((const EnvState<T, QC>*)h->envs, h->n_envs, h->B, h->scratch, h->scratch + nq, nullptr));
  CK(cudaGetLastError());
  if (qpos_host) CK(cudaMemcpy(qpos_host, h->scratch, nq * sizeof(double), cudaMemcpyDeviceToHost));
  if (qvel_host) CK(cudaMemcpy(qvel_host, h->scratch + nq, nv * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}

int gmj_probe_host(GmjHandle* h, double* probe_host) {
  if (!h || !probe_host) FAIL("null argument");
  DeviceGuard dg_(h->device);
  CK(cudaDeviceSynchronize());        // host-synchronising entry point: ordered after everything queued on any stream of the device
  size_t n = (size_t)h->n_envs * GMJ_PROBE_LEN;
  if (ensure_scratch(h, n)) return 1;
  int tb = 128, nb = (h->n_envs + tb - 1) / tb;
  GMJ_FOR_STATE(h, k_dump<T, QC><<<nb, tb>>>((const EnvState<T, QC>*)h->envs, h->n_envs, h->B, nullptr, nullptr, h->scratch));
#define GMJ_GAUGE(NN) GMJ_FOR_STATE(h, k_gauge<T, NN, QC><<<nb, tb>>>((const EnvState<T, QC>*)h->envs, (const DevModelS<T>*)h->model_s, h->n_envs, h->scratch))
  switch (h->N) { case 5: GMJ_GAUGE(5); break; case 8: GMJ_GAUGE(8); break; case 16: GMJ_GAUGE(16); break; case 24: GMJ_GAUGE(24); break; default: break; }
#undef GMJ_GAUGE
  CK(cudaGetLastError());
  CK(cudaMemcpy(probe_host, h->scratch, n * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}

// ------------------------------------------------------------------ allocation-free probe
// Scratch of gmj_probe_host_noalloc: lives in the library image (every GPU that loads the library holds these 5 MiB
// until exit), shared by every call and every handle on a device -- calls that overlap on the GPU race on it.
static __device__ __align__(256) double g_noalloc_scratch[(size_t)GMJ_NOALLOC_MAX_ENVS * GMJ_PROBE_LEN];
// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is looked up
// once per device, with that device current, and cached here.
#define GMJ_NOALLOC_MAX_DEVICES 64
static std::atomic<double*> g_noalloc_ptr[GMJ_NOALLOC_MAX_DEVICES];

int gmj_probe_host_noalloc(GmjHandle* h, double* probe_host) {
  if (!h || !probe_host) FAIL("null argument");
  if (h->n_envs > GMJ_NOALLOC_MAX_ENVS)
    FAIL("gmj_probe_host_noalloc: n_envs = " + std::to_string(h->n_envs) + " exceeds the static scratch (GMJ_NOALLOC_MAX_ENVS = " +
         std::to_string(GMJ_NOALLOC_MAX_ENVS) + "); use gmj_probe_host");
  if (h->device < 0 || h->device >= GMJ_NOALLOC_MAX_DEVICES) FAIL("gmj_probe_host_noalloc: device index out of range");
  DeviceGuard dg_(h->device);
  CK(cudaDeviceSynchronize());        // host-synchronising entry point (no allocation: a sync allocates nothing)
  double* scratch = g_noalloc_ptr[h->device].load(std::memory_order_acquire);
  if (!scratch) {
    double* ptr = nullptr;
    CK(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
    g_noalloc_ptr[h->device].store(ptr, std::memory_order_release);     // same address whichever thread wins
    scratch = ptr;
  }
  const size_t n = (size_t)h->n_envs * GMJ_PROBE_LEN;
  int tb = 128, nb = (h->n_envs + tb - 1) / tb;
  // k_dump zeroes every probe row before filling it, so nothing of the previous call survives in the rows read back
  GMJ_FOR_STATE(h, k_dump<T, QC><<<nb, tb>>>((const EnvState<T, QC>*)h->envs, h->n_envs, h->B, nullptr, nullptr, scratch));
#define GMJ_GAUGE(NN) GMJ_FOR_STATE(h, k_gauge<T, NN, QC><<<nb, tb>>>((const EnvState<T, QC>*)h->envs, (const DevModelS<T>*)h->model_s, h->n_envs, scratch))
  switch (h->N) { case 5: GMJ_GAUGE(5); break; case 8: GMJ_GAUGE(8); break; case 16: GMJ_GAUGE(16); break; case 24: GMJ_GAUGE(24); break; default: break; }
#undef GMJ_GAUGE
  CK(cudaGetLastError());
  CK(cudaMemcpy(probe_host, scratch, n * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}

#ifdef GMJ_PROF
// development build only: per-env cycle counters per phase, uint32[n_envs][12] to the host; clear != 0 zeroes them
int gmj_prof_host(GmjHandle* h, unsigned int* out_host, int clear) {
  if (!h || !out_host) FAIL("null argument");
  DeviceGuard dg_(h->device);
  CK(cudaDeviceSynchronize());
  unsigned int* d = nullptr;
  CK(cudaMalloc(&d, sizeof(unsigned int) * 12 * (size_t)h->n_envs));
  int tb = 128, nb = (h->n_envs + tb - 1) / tb;
  GMJ_FOR_STATE(h, k_prof<T, QC><<<nb, tb>>>((EnvState<T, QC>*)h->envs, h->n_envs, d, clear));
  CK(cudaGetLastError());
  CK(cudaMemcpy(out_host, d, sizeof(unsigned int) * 12 * (size_t)h->n_envs, cudaMemcpyDeviceToHost));
  CK(cudaFree(d));
  return 0;
}
#endif

int gmj_event_track_host(GmjHandle* h, int32_t* row_abs_host, float* last_host, float* cum_reward_host) {
  if (!h || !row_abs_host || !last_host || !cum_reward_host) FAIL("null argument");
  DeviceGuard dg_(h->device);
  CK(cudaDeviceSynchronize());
  constexpr int NE = GBE_COUNT + GLE_COUNT;
  const size_t n = (size_t)h->n_envs;
  // scratch counted in doubles: int32[n][NE][2] | float[n][NE] | float[n]
  const size_t bytes = n * NE * 2 * sizeof(int32_t) + n * NE * sizeof(float) + n * sizeof(float);
  if (ensure_scratch(h, (bytes + sizeof(double) - 1) / sizeof(double))) return 1;
  int32_t* ra = (int32_t*)h->scratch; float* la = (float*)(ra + n * NE * 2); float* cr = la + n * NE;
  int tb = 128, nb = (h->n_envs + tb - 1) / tb;
  GMJ_FOR_STATE(h, k_event_track<T, QC><<<nb, tb>>>((const EnvState<T, QC>*)h->envs, h->n_envs, ra, la, cr));
  CK(cudaGetLastError());
  CK(cudaMemcpy(row_abs_host, ra, n * NE * 2 * sizeof(int32_t), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(last_host, la, n * NE * sizeof(float), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(cum_reward_host, cr, n * sizeof(float), cudaMemcpyDeviceToHost));
  return 0;
}

int gmj_stats(GmjHandle* h, GmjStats* out) {
  if (!h || !out) FAIL("null argument");
  DeviceGuard dg_(h->device);
  CK(cudaDeviceSynchronize());        // host-synchronising entry point: ordered after everything queued on any stream of the device
  CK(cudaMemset(h->stats_dev, 0, 8 * sizeof(unsigned long long)));
  int tb = 128, nb = (h->n_envs + tb - 1) / tb;
  GMJ_FOR_STATE(h, k_stats<T, QC><<<nb, tb>>>((const EnvState<T, QC>*)h->envs, h->n_envs, h->stats_dev));
  CK(cudaGetLastError());
  unsigned long long v[8];
  CK(cudaMemcpy(v, h->stats_dev, sizeof(v), cudaMemcpyDeviceToHost));
  out->substeps = (int64_t)v[0]; out->unstable_resets = (int64_t)v[1]; out->contacts = (int64_t)v[2];
  out->rows = (int64_t)v[3]; out->solver_iters = (int64_t)v[4]; out->contact_overflow = (int64_t)v[5]; out->relaxed_solves = (int64_t)v[6]; out->capped_solves = (int64_t)v[7];
  int ef = 0;
  CK(cudaMemcpy(&ef, h->err_flag, sizeof(int), cudaMemcpyDeviceToHost));
  if (ef) { CK(cudaMemset(h->err_flag, 0, sizeof(int))); g_err = ef == 1 ? "an action index was out of range" : (ef == 2 ? "a spawn object index was out of range" : "spawn ranges give more candidates than the device search holds (1024 positions, 128 rotations)"); return 2; }
  return 0;
}

// ---- MjClass::configure_settings' auto-calibration on the host side of the C ABI (same control flow and arithmetic as
// gripper_mujoco_b200/calibration.py, which drives the oracle as well)
namespace {
struct CalibRun { std::vector<double> qpos; bool unstable; double si_wrist, gauge0; };
// one freshly reset scratch env with `st`, stepped n times through gmj_substeps (MjClass._backend in mjclass.py)
int calib_run(const GmjModel* model, const GmjSettings* st, int device, int precision, int n_substeps, int sensors, CalibRun* out, int* launches) {
  GmjHandle* h = nullptr;
  if (gmj_create(model, st, 1, device, precision, 0xC0FFEEu, &h)) return 1;
  int rc = gmj_substeps(h, n_substeps, sensors, nullptr);
  out->qpos.assign((size_t)model->B + 7, 0.0);
  std::vector<double> probe(GMJ_PROBE_LEN, 0.0);
  GmjStats sv; memset(&sv, 0, sizeof(sv));
  if (!rc) rc = gmj_get_state_host(h, out->qpos.data(), nullptr);
  if (!rc) rc = gmj_probe_host(h, probe.data());
  if (!rc) rc = gmj_stats(h, &sv) == 1;
  out->unstable = probe[GP_unstable] > 0 || sv.unstable_resets > 0;
  out->si_wrist = probe[GP_si_wrist]; out->gauge0 = probe[GP_gauge0];
  const std::string keep = g_err;
  gmj_destroy(h);
  if (rc) g_err = keep;
  (*launches)++;
  return rc;
}
float calib_yield_load(const GmjModel* m) {          // luke::calc_yield_point_load (src/myfunctions.cpp:3587-3606), float arithmetic
  const double t = m->finger_thickness, w = m->finger_width;
  const float inertia = (float)((w * pow(t, 3.0)) / 12.0);
  const float m_max = (215e6f * inertia) / (float)(0.5 * t);
  return m_max / (float)m->finger_length;
}
}  // namespace

int gmj_calibrate(const GmjModel* model, GmjSettings* s, int device, int precision, GmjCalib* io) {
  if (!model || !s || !io) FAIL("null argument to gmj_calibrate");
  const int N = model->N, NL = N + 2, B = model->B;
  io->launches = 0; io->highest_stable_timestep = 0;
  CalibRun r;
  // ---- luke::calibrate_reset (src/myfunctions.cpp:1450-1516): 0.8 s settle from the keyframe with every weld on
  {
    GmjSettings st = *s;
    st.general[GS_mujoco_timestep] = 2e-3; st.general[GS_base_position_noise] = 0.0; st.general[GS_curve_validation] = 0.0;
    st.general[GS_reset_qpos_base] = 0.0;
    for (int f = 0; f < 3; f++) { st.general[GS_reset_qpos_g0 + 2 * f] = model->qpos0[1 + f * NL]; st.general[GS_reset_qpos_g1 + 2 * f] = 0.0; }
    st.general[GS_reset_qpos_g6] = model->qpos0[B - 1];
    int steps = 400; bool ok = false;
    for (int attempt = 0; attempt < 5 && !ok; attempt++) {     // halve the timestep until the settle is stable (long fingers)
      if (calib_run(model, &st, device, precision, steps, 0, &r, &io->launches)) return 1;
      double mx = 0; bool fin = true;
      for (int i = 0; i < B; i++) { fin = fin && std::isfinite(r.qpos[i]); mx = fmax(mx, fabs(r.qpos[i])); }
      for (size_t i = 0; i < r.qpos.size(); i++) fin = fin && std::isfinite(r.qpos[i]);
      ok = !r.unstable && fin && mx < 10.0;
      if (!ok) { st.general[GS_mujoco_timestep] *= 0.5; steps *= 2; }
    }
    if (!ok) FAIL("calibrate_reset: no stable settle found");
    s->general[GS_reset_qpos_base] = r.qpos[0];
    for (int f = 0; f < 3; f++) { s->general[GS_reset_qpos_g0 + 2 * f] = r.qpos[1 + f * NL]; s->general[GS_reset_qpos_g1 + 2 * f] = r.qpos[2 + f * NL]; }
    s->general[GS_reset_qpos_g6] = r.qpos[B - 1];
  }
  // ---- MjClass::find_highest_stable_timestep (src/mjclass.cpp:4745-4850): float arithmetic as in the reference
  if (io->set_timestep) {
    const float coarse = 0.5e-3f, fine = 50e-6f, test_time = 1.0f, max_dt = 20e-3f;
    float nxt = 1.0e-3f; bool coarse_pass = true;
    GmjSettings st = *s;
    for (;;) {
      st.general[GS_mujoco_timestep] = (double)nxt;
      if (calib_run(model, &st, device, precision, (int)(test_time / nxt) + 1, 1, &r, &io->launches)) return 1;
      if (r.unstable) { coarse_pass = false; nxt = nxt - fine; }
      else if (coarse_pass) nxt = nxt + coarse;
      else break;
      if (nxt < fine) FAIL("no stable timestep found for the simulation");
      if (nxt > max_dt) { nxt = max_dt; coarse_pass = false; }
    }
    const float factor = nxt <= 3.0e-3f ? 0.8f : (nxt < 5.0e-3f ? 0.75f : 0.65f);
    float fin = nxt * factor;
    fin = (float)((double)(int)(fin * 1e6f) * 1e-6);
    s->general[GS_mujoco_timestep] = (double)fin;
    io->highest_stable_timestep = (double)nxt;
  }
  const float yl = calib_yield_load(model);
  io->yield_load = (double)yl;
  // ---- MjClass::calibrate_simulated_sensors (src/mjclass.cpp:4643-4676)
  if (io->calibrate_gauges) {
    const double norm = (double)(float)(s->general[GS_saturation_yield_factor] * (double)yl);
    GmjSettings st = *s;
    st.sensor[GSEN_wrist_Z][GSF_use_normalisation] = 0; st.sensor[GSEN_wrist_Z][GSF_use_noise] = 0; st.sensor[GSEN_wrist_Z][GSF_raw_value_offset] = 0;
    if (calib_run(model, &st, device, precision, (int)(0.3f / (float)st.general[GS_mujoco_timestep]), 1, &r, &io->launches)) return 1;
    s->sensor[GSEN_wrist_Z][GSF_raw_value_offset] = r.si_wrist;
    st = *s;
    st.general[GS_curve_validation] = 1.0; st.general[GS_tip_force_applied] = norm;
    bool ok = false;
    for (int attempt = 0; attempt < 5 && !ok; attempt++) {     // validate_curve_under_force retries at 0.8 x timestep
      if (calib_run(model, &st, device, precision, (int)(50.0f / (float)st.general[GS_mujoco_timestep]), 1, &r, &io->launches)) return 1;
      ok = !r.unstable;
      if (!ok) { st.general[GS_mujoco_timestep] *= 0.8; s->general[GS_mujoco_timestep] = st.general[GS_mujoco_timestep]; }
    }
    if (!ok) FAIL("curve validation unstable");
    s->sensor[GSEN_bending_gauge][GSF_normalise] = r.gauge0;
    s->general[GS_curve_validation] = 0.0;
    s->general[GS_sim_gauge_raw_to_N_factor] = norm / r.gauge0;
  }
  if (io->set_sim_steps) s->general[GS_sim_steps_per_action] = ceil(s->general[GS_time_for_action] / s->general[GS_mujoco_timestep]);
  io->timestep = s->general[GS_mujoco_timestep];
  io->sim_steps_per_action = (int)s->general[GS_sim_steps_per_action];
  io->wrist_raw_value_offset = s->sensor[GSEN_wrist_Z][GSF_raw_value_offset];
  io->bend_gauge_normalise = s->sensor[GSEN_bending_gauge][GSF_normalise];
  io->sim_gauge_raw_to_N_factor = s->general[GS_sim_gauge_raw_to_N_factor];
  io->reset_qpos[0] = s->general[GS_reset_qpos_base];
  for (int k = 0; k < 7; k++) io->reset_qpos[1 + k] = s->general[GS_reset_qpos_g0 + k];
  return 0;
}

int gmj_set_welds(GmjHandle* h, int mask_on, void* stream) {
  if (!h) FAIL("null handle");
  DeviceGuard dg_(h->device);
  int tb = 128, nb = (h->n_envs + tb - 1) / tb;
  GMJ_FOR_STATE(h, k_set_welds<T, QC><<<nb, tb, 0, (cudaStream_t)stream>>>((EnvState<T, QC>*)h->envs, h->n_envs, h->B, h->N, mask_on));
  CK(cudaGetLastError());
  return 0;
}

}  // extern "C"
