/* env.c -- oracle: MjClass / luke:: control flow around the physics step.
 * TEST INFRASTRUCTURE ONLY (see oracle.h). Each function cites the reference lines it restates.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "oracle.h"
#include "vec.h"

void or_kinematics(OrEnv* e);
void or_update_env(OrEnv* e);

/* ------------------------------------------------------------------ counter-based RNG
 * The reference draws from a process-global std::default_random_engine (src/mjclass.cpp:4,
 * src/mjclass.h:176-223). A batched environment cannot share one serial engine, so noise is
 * drawn from Philox4x32-10 keyed by (seed) and countered by (draw index, env id); the CUDA
 * path implements the same generator so noisy runs are comparable. */
static uint32_t mulhilo(uint32_t a, uint32_t b, uint32_t* hi) {
  uint64_t p = (uint64_t)a * b; *hi = (uint32_t)(p >> 32); return (uint32_t)p;
}
static uint32_t philox_first(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; r++) {
    uint32_t hi0, hi1, lo0 = mulhilo(0xD2511F53u, c0, &hi0), lo1 = mulhilo(0xCD9E8D57u, c2, &hi1);
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return c0;
}
static float or_uniform(OrEnv* e) {
  uint32_t x = philox_first(e->rng_ctr++, e->env_id, 0, 0, (uint32_t)e->seed, (uint32_t)(e->seed >> 32));
  return (float)(x >> 8) * (1.0f / 16777216.0f);
}

/* ------------------------------------------------------------------ welds
 * luke::set_constraint (src/myfunctions.cpp:1177-1279): anchor = pose of body2 relative to
 * body1 at activation: R12 = R1^T R2, v12 = R1 * (p2 - p1) (sic, the reference multiplies by R1,
 * not R1^T; identical for the symmetric base rotation and for coincident origins). */
void or_set_weld(OrEnv* e, int id, int on) {
  OrWeld* w = &e->weld[id];
  if (on) {
    double d[3];
    sub3(d, e->xpos[w->body2], e->xpos[w->body1]);
    mulv(w->relpos, e->xmat[w->body1], d);
    mulTm(w->relmat, e->xmat[w->body1], e->xmat[w->body2]);
  }
  w->on = on;
}

static void set_all_welds(OrEnv* e, int on) { for (int i = 0; i < 7; i++) or_set_weld(e, i, on); }

/* Target::prismatic_moving / revolute_moving / z_moving (src/customtypes.h:657-692) */
static int x_moving(const OrEnv* e) { return e->tgt_end.sx != e->tgt_next.sx; }
static int z_moving(const OrEnv* e) { return e->tgt_end.sz != e->tgt_next.sz; }
static int rev_moving(const OrEnv* e) {
  double a = (180.0 / M_PI) * or_gripper_th(&e->tgt_end), b = (180.0 / M_PI) * or_gripper_th(&e->tgt_next);
  return !(fabs(a - b) < 5e-1);
}

/* luke::update_constraints (src/myfunctions.cpp:2216-2284); revolute welds are never re-enabled
 * (disable_rev_constraint = true, :479) */
static void update_constraints(OrEnv* e) {
  int nx = x_moving(e), ny = rev_moving(e), nz = z_moving(e);
  if (nx != e->old_x) { for (int i = 0; i < 3; i++) or_set_weld(e, i, !nx); e->old_x = nx; }
  if (ny != e->old_y) { if (ny) for (int i = 3; i < 6; i++) or_set_weld(e, i, 0); e->old_y = ny; }
  if (nz != e->old_z) { or_set_weld(e, 6, !nz); e->old_z = nz; }
}

/* luke::update_all = update_stepper + update_objects (src/myfunctions.cpp:2110-2214),
 * ObjectHandler::apply_antiroll (src/objecthandler.cpp:1034-1059) */
static void update_all(OrEnv* e) {
  const double time_per_step = 10 / 2000.0;         /* num_steps / pulses_per_s, :532 */
  if (e->time > e->last_step_time + time_per_step) {
    update_constraints(e);
    e->last_step_time = e->time;
    or_gripper_step_to(&e->tgt_next, &e->tgt_end, 10);
  }
  if (e->has_object) {
    double* v = e->qvel + e->B;
    if (sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]) < 1e-6) for (int k = 0; k < 6; k++) v[k] = 0;
  }
}

/* ------------------------------------------------------------------ sensors */
static float apply_normalisation(const double* sn, float value) {   /* src/mjclass.h:155-174 */
  if (!sn[GSF_use_normalisation]) return value;
  float normalise = (float)sn[GSF_normalise];
  if (normalise <= 0) { return value < 0 ? -1.0f : 1.0f; }
  else if (value > normalise) return 1.0f;
  if (value < -normalise) return -1.0f;
  return value / normalise;
}

static float apply_noise(OrEnv* e, int sensor, float value, int i) { /* src/mjclass.h:176-223 */
  const double* sn = e->s.sensor[sensor];
  if (!sn[GSF_use_noise] || !e->s.general[GS_use_noise_seeded]) return value;
  const float eps = 1.1920929e-07f;
  float mu = e->rand_mu[sensor][i - 1];
  float noise_std = (float)sn[GSF_noise_std], noise_mag = (float)sn[GSF_noise_mag];
  if (noise_std < eps) {
    float noise = mu + noise_mag * (2 * or_uniform(e) - 1);
    value += noise;
  } else {
    float u1, u2;
    do { u1 = or_uniform(e); } while (u1 <= eps);
    u2 = or_uniform(e);
    float mag = (float)(noise_std * sqrt(-2.0 * log((double)u1)));
    float z0 = (float)(mag * cos((double)(6.2831853071795864769f * u2))) + mu;
    value += z0;
  }
  if (value > 1) value = 1; else if (value < -1) value = -1;
  return value;
}

static int ready_to_read(OrEnv* e, int sensor) {                     /* src/mjclass.h:225-241 */
  float rr = (float)e->s.sensor[sensor][GSF_read_rate];
  double between = 1 / rr;
  if (e->time > e->last_read_time[sensor] + between) { e->last_read_time[sensor] = e->time; return 1; }
  return 0;
}

/* luke::read_armadillo_gauge (src/myfunctions.cpp:2699-2795): cubic least squares through the
 * N+1 planar joint positions, evaluated at x = 50 mm, returned in mm. arma::polyfit solves the
 * Vandermonde system by QR; Householder QR here. */
float or_read_gauge(const OrEnv* e, int finger) {
  int N = e->m.N;
  double L = e->m.segment_length;
  double X[GMJ_MAXB + 1][4], y[GMJ_MAXB + 1], fx[GMJ_MAXB + 1];
  double cum = 0; fx[0] = 0; y[0] = 0;
  for (int i = 0; i < N; i++) {
    cum += e->qpos[3 + finger * (N + 2) + i];
    fx[i + 1] = fx[i] + L * cos(cum);
    y[i + 1] = y[i] + L * sin(cum);
  }
  int n = N + 1;
  for (int i = 0; i < n; i++) { X[i][0] = fx[i] * fx[i] * fx[i]; X[i][1] = fx[i] * fx[i]; X[i][2] = fx[i]; X[i][3] = 1; }
  for (int k = 0; k < 4; k++) {
    double nrm = 0;
    for (int i = k; i < n; i++) nrm += X[i][k] * X[i][k];
    nrm = sqrt(nrm);
    if (nrm == 0) continue;
    double alpha = X[k][k] > 0 ? -nrm : nrm;
    double v[GMJ_MAXB + 1], vn = 0;
    for (int i = k; i < n; i++) { v[i] = X[i][k]; }
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
  const double xp = 50e-3;   /* std::pow(xpos, order - i), written as products (same on the device path) */
  yv = (float)(yv + coeff[0] * (xp * xp * xp));
  yv = (float)(yv + coeff[1] * (xp * xp));
  yv = (float)(yv + coeff[2] * xp);
  yv = (float)(yv + coeff[3] * 1.0);
  return yv * 1000;
}

/* ObjectHandler::extract_forces_faster (src/objecthandler.cpp:737-992), classification by
 * precomputed geom class instead of geom-name prefixes (src/objecthandler.h:110-129) */
typedef struct {
  double obj_f[3][3], obj_palm[3], obj_gnd[3];       /* on the live object, world frame */
  double all_f[3][3], all_palm[3], gnd_f[3][3];
  double obj_f_loc[3][3], obj_palm_loc[3], all_f_loc[3][3], all_palm_loc[3], gnd_f_loc[3][3];
} OrForces;

static void extract_forces(const OrEnv* e, OrForces* F) {
  memset(F, 0, sizeof(*F));
  if (!e->has_object) return;                      /* live_objects.size() == 0 -> empty */
  if (e->script) {
    const double* c = e->script_cur;
    for (int k = 0; k < 3; k++) {
      for (int f = 0; f < 3; f++) { F->all_f_loc[f][k] = c[3 + 3 * f + k]; F->obj_f_loc[f][k] = c[15 + 3 * f + k]; F->gnd_f_loc[f][k] = c[30 + 3 * f + k]; }
      F->all_palm_loc[k] = c[12 + k]; F->obj_palm_loc[k] = c[24 + k]; F->obj_gnd[k] = c[27 + k];
    }
    return;
  }
  for (int c = 0; c < e->ncon; c++) {
    const OrContact* cn = &e->con[c];
    int a = cn->cls1, b = cn->cls2;
    double g[3] = { 0, 0, 0 };
    for (int k = 0; k < 3; k++) axpy3(g, cn->force[k], cn->frame + 3 * k);
    int with_obj = (a == GMJ_CLS_OBJECT || b == GMJ_CLS_OBJECT);
    int with_gnd = (a == GMJ_CLS_GROUND || b == GMJ_CLS_GROUND);
    int with_palm = (a == GMJ_CLS_PALM || b == GMJ_CLS_PALM);
    for (int f = 0; f < 3; f++) {
      int with_f = (a == GMJ_CLS_F1 + f || b == GMJ_CLS_F1 + f);
      if (with_obj && with_f) add3(F->obj_f[f], F->obj_f[f], g);
      if (with_f) { add3(F->all_f[f], F->all_f[f], g); if (with_gnd) add3(F->gnd_f[f], F->gnd_f[f], g); }
    }
    if (with_obj && with_palm) add3(F->obj_palm, F->obj_palm, g);
    if (with_obj && with_gnd) add3(F->obj_gnd, F->obj_gnd, g);
    if (with_palm) add3(F->all_palm, F->all_palm, g);
  }
  for (int f = 0; f < 3; f++) {
    const double* R = e->xmat[e->m.finger_body[f]];
    mulTv(F->obj_f_loc[f], R, F->obj_f[f]);
    mulTv(F->all_f_loc[f], R, F->all_f[f]);
    mulTv(F->gnd_f_loc[f], R, F->gnd_f[f]);
  }
  mulTv(F->obj_palm_loc, e->xmat[e->m.palm_body], F->obj_palm);
  mulTv(F->all_palm_loc, e->xmat[e->m.palm_body], F->all_palm);
}

/* MjClass::monitor_sensors (src/mjclass.cpp:741-898) */
static void monitor_sensors(OrEnv* e) {
  int got = 0; OrForces F;
  const double* g = e->s.general;
  if (ready_to_read(e, GSEN_bending_gauge)) {
    for (int f = 0; f < 3; f++) {
      float gv = e->script ? (float)e->script_cur[f] : or_read_gauge(e, f);
      or_win_add(&e->si_gauge[f], (float)((double)gv * g[GS_sim_gauge_raw_to_N_factor]));   /* gfloat * double member, src/mjclass.cpp:755-757 */
      gv = apply_normalisation(e->s.sensor[GSEN_bending_gauge], gv);
      gv = apply_noise(e, GSEN_bending_gauge, gv, f + 1);
      or_win_add(&e->w_gauge[f], gv);
    }
  }
  if (ready_to_read(e, GSEN_axial_gauge)) {
    if (!got) { extract_forces(e, &F); got = 1; }
    for (int f = 0; f < 3; f++) {
      float av = (float)F.all_f_loc[f][0];
      av = apply_normalisation(e->s.sensor[GSEN_axial_gauge], av);
      av = apply_noise(e, GSEN_axial_gauge, av, f + 1);
      or_win_add(&e->w_axial[f], av);
    }
  }
  if (ready_to_read(e, GSEN_palm)) {
    if (!got) { extract_forces(e, &F); got = 1; }
    float pv = (float)F.all_palm_loc[0];
    pv = (float)(pv * g[GS_palm_scale_factor]);
    or_win_add(&e->si_palm, pv);
    pv = apply_normalisation(e->s.sensor[GSEN_palm], pv);
    pv = apply_noise(e, GSEN_palm, pv, 1);
    or_win_add(&e->w_palm, pv);
  }
  if (ready_to_read(e, GSEN_wrist_Z)) {
    float z = (float)e->wrist_z;
    z -= (float)e->s.sensor[GSEN_wrist_Z][GSF_raw_value_offset];
    or_win_add(&e->si_wristz, z);
    z = apply_normalisation(e->s.sensor[GSEN_wrist_Z], z);
    z = apply_noise(e, GSEN_wrist_Z, z, 1);
    or_win_add(&e->w_wristz, z);
  }
}

/* MjClass::step (src/mjclass.cpp:504-530) */
void or_substep(OrEnv* e) {
  e->tip_force = e->s.general[GS_curve_validation] ? e->s.general[GS_tip_force_applied] : 0.0;
  if (e->s.general[GS_curve_validation]) {           /* apply_segment_force locks these welds */
    for (int i = 0; i < 6; i++) or_set_weld(e, i, 1);
  }
  or_physics_step(e);
  update_all(e);
  monitor_sensors(e);
}

/* calibrate_reset first call (src/myfunctions.cpp:1470-1476): before/step/after x n, welds off */
void or_settle(OrEnv* e, int nsteps, double dt) {
  double keep = e->s.general[GS_mujoco_timestep];
  e->s.general[GS_mujoco_timestep] = dt;
  for (int i = 0; i < nsteps; i++) { e->tip_force = 0; or_physics_step(e); update_all(e); }
  e->s.general[GS_mujoco_timestep] = keep;
  e->reset_qpos[0] = e->qpos[0];
  for (int f = 0; f < 3; f++) { e->reset_qpos[1 + 2 * f] = e->qpos[1 + f * (e->m.N + 2)]; e->reset_qpos[2 + 2 * f] = e->qpos[2 + f * (e->m.N + 2)]; }
  e->reset_qpos[7] = e->qpos[e->B - 1];
}

/* ------------------------------------------------------------------ lifecycle */
static void configure_settings(OrEnv* e) {          /* src/mjclass.cpp:97-314 (non-calibration part) */
  const GmjSettings* s = &e->s;
  int n = 0, cont = (int)s->general[GS_continous_actions];
  for (int a = 0; a < GACT_COUNT; a++) {
    if (!s->action[a][GAF_in_use]) continue;
    if (cont) { e->action_options[n++] = 3 * a + 2; }
    else { e->action_options[n++] = 3 * a; e->action_options[n++] = 3 * a + 1; }
  }
  if (s->general[GS_use_termination_action]) e->action_options[n++] = 3 * GACT_COUNT;
  e->n_actions = n;
  /* Settings::apply_noise_params (src/mjclass.cpp:5293-5346): 3 mean offsets per sensor, state
   * sensors then re-drawn with state_noise_mu */
  for (int k = 0; k < GSEN_COUNT; k++) {
    float mu = (float)(s->sensor[k][GSF_noise_overriden] ? s->sensor[k][GSF_noise_mu] : s->general[GS_sensor_noise_mu]);
    for (int j = 0; j < 3; j++) e->rand_mu[k][j] = mu * (2 * or_uniform(e) - 1);
  }
  static const int state_sensors[5] = { GSEN_motor_state, GSEN_base_XY, GSEN_base_Z, GSEN_base_yaw, GSEN_cartesian };
  for (int q = 0; q < 5; q++) {
    int k = state_sensors[q];
    float mu = (float)(s->sensor[k][GSF_noise_overriden] ? s->sensor[k][GSF_noise_mu] : s->general[GS_state_noise_mu]);
    for (int j = 0; j < 3; j++) e->rand_mu[k][j] = mu * (2 * or_uniform(e) - 1);
  }
  /* Settings::update_sensor_settings (src/mjclass.cpp:5236-5265, quirk Q11) */
  double tps = s->general[GS_mujoco_timestep] * (int)s->general[GS_sim_steps_per_action];
  for (int k = 0; k < GSEN_COUNT; k++) {
    int is_state = (k == GSEN_motor_state || k == GSEN_base_XY || k == GSEN_base_Z || k == GSEN_base_yaw || k == GSEN_cartesian);
    if (is_state) { e->prev_steps[k] = (int)s->general[GS_state_n_prev_steps]; e->readings_per_step[k] = 1; }
    else {
      e->prev_steps[k] = (int)s->general[GS_sensor_n_prev_steps];
      float t = (float)tps, rr = (float)s->sensor[k][GSF_read_rate];
      double rs = t * rr;
      e->readings_per_step[k] = (int)floor(rs);
    }
    e->total_readings[k] = 1 + e->readings_per_step[k] * e->prev_steps[k];
  }
}

void or_set_settings(OrEnv* e, const GmjSettings* s) {
  e->s = *s;
  e->reset_qpos[0] = s->general[GS_reset_qpos_base];
  for (int k = 0; k < 7; k++) e->reset_qpos[1 + k] = s->general[GS_reset_qpos_g0 + k];
}

OrEnv* or_create(const GmjModel* m, const GmjSettings* s) {
  OrEnv* e = (OrEnv*)calloc(1, sizeof(OrEnv));
  e->m = *m;
  e->B = m->B; e->nv = m->B + 6; e->nq = m->B + 7;
  or_set_settings(e, s);
  static const int wb[7][2] = { {0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0} };
  (void)wb;
  for (int f = 0; f < 3; f++) {
    int ip = 1 + f * (m->N + 2);
    e->weld[f].body1 = 0; e->weld[f].body2 = ip;              /* pris{f}_weld: base <-> intermediate */
    e->weld[3 + f].body1 = ip + 1; e->weld[3 + f].body2 = ip; /* rev{f}_weld: finger <-> intermediate */
  }
  e->weld[6].body1 = 0; e->weld[6].body2 = m->B - 1;          /* palm_weld */
  or_reset(e, 0, 0, 0);
  return e;
}

void or_destroy(OrEnv* e) { free(e); }
int or_struct_size(void) { return (int)sizeof(OrEnv); }

static void reset_windows(OrEnv* e) {
  for (int k = 0; k < 3; k++) { or_win_reset(&e->w_gauge[k]); or_win_reset(&e->w_axial[k]); or_win_reset(&e->w_motor[k]); or_win_reset(&e->si_gauge[k]); }
  or_win_reset(&e->w_palm); or_win_reset(&e->w_wristz); or_win_reset(&e->w_basexy[0]); or_win_reset(&e->w_basexy[1]);
  or_win_reset(&e->w_basez); or_win_reset(&e->w_baseyaw); or_win_reset(&e->si_palm); or_win_reset(&e->si_wristz);
  for (int k = 0; k < 12; k++) or_win_reset(&e->w_cart[k]);
}

/* luke::reset (src/myfunctions.cpp:544-570) + MjClass::reset (src/mjclass.cpp:434-486) */
void or_reset(OrEnv* e, uint64_t seed, uint32_t env_id, uint32_t episode) {
  const GmjModel* m = &e->m;
  int B = e->B;
  e->seed = seed; e->env_id = env_id; e->rng_ctr = episode << 12;
  or_gripper_reset(&e->tgt_end); or_gripper_reset(&e->tgt_next);
  e->base_x = e->base_y = e->base_z = e->base_yaw = 0;
  for (int i = 0; i < 7; i++) e->weld[i].on = 0;
  /* mj_resetData + keyframe "initial pose"; live object back to its parked pose (dead group) */
  e->time = 0; e->last_step_time = 0;
  memset(e->qvel, 0, sizeof(e->qvel));
  for (int i = 0; i < B; i++) e->qpos[i] = m->qpos0[i];
  e->has_object = 0;
  double* q = e->qpos + B;
  q[0] = 0; q[1] = 2; q[2] = 1; q[3] = 0; q[4] = 0; q[5] = 0; q[6] = 1;
  or_kinematics(e);                                   /* mj_forward at the keyframe pose */
  e->ncon = 0;
  /* calibrate_reset: write the stored equilibrium (no mj_forward follows, so the weld anchors
   * below come from the keyframe-pose xpos/xmat, src/myfunctions.cpp:565-569) */
  e->qpos[0] = e->reset_qpos[0];
  for (int f = 0; f < 3; f++) { e->qpos[1 + f * (m->N + 2)] = e->reset_qpos[1 + 2 * f]; e->qpos[2 + f * (m->N + 2)] = e->reset_qpos[2 + 2 * f]; }
  e->qpos[B - 1] = e->reset_qpos[7];
  set_all_welds(e, 1);
  e->old_x = e->old_y = e->old_z = 0;                 /* quirk Q12: per-env, cleared at reset */
  /* MjClass::reset */
  reset_windows(e);
  memset(e->last_read_time, 0, sizeof(e->last_read_time));
  memset(e->bin, 0, sizeof(e->bin)); memset(e->lin, 0, sizeof(e->lin));
  e->cumulative_reward = 0; e->num_action_steps = 0; e->termination_signal_sent = 0;
  e->peak_finger_lateral_force = 0; e->start_z = 0;
  e->unstable = 0; e->nrow_prev = 0; e->tip_force = 0; e->wrist_z = 0;
  configure_settings(e);
  /* random_base_Z_movement (src/mjclass.cpp:1423-1434) -> set_base_to_Z_position
   * (src/myfunctions.cpp:2610-2640; float argument) */
  double size = e->s.general[GS_base_position_noise];
  double zmove = (e->s.general[GS_use_noise_seeded] && size > 0) ? (-size + 2 * size * (double)or_uniform(e)) : 0.0;
  float zf = (float)zmove;
  e->base_z = zf;
  if (e->base_z > e->s.general[GS_base_z_max]) e->base_z = e->s.general[GS_base_z_max];
  if (e->base_z < e->s.general[GS_base_z_min]) e->base_z = e->s.general[GS_base_z_min];
  e->qpos[0] = e->base_z + e->reset_qpos[0];
}

/* MjClass::spawn_object (src/mjclass.cpp:2352-2420) + ObjectHandler::spawn_object
 * (src/objecthandler.cpp:403-428): pose quaternion = keyframe (0,0,0,1) composed with a
 * rotation of -zrot, z = keyframe z + 1e-6, zero velocity, collision group -> main */
void or_spawn(OrEnv* e, const GmjObject* o, double x, double y, double zrot) {
  e->obj = *o; e->has_object = 1;
  double* q = e->qpos + e->B;
  q[0] = x; q[1] = y; q[2] = o->z_rest + 1e-6;
  q[3] = sin(-zrot / 2.0); q[4] = 0; q[5] = 0; q[6] = cos(-zrot / 2.0);
  for (int k = 0; k < 6; k++) e->qvel[e->B + k] = 0;
  e->start_z = q[2];
  e->nrow_prev = 0;
  if (e->script) { e->script_cur[40] = q[0]; e->script_cur[41] = q[1]; e->script_cur[42] = q[2]; return; }
  or_forward(e);
}

/* ------------------------------------------------------------------ spawn_into_scene (src/mjclass.cpp:2475-2654)
 * Fingertip boxes as Env::reset takes them right after luke::reset (src/mjclass.h:893-904,
 * luke::get_finger_hook_locations src/myfunctions.cpp:3717-3761): target at home, base at the origin. */
static void fingertip_boxes(const OrEnv* e, double tips[3][8]) {
  OrGripper home; or_gripper_reset(&home);
  const double hook_th = e->m.hook_angle_deg * (M_PI / 180.000);
  const double PI_23 = M_PI * (2.0 / 3.0);
  for (int i = 0; i < 3; i++) {
    double ang = i * PI_23;
    double hook_x = 0.5 * e->m.hook_length * sin(hook_th);
    double px = -(home.x - hook_x) * sin(ang) + 0.0, py = -(home.x - hook_x) * cos(ang) + 0.0;
    or_box2d_init(tips[i], px, py, e->m.finger_width, e->m.hook_length, -ang);
  }
}
/* the accept test of ONE candidate pose (:2594-2640): outer bounds (SpawnParams defaults +-100 m), then Box2d SAT with
 * a 1 mm gap against the three fingertip boxes; single live object, so no object-object test */
int or_spawn_ok(const OrEnv* e, double x, double y, double rot, const double* bbox) {
  double ours[8], tips[3][8];
  or_box2d_init(ours, x, y, bbox[0], bbox[1], rot);
  for (int i = 0; i < 4; i++) if (ours[i] < -100 || ours[i] > 100 || ours[4 + i] < -100 || ours[4 + i] > 100) return 0;
  fingertip_boxes(e, tips);
  for (int i = 0; i < 3; i++) if (or_box2d_overlap(ours, tips[i], 1e-3)) return 0;
  return 1;
}
/* Candidate grid (2 mm steps within +-xrange / +-yrange, pi/30 steps within +-rotrange), both lists shuffled, walked
 * pairwise for max(n_xy, n_rot) tries, first accepted pose spawns; 0 when none is (the caller retries,
 * rl/env/MjEnv.py:1216-1244). std::shuffle on the global engine becomes a Fisher-Yates walk on the counter RNG
 * (same on the device): j = (int)(u * (i + 1)) for i = n-1 .. 1. */
#define OR_SPAWN_MAXXY 1024
#define OR_SPAWN_MAXROT 128
static void fy_shuffle(OrEnv* e, unsigned short* a, int n) {
  for (int i = n - 1; i >= 1; i--) {
    int j = (int)(or_uniform(e) * (float)(i + 1)); if (j > i) j = i;
    unsigned short t = a[i]; a[i] = a[j]; a[j] = t;
  }
}
int or_spawn_into_scene(OrEnv* e, const GmjObject* o, double xpos, double ypos, double zrot, double xrange, double yrange,
                        double rotrange) {
  const double inc = 2e-3, rinc = M_PI / 30.0;
  int num_x = (int)((2 * xrange) / inc) + 1, num_y = (int)((2 * yrange) / inc) + 1, num_r = (int)((2 * rotrange) / rinc) + 1;
  int nxy = num_x * num_y;
  if (nxy > OR_SPAWN_MAXXY || num_r > OR_SPAWN_MAXROT) return -1;
  unsigned short pxy[OR_SPAWN_MAXXY], pr[OR_SPAWN_MAXROT];
  for (int i = 0; i < nxy; i++) pxy[i] = (unsigned short)i;
  for (int i = 0; i < num_r; i++) pr[i] = (unsigned short)i;
  if (nxy > 1) fy_shuffle(e, pxy, nxy);
  if (num_r > 1) fy_shuffle(e, pr, num_r);
  int tries = nxy > num_r ? nxy : num_r;
  for (int i = 0; i < tries; i++) {
    int k = pxy[i % nxy], ix = k / num_y, iy = k % num_y, ir = pr[i % num_r];
    double x = num_x > 1 ? -xrange + ix * inc + xpos : xpos;
    double y = num_y > 1 ? -yrange + iy * inc + ypos : ypos;
    double r = num_r > 1 ? -rotrange + ir * rinc + zrot : zrot;
    if (or_spawn_ok(e, x, y, r, o->bbox)) { or_spawn(e, o, x, y, r); return 1; }
  }
  return 0;
}
/* MjEnv._spawn_object on one env (rl/env/MjEnv.py:1216-1244): up to `tries` spawn_into_scene attempts, then the "old
 * method" -- a pose on the millimetre grid within +-fb_xy_mm with a base angle 0 / 60 / 120 degrees +-fb_rot_deg and no
 * collision test (skipped when fb_xy_mm < 0). Returns 1 (search), 2 (fallback) or 0. Same draws as the device. */
int or_spawn_retry(OrEnv* e, const GmjObject* o, double xrange, double rotrange, int tries, int fb_xy_mm, int fb_rot_deg) {
  for (int a = 0; a < tries; a++) {
    int ok = or_spawn_into_scene(e, o, 0.0, 0.0, 0.0, xrange, xrange, rotrange);
    if (ok != 0) return ok;
  }
  if (fb_xy_mm < 0) return 0;
  int xm = (int)(or_uniform(e) * (float)(2 * fb_xy_mm + 1)); if (xm > 2 * fb_xy_mm) xm = 2 * fb_xy_mm;
  int ym = (int)(or_uniform(e) * (float)(2 * fb_xy_mm + 1)); if (ym > 2 * fb_xy_mm) ym = 2 * fb_xy_mm;
  int an = (int)(or_uniform(e) * (float)(2 * fb_rot_deg + 1)); if (an > 2 * fb_rot_deg) an = 2 * fb_rot_deg;
  int bi = (int)(or_uniform(e) * 3.0f); if (bi > 2) bi = 2;
  or_spawn(e, o, (xm - fb_xy_mm) * 1e-3, (ym - fb_xy_mm) * 1e-3, (60 * bi + an - fb_rot_deg) * (M_PI / 180.0));
  return 2;
}
/* object index draw of the batched env (device: same draw, before the pose search) */
int or_draw_object(OrEnv* e, int n_objects) {
  int oi = (int)(or_uniform(e) * n_objects); if (oi >= n_objects) oi = n_objects - 1;
  return oi;
}

void or_set_fast(OrEnv* e, int on) { e->fast_solver = on; }
void or_set_self_collision(OrEnv* e, int on) { e->self_collision = on; }

void or_set_script(OrEnv* e, const double* rows, int nrows) {
  e->script = rows; e->script_rows = nrows; e->script_cursor = 0;
  memset(e->script_cur, 0, sizeof(e->script_cur));
}

/* base target clamp shared by move_base_target_m / lift_base_to_height (src/myfunctions.cpp:2444-2529) */
static int clamp_base(OrEnv* e) {
  const double* g = e->s.general; int wl = 1;
  if (e->base_x > g[GS_base_x_max]) { e->base_x = g[GS_base_x_max]; wl = 0; }
  if (e->base_x < g[GS_base_x_min]) { e->base_x = g[GS_base_x_min]; wl = 0; }
  if (e->base_y > g[GS_base_y_max]) { e->base_y = g[GS_base_y_max]; wl = 0; }
  if (e->base_y < g[GS_base_y_min]) { e->base_y = g[GS_base_y_min]; wl = 0; }
  if (e->base_z > g[GS_base_z_max]) { e->base_z = g[GS_base_z_max]; wl = 0; }
  if (e->base_z < g[GS_base_z_min]) { e->base_z = g[GS_base_z_min]; wl = 0; }
  return wl;
}

/* ActionSetting::call_action_function (src/mjclass.h:546-571) */
static int call_action(OrEnv* e, int a, double call_value) {
  call_value *= (int)e->s.action[a][GAF_sign];
  OrGripper* t = &e->tgt_end;
  switch (a) {
    case GACT_gripper_X: return or_gripper_set_xyz_m(t, t->x + call_value, t->y, t->z);
    case GACT_gripper_Y: return or_gripper_set_xyz_m(t, t->x, t->y + call_value, t->z);
    case GACT_gripper_Z: return or_gripper_set_xyz_m(t, t->x, t->y, t->z + call_value);
    case GACT_gripper_prismatic_X: return or_gripper_set_xyz_m_rad(t, t->x + call_value, t->th, t->z);
    case GACT_gripper_revolute_Y: return or_gripper_set_xyz_m_rad(t, t->x, t->th + call_value, t->z);
    case GACT_base_X: e->base_x += call_value; return clamp_base(e);
    case GACT_base_Y: e->base_y += call_value; return clamp_base(e);
    case GACT_base_Z: e->base_z += call_value; return clamp_base(e);
    case GACT_base_roll: case GACT_base_pitch: case GACT_base_yaw: {
      /* move_base_target_rad only ever applies its third argument (yaw); roll/pitch actions pass 0 there */
      int wl = 1; const double* g = e->s.general;
      if (a == GACT_base_yaw) e->base_yaw += call_value;
      if (e->base_yaw > g[GS_base_yaw_max]) { e->base_yaw = g[GS_base_yaw_max]; wl = 0; }
      if (e->base_yaw < g[GS_base_yaw_min]) { e->base_yaw = g[GS_base_yaw_min]; wl = 0; }
      return wl;
    }
  }
  return 1;
}

/* MjClass::set_action (src/mjclass.cpp:1528-1630). Returns 0 ok, -1 out-of-range action. */
int or_set_action(OrEnv* e, int action, float fraction) {
  int wl = 1;
  e->termination_signal_sent = 0;
  if (action < 0 || action >= e->n_actions) return -1;
  int code = e->action_options[action];
  const GmjSettings* s = &e->s;
  if (code == 3 * GACT_COUNT) {
    float value = s->general[GS_continous_actions] ? fraction : 1.0f;
    if (value > (float)s->general[GS_termination_threshold]) {
      e->termination_signal_sent = 1;
      if (s->general[GS_lift_on_termination]) {
        e->base_z = -s->general[GS_base_z_max];       /* lift_base_to_height(base_max_[2]) */
        if (e->base_z > s->general[GS_base_z_max]) e->base_z = s->general[GS_base_z_max];
        if (e->base_z < s->general[GS_base_z_min]) e->base_z = s->general[GS_base_z_min];
        int S = (int)s->general[GS_sim_steps_per_action];
        for (int i = 0; i < S * 2; i++) or_substep(e);  /* quirk Q8 */
      }
    }
    wl = 1;
  } else {
    int a = code / 3, kind = code % 3;
    double v = s->action[a][GAF_value];
    if (kind == 0) wl = call_action(e, a, v);
    else if (kind == 1) wl = call_action(e, a, -1 * v);
    else {
      wl = call_action(e, a, v * fraction);
      e->lin[GLE_action_penalty_lin].value += fabsf(fraction);
      e->lin[GLE_action_penalty_sq].value += fraction * fraction;
    }
  }
  /* luke::get_fingertip_z_height (src/myfunctions.cpp:3608-3620), float arithmetic */
  {
    float bmin = (float)s->general[GS_base_z_min];
    float straight = (float)(-s->general[GS_base_z_min] - e->base_z);
    float tip_lift = (float)(e->m.finger_length * (1 - cos(or_gripper_th(&e->tgt_end))));
    float above = straight + tip_lift;
    float h = (float)((double)above + s->general[GS_base_z_min]);
    (void)bmin;
    if (h < s->general[GS_fingertip_min_mm] * 1e-3) wl = 0;
  }
  e->bin[GBE_exceed_limits].value += !wl;
  if (e->bin[GBE_exceed_limits].value != 0) e->bin[GBE_exceed_limits].value = 1;  /* bool += */
  return 0;
}

/* MjClass::sense_gripper_state (src/mjclass.cpp:900-964) */
static void sense_gripper_state(OrEnv* e) {
  const double* g = e->s.general;
  float sx = or_normalise_between((float)e->tgt_end.x, (float)49e-3, (float)134e-3);
  float sy = or_normalise_between((float)e->tgt_end.y, (float)49e-3, (float)134e-3);
  float sz = or_normalise_between((float)e->tgt_end.z, (float)0e-3, (float)165e-3);
  float bx = or_normalise_between((float)e->base_x, (float)g[GS_base_x_min], (float)g[GS_base_x_max]);
  float by = or_normalise_between((float)e->base_y, (float)g[GS_base_y_min], (float)g[GS_base_y_max]);
  float bz = or_normalise_between((float)e->base_z, (float)g[GS_base_z_min], (float)g[GS_base_z_max]);
  float byaw = or_normalise_between((float)e->base_yaw, (float)g[GS_base_yaw_min], (float)g[GS_base_yaw_max]);
  /* JointStates fields are double: the normalised float is stored to double, noise is applied
   * on a float copy and stored back, then add() narrows to float again */
  sx = apply_noise(e, GSEN_motor_state, sx, 1);
  sy = apply_noise(e, GSEN_motor_state, sy, 2);
  sz = apply_noise(e, GSEN_motor_state, sz, 3);
  bx = apply_noise(e, GSEN_base_XY, bx, 1);
  by = apply_noise(e, GSEN_base_XY, by, 2);
  bz = apply_noise(e, GSEN_base_Z, bz, 1);
  byaw = apply_noise(e, GSEN_base_yaw, byaw, 1);
  or_win_add(&e->w_motor[0], sx); or_win_add(&e->w_motor[1], sy); or_win_add(&e->w_motor[2], sz);
  or_win_add(&e->w_basexy[0], bx); or_win_add(&e->w_basexy[1], by);
  or_win_add(&e->w_basez, bz); or_win_add(&e->w_baseyaw, byaw);
  /* cartesian contact points for MAT: luke::get_fingerend_and_palm_xyz (src/myfunctions.cpp:3622-3688) */
  double F[3] = { or_win_read(&e->si_gauge[0], 0), or_win_read(&e->si_gauge[1], 0), or_win_read(&e->si_gauge[2], 0) };
  double palmF = or_win_read(&e->si_palm, 0);
  double fing_x = e->tgt_end.x, fing_th = or_gripper_th(&e->tgt_end), palm_z = e->tgt_end.z;
  double L = e->m.finger_length;
  double untilted = e->m.fingertip_clearance - e->base_z;
  double tilted = untilted + L * (1 - cos(fing_th));
  const double ft = 0.2, PI_23 = M_PI * (2.0 / 3.0);
  for (int i = 0; i < 3; i++) {
    double ang = i * PI_23;
    double tilt_x = fing_x - L * sin(fing_th);
    double defl = F[i] * (L * L * L) / (3 * e->m.EI);
    double fx = tilt_x + defl * cos(fing_th);
    double px = -fx * sin(ang + e->base_yaw) + e->base_x;
    double py = -fx * cos(ang + e->base_yaw) + e->base_y;
    double pz = tilted - defl * sin(fing_th);
    int on = fabs(F[i]) > ft;
    or_win_add(&e->w_cart[3 * i + 0], on ? (float)px : 0.0f);
    or_win_add(&e->w_cart[3 * i + 1], on ? (float)py : 0.0f);
    or_win_add(&e->w_cart[3 * i + 2], on ? (float)pz : 0.0f);
  }
  int on = palmF > ft;
  or_win_add(&e->w_cart[9], on ? (float)e->base_x : 0.0f);
  or_win_add(&e->w_cart[10], on ? (float)e->base_y : 0.0f);
  or_win_add(&e->w_cart[11], on ? (float)(untilted + 165e-3 - palm_z) : 0.0f);
}

/* MjClass::action_step (src/mjclass.cpp:1483-1508) */
void or_action_step(OrEnv* e) {
  int S = (int)e->s.general[GS_sim_steps_per_action];
  for (int i = 0; i < S; i++) or_substep(e);
  sense_gripper_state(e);
  or_update_env(e);
  e->num_action_steps += 1;
}

/* exported for events.c */
void or_extract_forces_flat(const OrEnv* e, double* out) {
  OrForces F; extract_forces(e, &F);
  memcpy(out, &F, sizeof(F));
}
int or_forces_size(void) { return (int)(sizeof(OrForces) / sizeof(double)); }

/* flat state view for parity checks (mirrors OrPeek in oracle/binding.py) */
typedef struct {
  double qpos[96], qvel[96], time, wrist_z;
  int ncon, nrow, niter, unstable;
  double tgt_end[4], tgt_next[4];
  int tgt_steps[6];
  double base_z;
  int weld_on[7], pad;
  double reset_qpos[8];
  float si[5], pad2;
  double con_dist[24], con_force[72], con_pos[72], con_normal[72];
  int con_cls[48];
  int64_t substeps, contacts, rows, iters;
} OrPeek;

void or_peek(const OrEnv* e, OrPeek* p) {
  memset(p, 0, sizeof(*p));
  for (int i = 0; i < e->nq && i < 96; i++) p->qpos[i] = e->qpos[i];
  for (int i = 0; i < e->nv && i < 96; i++) p->qvel[i] = e->qvel[i];
  p->time = e->time; p->wrist_z = e->wrist_z;
  p->ncon = e->ncon; p->nrow = e->nrow; p->niter = e->niter; p->unstable = e->unstable;
  p->tgt_end[0] = e->tgt_end.x; p->tgt_end[1] = e->tgt_end.y; p->tgt_end[2] = e->tgt_end.z; p->tgt_end[3] = e->tgt_end.th;
  p->tgt_next[0] = e->tgt_next.x; p->tgt_next[1] = e->tgt_next.y; p->tgt_next[2] = e->tgt_next.z; p->tgt_next[3] = e->tgt_next.th;
  p->tgt_steps[0] = e->tgt_end.sx; p->tgt_steps[1] = e->tgt_end.sy; p->tgt_steps[2] = e->tgt_end.sz;
  p->tgt_steps[3] = e->tgt_next.sx; p->tgt_steps[4] = e->tgt_next.sy; p->tgt_steps[5] = e->tgt_next.sz;
  p->base_z = e->base_z;
  for (int i = 0; i < 7; i++) p->weld_on[i] = e->weld[i].on;
  for (int i = 0; i < 8; i++) p->reset_qpos[i] = e->reset_qpos[i];
  for (int k = 0; k < 3; k++) p->si[k] = or_win_read(&e->si_gauge[k], 0);
  p->si[3] = or_win_read(&e->si_palm, 0); p->si[4] = or_win_read(&e->si_wristz, 0);
  for (int c = 0; c < e->ncon; c++) {
    p->con_dist[c] = e->con[c].dist;
    for (int k = 0; k < 3; k++) { p->con_force[3 * c + k] = e->con[c].force[k]; p->con_pos[3 * c + k] = e->con[c].pos[k]; p->con_normal[3 * c + k] = e->con[c].frame[k]; }
    p->con_cls[2 * c] = e->con[c].cls1; p->con_cls[2 * c + 1] = e->con[c].cls2;
  }
  p->substeps = e->st_substeps; p->contacts = e->st_contacts; p->rows = e->st_rows; p->iters = e->st_iters;
}
