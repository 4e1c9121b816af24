/* events.c -- oracle: update_env, event counters, reward, done, observation.
 * TEST INFRASTRUCTURE ONLY (see oracle.h). Restates src/mjclass.cpp:966-1346, 1632-1959,
 * 3000-3049, 5437-5528, 5594-5615 for a single live object.
 */
#include <math.h>
#include <string.h>
#include "oracle.h"

void or_extract_forces_flat(const OrEnv* e, double* out);

typedef struct {
  double obj_f[3][3], obj_palm[3], obj_gnd[3];
  double all_f[3][3], all_palm[3], gnd_f[3][3];
  double obj_f_loc[3][3], obj_palm_loc[3], all_f_loc[3][3], all_palm_loc[3], gnd_f_loc[3][3];
} Forces;

static float mag3(const double* v) { return (float)sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }

/* update_events (src/mjclass.cpp:5437-5469) */
static void update_events(OrEnv* e) {
  for (int k = 0; k < GBE_COUNT; k++) {
    OrEvent* ev = &e->bin[k];
    int v = ev->value != 0;
    ev->row = ev->row * v + v;
    ev->abs_ += v;
    ev->last_value = (float)v;
    ev->active_sum = ev->row != 0;
    ev->value = 0;
  }
  for (int k = 0; k < GLE_COUNT; k++) {
    OrEvent* ev = &e->lin[k];
    float mn = (float)e->s.linear[k][GLF_min], ov = (float)e->s.linear[k][GLF_overshoot];
    int active = 0;
    if (ev->value > mn && (ev->value < ov || ov < 0)) active = 1;
    ev->row = ev->row * active + active;
    ev->abs_ += active;
    ev->last_value = ev->value;
    ev->active_sum = active;
    ev->value = 0.0f;
  }
}

/* MjClass::update_env (src/mjclass.cpp:966-1346) */
void or_update_env(OrEnv* e) {
  const double* g = e->s.general;
  const float ftol = 1e-5f;
  Forces F;
  or_extract_forces_flat(e, (double*)&F);
  if (!e->has_object) {
    /* the reference throws when no object is live (:989-993); keep counting steps only */
    e->bin[GBE_step_num].value = 1;
    update_events(e);
    return;
  }
  const double* q = e->qpos + e->B;
  /* luke::get_object_XY_relative_to_gripper: object xy minus base target xy */
  double rx = q[0] - e->base_x, ry = q[1] - e->base_y;
  float distance = (float)sqrt(rx * rx + ry * ry);
  float f1 = mag3(F.obj_f_loc[0]), f2 = mag3(F.obj_f_loc[1]), f3 = mag3(F.obj_f_loc[2]);
  float pm = mag3(F.obj_palm_loc), gm = mag3(F.obj_gnd);
  float palm_axial = (float)F.obj_palm_loc[0];
  float lift_height = (float)(q[2] - e->start_z);
  float avg_finger = (float)(0.33333 * (f1 + f2 + f3));
  float peak_lat = (float)(-1 * fmin(F.obj_f_loc[0][1], fmin(F.obj_f_loc[1][1], F.obj_f_loc[2][1])));
  float ov_avg = 0, ov_palm = 0, ov_lift = 0;          /* ObjValues blank */
  if (avg_finger > ov_avg) ov_avg = avg_finger;
  if (palm_axial > ov_palm) ov_palm = palm_axial;
  if (peak_lat > e->peak_finger_lateral_force) e->peak_finger_lateral_force = peak_lat;
  if (lift_height > ov_lift) ov_lift = lift_height;
  float gripper_z_height = (float)(-1 * e->base_z);
  float peak_axial = (float)(-1 * fmin(F.gnd_f_loc[0][0], fmin(F.gnd_f_loc[1][0], F.gnd_f_loc[2][0])));
  float g1 = or_win_read(&e->si_gauge[0], 0), g2 = or_win_read(&e->si_gauge[1], 0), g3 = or_win_read(&e->si_gauge[2], 0);
  float last_palm = or_win_read(&e->si_palm, 0), last_wrist = or_win_read(&e->si_wristz, 0);
  float max_gauge = fmaxf(fmaxf(g1, g2), g3);
  float avg_gauge = (float)((1.0 / 3.0) * (g1 + g2 + g3));

  OrEvent* b = e->bin; OrEvent* l = e->lin;
  b[GBE_step_num].value = 1;
  int lifted = 0, oob = 0, l2h = 0, th = 0, stable = 0;
  /* quirk Q4: obj.peak_finger_axial_force is never written (0) */
  if (gm < ftol && 0.0f < ftol) { b[GBE_lifted].value = 1; lifted = 1; }
  double od = g[GS_oob_distance];
  if (q[0] > od || q[0] < -od || q[1] > od || q[1] < -od) { b[GBE_oob].value = 1; oob = 1; }
  if (ov_lift > g[GS_lift_height] - ftol && lifted && !oob) { b[GBE_lifted_to_height].value = 1; l2h = 1; }
  if (l2h && gripper_z_height > g[GS_gripper_target_height] - ftol) { b[GBE_target_height].value = 1; th = 1; }
  if (f1 > ftol || f2 > ftol || f3 > ftol || pm > ftol) b[GBE_object_contact].value = 1;
  if (f1 > g[GS_stable_finger_force] && f2 > g[GS_stable_finger_force] && f3 > g[GS_stable_finger_force] &&
      f1 < g[GS_stable_finger_force_lim] && f2 < g[GS_stable_finger_force_lim] && f3 < g[GS_stable_finger_force_lim] &&
      pm > g[GS_stable_palm_force] && pm < g[GS_stable_palm_force_lim] && b[GBE_lifted].value) {
    b[GBE_object_stable].value = 1; stable = 1;
  }
  if (stable && th) b[GBE_stable_height].value = 1;
  if (e->termination_signal_sent) {
    if (e->s.binary[GBE_lifted_termination][GBF_done]) {
      if (l2h) b[GBE_lifted_termination].value = 1; else b[GBE_failed_termination].value = 1;
    } else {
      if (b[GBE_stable_height].value) b[GBE_stable_termination].value = 1; else b[GBE_failed_termination].value = 1;
    }
  }
  if (distance < g[GS_XY_distance_threshold]) b[GBE_within_XY_distance].value = 1;
  double closest = distance;
  /* dropped (src/mjclass.cpp:1249-1255, quirk Q7); value is a bool */
  {
    int dr = b[GBE_dropped].row, lv = b[GBE_lifted].value != 0, lr = b[GBE_lifted].row;
    int v = ((!dr) * (!lv) * lr) ? 1 : (lv ? 0 : (dr ? dr + 1 : 0));
    b[GBE_dropped].value = v != 0;
  }
  l[GLE_exceed_axial].value = peak_axial;
  l[GLE_exceed_lateral].value = e->peak_finger_lateral_force;
  l[GLE_palm_force].value = ov_palm * (b[GBE_lifted].value != 0);
  l[GLE_exceed_palm].value = ov_palm;
  l[GLE_finger_force].value = ov_avg;
  l[GLE_finger1_force].value = f1; l[GLE_finger2_force].value = f2; l[GLE_finger3_force].value = f3;
  l[GLE_ground_force].value = gm;
  l[GLE_good_bend_sensor].value = avg_gauge;
  l[GLE_exceed_bend_sensor].value = max_gauge; l[GLE_dangerous_bend_sensor].value = max_gauge;
  l[GLE_good_palm_sensor].value = last_palm; l[GLE_exceed_palm_sensor].value = last_palm;
  l[GLE_dangerous_palm_sensor].value = last_palm;
  l[GLE_exceed_wrist_sensor].value = last_wrist; l[GLE_dangerous_wrist_sensor].value = last_wrist;
  float denom = (float)(e->n_actions - (int)g[GS_use_termination_action]);
  l[GLE_action_penalty_lin].value /= denom;
  l[GLE_action_penalty_sq].value /= denom;
  l[GLE_object_XY_distance].value = (float)(-closest);
  /* successful_grasp metric (:1298-1316) */
  for (int k = 0; k < GBE_COUNT; k++) {
    if (b[k].value != 0 && (float)e->s.binary[k][GBF_reward] >= (1.0 - 1e-5) && e->s.binary[k][GBF_done] &&
        b[k].row + 1 >= (int)e->s.binary[k][GBF_trigger])
      b[GBE_successful_grasp].value = 1;
  }
  update_events(e);
}

/* calc_rewards (src/mjclass.cpp:5471-5528) + MjClass::reward (:3000-3049) */
float or_reward(OrEnv* e) {
  float reward = 0;
  for (int k = 0; k < GBE_COUNT; k++)
    if (e->bin[k].row >= (int)e->s.binary[k][GBF_trigger]) reward += (float)e->s.binary[k][GBF_reward];
  for (int k = 0; k < GLE_COUNT; k++)
    if (e->lin[k].row >= (int)e->s.linear[k][GLF_trigger]) {
      float fr = or_linear_reward(e->lin[k].last_value, (float)e->s.linear[k][GLF_min],
                                  (float)e->s.linear[k][GLF_max], (float)e->s.linear[k][GLF_overshoot]);
      float sc = (float)e->s.linear[k][GLF_reward] * fr;
      reward += sc;
    }
  float lo = (float)e->s.general[GS_reward_cap_lower_bound], hi = (float)e->s.general[GS_reward_cap_upper_bound];
  e->cumulative_reward += reward;
  if (e->cumulative_reward < lo && e->s.general[GS_cap_reward]) {
    reward += lo - e->cumulative_reward; e->cumulative_reward = lo;
  }
  if (e->cumulative_reward > hi && e->s.general[GS_cap_reward]) {
    reward += hi - e->cumulative_reward; e->cumulative_reward = hi;
  }
  return reward;
}

/* MjClass::is_done (src/mjclass.cpp:1632-1698) */
int or_is_done(OrEnv* e) {
  const float ftol = 1e-5f;
  if (e->unstable) return 1;   /* BADQACC analogue: the env must be reset */
  for (int k = 0; k < GBE_COUNT; k++) {
    int done = (int)e->s.binary[k][GBF_done];
    if (done && e->bin[k].row >= done) return 1;
  }
  for (int k = 0; k < GLE_COUNT; k++) {
    int done = (int)e->s.linear[k][GLF_done];
    if (done && e->lin[k].row >= done) return 1;
  }
  if (e->s.general[GS_cap_reward] && e->s.general[GS_quit_if_cap_exceeded]) {
    if (e->cumulative_reward - ftol < (float)e->s.general[GS_reward_cap_lower_bound]) return 1;
    if (e->cumulative_reward + ftol > (float)e->s.general[GS_reward_cap_upper_bound]) return 1;
  }
  return 0;
}

/* EventTrack::vectorise (src/mjclass.cpp:5594-5615) */
void or_event_state(OrEnv* e, float* out) {
  int n = 0;
  for (int k = 0; k < GBE_COUNT; k++) out[n++] = (float)e->bin[k].row;
  for (int k = 0; k < GLE_COUNT; k++) { out[n++] = (float)e->lin[k].row; out[n++] = e->lin[k].last_value; }
}

/* MjClass::get_observation (src/mjclass.cpp:1700-1959); stream order fixed, quirk Q2 noted:
 * wrist-Z is sampled with the wrist_sensor_XY Sensor object's prev_steps/readings_per_step */
int or_get_observation(OrEnv* e, float* obs) {
  const GmjSettings* s = &e->s;
  int n = 0;
  int smode = (int)s->general[GS_sensor_sample_mode], tmode = (int)s->general[GS_state_sample_mode];
  if (tmode == GSM_scaled_change_sq) { smode = GSM_scaled_change_sq; tmode = GSM_sign; }  /* quirk Q2: :204-207 assigns sampleFcnPtr */
#define SAMP(SEN, MODE, WIN) n += or_sample(MODE, e->prev_steps[SEN], e->readings_per_step[SEN], WIN, obs + n)
  if (s->sensor[GSEN_bending_gauge][GSF_in_use]) for (int k = 0; k < 3; k++) SAMP(GSEN_bending_gauge, smode, &e->w_gauge[k]);
  if (s->sensor[GSEN_axial_gauge][GSF_in_use]) for (int k = 0; k < 3; k++) SAMP(GSEN_axial_gauge, smode, &e->w_axial[k]);
  if (s->sensor[GSEN_palm][GSF_in_use]) SAMP(GSEN_palm, smode, &e->w_palm);
  if (s->sensor[GSEN_wrist_XY][GSF_in_use]) { OrWindow z; or_win_reset(&z); SAMP(GSEN_wrist_XY, smode, &z); SAMP(GSEN_wrist_XY, smode, &z); }
  if (s->sensor[GSEN_wrist_Z][GSF_in_use]) SAMP(GSEN_wrist_XY, smode, &e->w_wristz);
  if (s->sensor[GSEN_motor_state][GSF_in_use]) for (int k = 0; k < 3; k++) SAMP(GSEN_motor_state, tmode, &e->w_motor[k]);
  if (s->sensor[GSEN_base_XY][GSF_in_use]) for (int k = 0; k < 2; k++) SAMP(GSEN_base_XY, tmode, &e->w_basexy[k]);
  if (s->sensor[GSEN_base_Z][GSF_in_use]) SAMP(GSEN_base_Z, tmode, &e->w_basez);
  if (s->sensor[GSEN_base_yaw][GSF_in_use]) SAMP(GSEN_base_yaw, tmode, &e->w_baseyaw);
  if (s->sensor[GSEN_cartesian][GSF_in_use]) for (int k = 0; k < 12; k++) SAMP(GSEN_cartesian, GSM_change, &e->w_cart[k]);
#undef SAMP
  return n;
}
