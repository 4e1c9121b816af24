/* kin.c -- oracle: motor-space kinematics, sliding windows, samplers, reward helpers, Box2d.
 * TEST INFRASTRUCTURE ONLY (see oracle.h). Every function cites the reference lines it restates.
 */
#include <math.h>
#include <string.h>
#include "oracle.h"

/* ---- constants of luke::Gripper (src/gripper.h:29-65) */
#define G_LEADSCREW 35e-3
#define G_XY_MIN 49e-3
#define G_XY_MAX 134e-3
#define G_Z_MIN 0.0
#define G_Z_MAX 165e-3
#define G_LIMIT_TOL 1e-4
#define G_XY_STEP (4.0 / (1.0 * 400 * 1e3))
#define G_Z_STEP (4.8768 / (1 * 400 * 1e3))
#define G_XY_HOME (G_XY_MAX - 1.0 * (4 * 1e-3 / 1.0))
#define G_Z_HOME (G_Z_MIN + 1.0 * (4.8768 * 1e-3 / 1))
#define G_FINGER_LENGTH 235e-3
#define G_HOOK_LENGTH 35.0e-3

static const double g_to_rad = M_PI / 180.0;

static double calc_th(double x, double y) { return asin((y - x) / G_LEADSCREW); } /* gripper.h:84 (sign=+1) */
static double calc_y(double x, double th) { return x + G_LEADSCREW * sin(th); }    /* gripper.h:83 */

double or_gripper_th(const OrGripper* g) { return calc_th(g->x, g->y); }           /* get_th_rad, gripper.h:102 */

/* Gripper::update_xy (src/gripper.cpp:6-90) */
int or_gripper_update_xy(OrGripper* g) {
  int wl = 1;
  const double th_min = -40 * g_to_rad, th_max = 40 * g_to_rad, fingertip_radius_min = -1;
  if (g->x > G_XY_MAX + G_LIMIT_TOL) { g->x = G_XY_MAX; wl = 0; }
  if (g->x < G_XY_MIN - G_LIMIT_TOL) { g->x = G_XY_MIN; wl = 0; }
  if (g->y > G_XY_MAX + G_LIMIT_TOL) { g->y = G_XY_MAX; wl = 0; }
  if (g->y < G_XY_MIN - G_LIMIT_TOL) { g->y = G_XY_MIN; wl = 0; }
  double new_th = calc_th(g->x, g->y);
  if (new_th > th_max) { new_th = th_max; g->y = calc_y(g->x, th_max); wl = 0; }
  if (new_th < th_min) { new_th = th_min; g->y = calc_y(g->x, th_min); wl = 0; }
  g->th = new_th;
  /* fingertip radius cap (gripper.h:88-94); disabled by fingertip_radius_min = -1 but evaluated */
  const double hyp = sqrt(G_FINGER_LENGTH * G_FINGER_LENGTH + G_HOOK_LENGTH * G_HOOK_LENGTH);
  const double rest = atan(G_HOOK_LENGTH / G_FINGER_LENGTH);
  double th_lim = asin((fingertip_radius_min - g->x) / hyp) + rest;
  if (g->th < th_lim) { g->y = calc_y(g->x, th_lim); g->th = th_lim; wl = 0; }
  g->sx = (int)round((G_XY_MAX - g->x) / G_XY_STEP);
  g->sy = (int)round((G_XY_MAX - g->y) / G_XY_STEP);
  return wl;
}

/* Gripper::update_z (src/gripper.cpp:92-118) */
int or_gripper_update_z(OrGripper* g) {
  int wl = 1;
  if (g->z > G_Z_MAX + G_LIMIT_TOL) { g->z = G_Z_MAX; wl = 0; }
  if (g->z < G_Z_MIN - G_LIMIT_TOL) { g->z = G_Z_MIN; wl = 0; }
  g->sz = (int)round(g->z / G_Z_STEP);
  return wl;
}

/* Gripper::update (src/gripper.cpp:120-131) */
static int gripper_update(OrGripper* g) {
  int a = or_gripper_update_xy(g);
  int b = or_gripper_update_z(g);
  return a && b;
}

void or_gripper_reset(OrGripper* g) {            /* gripper.h:78 */
  g->x = G_XY_HOME; g->y = G_XY_HOME; g->z = G_Z_HOME;
  gripper_update(g);
}

int or_gripper_set_xyz_m(OrGripper* g, double x, double y, double z) {   /* gripper.h:151-153 */
  g->x = x; g->y = y; g->z = z;
  return gripper_update(g);
}

int or_gripper_set_xyz_m_rad(OrGripper* g, double x, double th, double z) { /* gripper.h:158-161 */
  g->x = x;
  g->y = calc_y(g->x, th);
  int in_lim = or_gripper_update_xy(g);
  g->z = z;
  return or_gripper_update_z(g) ? in_lim : 0;
}

static int set_xyz_step(OrGripper* g, int xs, int ys, int zs) {          /* gripper.h:162-165 */
  g->x = G_XY_MAX - G_XY_STEP * xs; g->y = G_XY_MAX - G_XY_STEP * ys;
  int in_lim = or_gripper_update_xy(g);
  g->z = G_Z_STEP * zs;
  return or_gripper_update_z(g) ? in_lim : 0;
}

/* Gripper::step_to (src/gripper.cpp:158-228) */
int or_gripper_step_to(OrGripper* g, const OrGripper* t, int num) {
  int finished = 1;
  int d[3] = { t->sx - g->sx, t->sy - g->sy, t->sz - g->sz };
  for (int k = 0; k < 3; k++) {
    if (d[k] < 0) { if (-d[k] > num) { d[k] = -num; finished = 0; } }
    else { if (d[k] > num) { d[k] = num; finished = 0; } }
  }
  set_xyz_step(g, g->sx + d[0], g->sy + d[1], g->sz + d[2]);
  return finished;
}

/* ---- luke::SlidingWindow<float> (src/slidingwindow.h:9-55); depth OR_WIN instead of 1000:
 * the hot path never reads further back than total_readings-1 = 6. */
void or_win_reset(OrWindow* w) { memset(w->v, 0, sizeof(w->v)); w->i = -1; }
void or_win_add(OrWindow* w, float v) {
  w->i += 1;
  if (w->i > OR_WIN - 1) w->i = 0;
  w->v[w->i] = v;
}
float or_win_read(const OrWindow* w, int n) {
  int idx = w->i - n;
  while (idx < 0) idx += OR_WIN;
  return w->v[idx];
}

/* ---- MjType::Sensor samplers (src/mjclass.h:278-453). Returns the output length. */
int or_sample(int mode, int p, int r, const OrWindow* w, float* out) {
  int T = 1 + r * p;
  if (mode == GSM_raw) {                                 /* raw_sample: data.read(T-1) */
    int n = T - 1;
    for (int j = n - 1, k = 0; j >= 0; j--, k++) out[k] = or_win_read(w, j);
    return n;
  }
  out[0] = or_win_read(w, T - 1);
  for (int i = 0; i < p; i++) {
    int first = T - 1 - i * r;
    out[2 * i + 2] = or_win_read(w, first - r);
    float a = out[2 * i], b = out[2 * i + 2], v = 0.0f;
    switch (mode) {
      case GSM_change: v = b - a; break;
      case GSM_average: {
        v = 0;
        for (int j = 0; j < r + 1; j++) v += or_win_read(w, first - j);
        v /= r + 1;
      } break;
      case GSM_median: {
        float vals[OR_WIN]; int n = r + 1;
        for (int j = 0; j < n; j++) vals[j] = or_win_read(w, first - j);
        for (int x = 1; x < n; x++) { float t = vals[x]; int y = x - 1;
          while (y >= 0 && vals[y] > t) { vals[y + 1] = vals[y]; y--; } vals[y + 1] = t; }
        int h = n / 2;
        v = vals[h];
        if (!(n & 1)) v = (float)((vals[h - 1] + v) / 2.0);
      } break;
      case GSM_sign: {
        const float tol = 1e-6f; float c = b - a;
        v = (c > tol) ? 1.0f : ((c < -tol) ? -1.0f : 0.0f);
      } break;
      case GSM_scaled_change: {
        v = (b - a) / 0.07f;
        if (v > 1.0) v = 1.0f; else if (v < -1.0) v = -1.0f;
      } break;
      case GSM_scaled_change_sq: {
        const float m2 = 0.10f; const float fac = (float)(1.0 / (m2 * m2));
        float c = fabsf(b - a);
        v = (b - a) * c * fac;
        if (v > 1.0) v = 1.0f; else if (v < -1.0) v = -1.0f;
      } break;
      default: v = 0; break;
    }
    out[2 * i + 1] = v;
  }
  return 2 * p + 1;
}

/* linear_reward (src/mjclass.cpp:4866-4894) */
float or_linear_reward(float val, float mn, float mx, float overshoot) {
  if (val < mn) return 0.0f;
  if (val > mx) {
    if (overshoot < mx) return 1.0f;
    if (val > overshoot) return 0.0f;
    mn = 0; mx = overshoot - mx; val = overshoot - val;
  }
  return (val - mn) / (mx - mn);
}

/* normalise_between (src/mjclass.cpp:4896-4904) */
float or_normalise_between(float val, float mn, float mx) {
  if (val > mx) return 1.0f;
  else if (val < mn) return -1.0f;
  return 2 * (val - mn) / (mx - mn) - 1;
}

/* ---- luke::Box2d (src/customtypes.h:35-172); box = x0..x3,y0..y3 */
void or_box2d_init(double b[8], double cx, double cy, double w, double h, double theta) {
  double hw = w / 2.0, hh = h / 2.0;
  b[0] = cx - hw; b[4] = cy - hh; b[1] = cx + hw; b[5] = cy - hh;
  b[2] = cx + hw; b[6] = cy + hh; b[3] = cx - hw; b[7] = cy + hh;
  double mx = (b[0] + b[1] + b[2] + b[3]) / 4.0, my = (b[4] + b[5] + b[6] + b[7]) / 4.0;
  for (int i = 0; i < 4; i++) {
    double nx = mx + (b[i] - mx) * cos(theta) - (b[4 + i] - my) * sin(theta);
    double ny = my + (b[i] - mx) * sin(theta) + (b[4 + i] - my) * cos(theta);
    b[i] = nx; b[4 + i] = ny;
  }
}

int or_box2d_overlap(const double a[8], const double o[8], double min_distance) {
  int contains_other = 1;
  for (int i = 0; i < 4; i++) {
    int j = (i + 1) % 4;
    double ex = a[j] - a[i], ey = a[4 + j] - a[4 + i];
    double px = -ey, py = ex;
    double len = sqrt(px * px + py * py);
    px /= len; py /= len;
    double min1 = a[0] * px + a[4] * py, max1 = min1;
    double min2 = o[0] * px + o[4] * py, max2 = min2;
    for (int k = 1; k < 4; k++) {
      double p1 = a[k] * px + a[4 + k] * py, p2 = o[k] * px + o[4 + k] * py;
      if (p1 < min1) min1 = p1;
      if (p1 > max1) max1 = p1;
      if (p2 < min2) min2 = p2;
      if (p2 > max2) max2 = p2;
    }
    if (max1 + min_distance < min2 || max2 + min_distance < min1) return 0;
    if (max1 < min2 || max2 < min1) contains_other = 0;
  }
  return contains_other;
}
