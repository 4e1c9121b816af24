/* oracle.h -- CPU oracle: single-environment restatement of the reference hot path.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under gripper_mujoco_b200/ may include, link or call this;
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * PARITY STATUS
 *  - Motor kinematics, limits, stepper, sliding windows: PINNED against the reference's own src/gripper.cpp and
 *    src/slidingwindow.h compiled verbatim into oracle/_ref/libgripper_ref.so (tests/test_oracle_golden.py).
 *  - The whole non-physics tail (set_action, the action_step loop's sensor sampling with noise, update_env, events,
 *    reward, is_done, get_observation in every sample mode, spawn_into_scene's search and Box2d SAT): PINNED against
 *    the reference's own src/mjclass.cpp compiled in place over stub MuJoCo headers and driven by scripted physics
 *    observables (oracle/ref_tail_shim.cpp, oracle/_ref/libmjclass_ref.so); or_set_script() runs this restatement on
 *    the same script and tests/test_ref_tail.py compares with tests/golden/tail_golden.npz bit for bit.
 *  - Physics (mj_step1 / mj_step2): the arithmetic lives in MuJoCo 2.1.5 (libmujoco, not vendored, not installable
 *    here). It is restated from MuJoCo's published algorithm for this scene class. PARITY UNPINNED for physics: no
 *    MuJoCo trajectories exist in this container (route to a pin: tools/mujoco_pin.py).
 *  - or_set_fast(): timed-baseline mode for bench.py's CPU arms (direct active-set solve, skyline Cholesky); equals
 *    the checker mode to 1e-16 (tests/test_host_logic.py).
 *
 * double = mjtNum, float = luke::gfloat (src/customtypes.h:22).
 */
#ifndef ORACLE_H_
#define ORACLE_H_

#include <stdint.h>
#include "../include/gmj_types.h"

#ifdef __cplusplus
extern "C" {
#endif

#define OR_MAXV (GMJ_MAXB + 6)
#define OR_MAXROW (6 * 7 + GMJ_MAXB + 4 * GMJ_MAXCON)
#define OR_SCRIPT_ROWLEN 46         /* gauges 3 | all f1-3 loc 9 | all palm loc 3 | obj f1-3 loc 9 | obj palm loc 3 | obj gnd 3 | gnd f1-3 loc 9 | wrist z | obj xyz 3 | spare 3 */
#define OR_WIN 16                 /* ring depth kept (reference keeps 1000, reads <= 7) */

/* luke::Gripper (src/gripper.h:13-110) */
typedef struct { double x, y, z, th; int sx, sy, sz; } OrGripper;

typedef struct { float v[OR_WIN]; int i; } OrWindow;   /* luke::SlidingWindow<float> */

typedef struct {
  double pos[3];        /* contact point (world) */
  double frame[9];      /* rows: normal (geom1 -> geom2), tangent1, tangent2 */
  double dist;          /* < 0 = penetration */
  double mu;
  int body1, body2;     /* -1 world, 0..B-1 gripper body, B = object */
  int cls1, cls2;       /* GMJ_CLS_* of the two geoms */
  int row;              /* first efc row */
  double force[3];      /* contact-frame force (normal, t1, t2), mj_contactForce convention */
} OrContact;

typedef struct {
  int on;
  int body1, body2;
  double relpos[3];     /* anchor of body2 origin in body1 frame */
  double relmat[9];     /* R1^T R2 at activation */
} OrWeld;

typedef struct { int row, abs_; float last_value; int active_sum; float value; } OrEvent;

typedef struct OrEnv {
  GmjModel m;
  GmjSettings s;
  GmjObject obj;
  int has_object;
  int nv, nq, B;
  /* ---- mjData-like state */
  double qpos[OR_MAXV + 1], qvel[OR_MAXV];
  double time;
  double tip_force;               /* calibration tip load on the 3 last segments (N) */
  /* derived each step */
  double xpos[GMJ_MAXB][3], xmat[GMJ_MAXB][9], xaxis[GMJ_MAXB][3], xanchor[GMJ_MAXB][3];
  double opos[3], omat[9];
  int ncon, nrow, niter;
  OrContact con[GMJ_MAXCON];
  double wrist_z;                 /* force sensor z (src/mjclass.cpp:878, quirk Q3) */
  int unstable;                   /* BADQACC analogue (src/myfunctions.cpp:4233-4242) */
  double qacc[OR_MAXV];
  double efc_force_prev[OR_MAXROW];
  int nrow_prev;
  /* ---- luke:: globals made per-env */
  OrGripper tgt_end, tgt_next;    /* target_.end / target_.next */
  double base_z, base_x, base_y, base_yaw; /* target_.base */
  double last_step_time;
  int old_x, old_y, old_z;        /* update_constraints edge state (quirk Q12) */
  OrWeld weld[7];                 /* pris1-3, rev1-3, palm (xml_script.py:332-371) */
  double reset_qpos[8];           /* calibrate_reset equilibrium: base + 7 motor joints */
  /* ---- MjClass members */
  int n_actions, action_options[32];
  double last_read_time[GSEN_COUNT];
  int prev_steps[GSEN_COUNT], readings_per_step[GSEN_COUNT], total_readings[GSEN_COUNT];
  float rand_mu[GSEN_COUNT][3];
  OrWindow w_gauge[3], w_axial[3], w_palm, w_wristz;     /* sim_sensors_  (normalised) */
  OrWindow w_motor[3], w_basexy[2], w_basez, w_baseyaw, w_cart[12];
  OrWindow si_gauge[3], si_palm, si_wristz;              /* sim_sensors_SI_ */
  OrEvent bin[GBE_COUNT], lin[GLE_COUNT];                /* env_.cnt */
  float cumulative_reward;
  int num_action_steps;
  int termination_signal_sent;
  float peak_finger_lateral_force;                       /* env_.grp (quirk Q4: starts at 0) */
  double start_z;                                        /* env_.obj[0].start_qpos.z */
  /* counter-based RNG (replaces the process-global std::default_random_engine) */
  uint64_t seed; uint32_t env_id; uint32_t rng_ctr;
  /* stats */
  int64_t st_substeps, st_contacts, st_rows, st_iters, st_unstable;
  /* scripted physics observables (tests/test_ref_tail.py): when set, or_physics_step only advances the clock and
   * loads the next row, and the tail reads gauges / force sums / wrist force / object position from it -- the same
   * rows oracle/ref_tail_shim.cpp feeds to the reference's own compiled MjClass */
  const double* script; int script_rows, script_cursor;
  double script_cur[OR_SCRIPT_ROWLEN];
  /* 1: the timed CPU baseline's solver -- straight to the active-set pivoting solve from the warm start, without the
   * Gauss-Seidel sweeps to 1e-13 the parity oracle runs first (same unique solution; bench.py's CPU arms use it so the
   * baseline is not handicapped by the checker's belt and braces) */
  int fast_solver;
  /* 1: gripper <-> gripper contacts as well (finger <-> finger, hook <-> hook, segment <-> palm; the reference's single
   * collision group, src/objecthandler.cpp:135-212, with MuJoCo's same-body and parent-child filters). OFF by default:
   * the CUDA path does not simulate them, and the oracle is its checker. Used by
   * profiles/r2_m9_effect.py to measure what the gap is worth in returns and success rates. */
  int self_collision;
} OrEnv;

/* ---- gripper motor space (src/gripper.h, src/gripper.cpp) */
void or_gripper_reset(OrGripper* g);
int or_gripper_update_xy(OrGripper* g);
int or_gripper_update_z(OrGripper* g);
int or_gripper_set_xyz_m(OrGripper* g, double x, double y, double z);
int or_gripper_set_xyz_m_rad(OrGripper* g, double x, double th, double z);
int or_gripper_step_to(OrGripper* g, const OrGripper* target, int num);
double or_gripper_th(const OrGripper* g);

/* ---- sliding window + samplers (src/slidingwindow.h, src/mjclass.h:278-453) */
void or_win_reset(OrWindow* w);
void or_win_add(OrWindow* w, float v);
float or_win_read(const OrWindow* w, int n);
int or_sample(int mode, int prev_steps, int readings_per_step, const OrWindow* w, float* out);

/* ---- reward helpers (src/mjclass.cpp:4866-4904) */
float or_linear_reward(float val, float mn, float mx, float overshoot);
float or_normalise_between(float val, float mn, float mx);

/* ---- spawn geometry (src/customtypes.h:35-172) */
int or_box2d_overlap(const double a[8], const double b[8], double min_distance);
void or_box2d_init(double box[8], double cx, double cy, double w, double h, double theta);

/* ---- env */
OrEnv* or_create(const GmjModel* m, const GmjSettings* s);
void or_destroy(OrEnv* e);
void or_set_settings(OrEnv* e, const GmjSettings* s);
void or_reset(OrEnv* e, uint64_t seed, uint32_t env_id, uint32_t episode);
void or_spawn(OrEnv* e, const GmjObject* o, double x, double y, double zrot);
int or_set_action(OrEnv* e, int action, float fraction);
void or_substep(OrEnv* e);                 /* MjClass::step */
void or_physics_step(OrEnv* e);            /* luke::before_step/step/after_step only */
void or_action_step(OrEnv* e);
int or_get_observation(OrEnv* e, float* obs);
float or_reward(OrEnv* e);
int or_is_done(OrEnv* e);
void or_event_state(OrEnv* e, float* out57);
void or_settle(OrEnv* e, int nsteps, double dt);   /* calibrate_reset first call */
float or_read_gauge(const OrEnv* e, int finger);   /* read_armadillo_gauge */
void or_forward(OrEnv* e);                 /* FK + collision only (mj_forward subset) */
void or_set_weld(OrEnv* e, int id, int on);
void or_set_self_collision(OrEnv* e, int on);
int or_struct_size(void);
int or_spawn_ok(const OrEnv* e, double x, double y, double rot, const double* bbox3);
int or_spawn_into_scene(OrEnv* e, const GmjObject* o, double x, double y, double zrot, double xrange, double yrange, double rotrange);
int or_spawn_retry(OrEnv* e, const GmjObject* o, double xrange, double rotrange, int tries, int fb_xy_mm, int fb_rot_deg);
int or_draw_object(OrEnv* e, int n_objects);
void or_set_fast(OrEnv* e, int on);
void or_set_script(OrEnv* e, const double* rows, int nrows);

#ifdef __cplusplus
}
#endif
#endif
