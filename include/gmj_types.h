/* gmj_types.h -- plain-C data layout shared by the host (Python/ctypes), the CUDA
 * library (libgripper_b200.so) and the CPU oracle (test infrastructure only).
 *
 * Everything here is data, no behaviour. Citations are path:line in the
 * reference tree (lukebeddow/gripper-mujoco).
 *
 * Dof / body order of the gripper (reference keyframe order,
 * description/mujoco/xml_script.py:216-223):
 *   0                      base z slide (world_to_base)
 *   1 + f*(N+2)            finger f prismatic      (f = 0,1,2)
 *   2 + f*(N+2)            finger f revolute
 *   2 + f*(N+2) + k        finger f segment joint k (k = 1..N)
 *   3N + 7                 palm prismatic
 * followed by the single live object's free joint (6 dofs / 7 qpos).
 */
#ifndef GMJ_TYPES_H_
#define GMJ_TYPES_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GMJ_MAXB 88   /* max gripper bodies = dofs (N=24 -> 80)            */
#define GMJ_MAXG 96   /* max gripper collision geoms (3N segs + 3 hooks + palm) */

/* geom types */
enum { GMJ_GEOM_ELLIPSOID = 0, GMJ_GEOM_BOX = 1, GMJ_GEOM_CYLINDER = 2 };
/* geom classes for the force tally (src/objecthandler.h:110-129) */
enum { GMJ_CLS_NONE = 0, GMJ_CLS_F1 = 1, GMJ_CLS_F2 = 2, GMJ_CLS_F3 = 3,
       GMJ_CLS_PALM = 4, GMJ_CLS_GROUND = 5, GMJ_CLS_OBJECT = 6 };

typedef struct GmjModel {
  int32_t N;                 /* segments per finger                              */
  int32_t nb;                /* base dofs (1 supported: world_to_base)            */
  int32_t B;                 /* gripper bodies = gripper dofs = nb + 7 + 3N       */
  int32_t ngeom;             /* gripper collision geoms                           */
  int32_t parent[GMJ_MAXB];  /* parent body, -1 = world                           */
  int32_t jtype[GMJ_MAXB];   /* 0 = slide, 1 = hinge                              */
  int32_t group[GMJ_MAXB];   /* 0 base, 1..3 finger, 4 palm                       */
  double pos[GMJ_MAXB][3];   /* body frame origin in parent frame                 */
  double quat[GMJ_MAXB][4];  /* body frame orientation in parent frame (w,x,y,z)  */
  double axis[GMJ_MAXB][3];  /* joint axis in body frame                          */
  double mass[GMJ_MAXB];
  double ipos[GMJ_MAXB][3];  /* COM in body frame                                 */
  double inertia[GMJ_MAXB][6]; /* xx yy zz xy xz yz about COM, body frame         */
  double stiffness[GMJ_MAXB];
  double damping[GMJ_MAXB];
  double range_lo[GMJ_MAXB];
  double range_hi[GMJ_MAXB];
  double qpos0[GMJ_MAXB];    /* keyframe "initial pose"                           */
  /* gripper collision geoms */
  int32_t g_body[GMJ_MAXG];
  int32_t g_type[GMJ_MAXG];
  int32_t g_class[GMJ_MAXG];
  double g_pos[GMJ_MAXG][3];
  double g_quat[GMJ_MAXG][4];
  double g_size[GMJ_MAXG][3]; /* box: half extents; cylinder: (r, r, half length along local z) */
  double g_friction[GMJ_MAXG];
  double ground_friction;
  /* MuJoCo option defaults (description/mujoco/xacro/mujoco.xacro:6-12 sets none) */
  double gravity;            /* 9.81, along -z                                    */
  double solref[2];          /* 0.02, 1                                           */
  double solimp[5];          /* 0.9 0.95 0.001 0.5 2                              */
  /* finger dimensions (description/config/gripper.yaml:13-22) */
  double finger_length, finger_width, finger_thickness, finger_E;
  double hook_length, hook_angle_deg, fingertip_clearance, segment_length;
  double EI;
  /* PD gains: gripper (x, th, z) and base z (src/myfunctions.cpp:273-296) */
  double kp[3], kd[3], base_kp, base_kd;
  /* tip bodies (last segment of each finger) and the initial segment matrices used
   * by the calibration tip load (src/myfunctions.cpp:1642-1727) */
  int32_t tip_body[3];
  int32_t finger_body[3];    /* body "finger_i" (local force frames)              */
  int32_t palm_body;
  int32_t pad_;
  double tip_mat[3][9];
} GmjModel;

/* one graspable object (description/mujoco/object_sets/build_object_set.py:255-343) */
typedef struct GmjObject {
  int32_t type;              /* GMJ_GEOM_*                                        */
  int32_t pad_;
  double size[3];            /* inner shape: box half extents / (r, r, half h) / radii */
  double round;              /* fillet radius (sphere swept around the inner shape) */
  double quat[4];            /* shape frame in body frame (mesh refquat)          */
  double mass;
  double inertia[3];         /* diagonal in body frame                            */
  double friction;           /* sliding friction of the object geom               */
  double z_rest;             /* keyframe z (rest height + 1e-4)                   */
  double bbox[3];            /* "Task object j" numeric: x y z sizes              */
} GmjObject;

/* ---- settings: flat arrays, index enums follow src/simsettings.h order ---- */
enum {
  GS_mujoco_timestep = 0, GS_curve_validation, GS_tip_force_applied, GS_random_seed,
  GS_time_for_action, GS_saturation_yield_factor,
  GS_sensor_sample_mode, GS_state_sample_mode, GS_sensor_n_prev_steps, GS_state_n_prev_steps,
  GS_sensor_noise_mag, GS_sensor_noise_mu, GS_sensor_noise_std,
  GS_state_noise_mag, GS_state_noise_mu, GS_state_noise_std,
  GS_base_position_noise, GS_palm_scale_factor,
  GS_oob_distance, GS_lift_height, GS_gripper_target_height,
  GS_stable_finger_force, GS_stable_palm_force, GS_stable_finger_force_lim, GS_stable_palm_force_lim,
  GS_XY_distance_threshold,
  GS_cap_reward, GS_quit_if_cap_exceeded, GS_reward_cap_lower_bound, GS_reward_cap_upper_bound,
  GS_continous_actions, GS_use_termination_action, GS_termination_threshold, GS_lift_on_termination,
  GS_sim_steps_per_action, GS_fingertip_min_mm,
  /* derived / calibration scalars (src/mjclass.cpp:274, 4659-4667) */
  GS_sim_gauge_raw_to_N_factor,
  /* base limits (src/myfunctions.cpp:226-241) */
  GS_base_z_min, GS_base_z_max, GS_base_x_min, GS_base_x_max, GS_base_y_min, GS_base_y_max,
  GS_base_yaw_min, GS_base_yaw_max,
  GS_reset_qpos_base,        /* calibrate_reset equilibrium of the base joint      */
  GS_reset_qpos_g0, GS_reset_qpos_g1, GS_reset_qpos_g2, GS_reset_qpos_g3, GS_reset_qpos_g4,
  GS_reset_qpos_g5, GS_reset_qpos_g6, /* ... and of the 7 motor joints (src/myfunctions.cpp:1476-1495) */
  GS_use_noise_seeded,       /* 1: counter-based noise enabled on device            */
  GS_COUNT_ = 56
};
/* sensors in X-macro order (src/simsettings.h:106-116) */
enum { GSEN_motor_state = 0, GSEN_base_Z, GSEN_base_XY, GSEN_base_yaw, GSEN_bending_gauge,
       GSEN_axial_gauge, GSEN_palm, GSEN_wrist_XY, GSEN_wrist_Z, GSEN_cartesian, GSEN_COUNT = 10 };
/* per-sensor fields */
enum { GSF_in_use = 0, GSF_normalise, GSF_read_rate, GSF_use_normalisation, GSF_use_noise,
       GSF_raw_value_offset, GSF_noise_mag, GSF_noise_mu, GSF_noise_std, GSF_noise_overriden,
       GSF_COUNT = 10 };
/* actions in X-macro order (src/simsettings.h:119-129) */
enum { GACT_gripper_X = 0, GACT_gripper_prismatic_X, GACT_gripper_Y, GACT_gripper_revolute_Y,
       GACT_gripper_Z, GACT_base_X, GACT_base_Y, GACT_base_Z, GACT_base_roll, GACT_base_pitch,
       GACT_base_yaw, GACT_COUNT = 11 };
enum { GAF_in_use = 0, GAF_continous, GAF_value, GAF_sign, GAF_COUNT = 4 };
/* binary events (src/simsettings.h:138-152) */
enum { GBE_step_num = 0, GBE_lifted, GBE_oob, GBE_dropped, GBE_lifted_to_height, GBE_target_height,
       GBE_exceed_limits, GBE_object_contact, GBE_object_stable, GBE_stable_height,
       GBE_stable_termination, GBE_lifted_termination, GBE_failed_termination,
       GBE_successful_grasp, GBE_within_XY_distance, GBE_COUNT = 15 };
enum { GBF_reward = 0, GBF_done, GBF_trigger, GBF_COUNT = 3 };
/* linear events (src/simsettings.h:159-183) */
enum { GLE_finger_force = 0, GLE_palm_force, GLE_exceed_axial, GLE_exceed_lateral, GLE_exceed_palm,
       GLE_good_bend_sensor, GLE_good_palm_sensor, GLE_exceed_bend_sensor, GLE_exceed_wrist_sensor,
       GLE_exceed_palm_sensor, GLE_dangerous_bend_sensor, GLE_dangerous_wrist_sensor,
       GLE_dangerous_palm_sensor, GLE_action_penalty_lin, GLE_action_penalty_sq,
       GLE_finger1_force, GLE_finger2_force, GLE_finger3_force, GLE_ground_force, GLE_grasp_metric,
       GLE_object_XY_distance, GLE_COUNT = 21 };
enum { GLF_reward = 0, GLF_done, GLF_trigger, GLF_min, GLF_max, GLF_overshoot, GLF_COUNT = 6 };

typedef struct GmjSettings {
  double general[GS_COUNT_];
  double sensor[GSEN_COUNT][GSF_COUNT];
  double action[GACT_COUNT][GAF_COUNT];
  double binary[GBE_COUNT][GBF_COUNT];
  double linear[GLE_COUNT][GLF_COUNT];
} GmjSettings;

/* sample modes (src/mjclass.h MjType::Sample) */
enum { GSM_raw = 0, GSM_change, GSM_average, GSM_median, GSM_sign, GSM_scaled_change,
       GSM_scaled_change_sq };

#define GMJ_EVENT_STATE_LEN 57   /* 15 + 2*21, src/mjclass.cpp:5594-5615 */
#define GMJ_MAX_OBS 160          /* 19+ streams * (2*prev_steps+1) upper bound */
#define GMJ_MAXROW 48           /* constraint rows kept per env per substep; a contact whose 4 rows
                                    do not fit is dropped together with all later contacts */
#define GMJ_MAXCON 10            /* contacts kept per env per substep */

/* per-env parity dump (qpos/qvel in reference order, sensors, contact stats) */
typedef struct GmjStateView {
  int32_t nq, nv, nwin;
  int32_t pad_;
} GmjStateView;

/* MjClass::configure_settings' auto-calibration (src/mjclass.cpp:225-314) as one call: which parts to run (the
 * reference's auto_* settings) and what came out. The results are also written into the GmjSettings passed along. */
typedef struct GmjCalib {
  int32_t set_timestep;        /* in: auto_set_timestep     -> find_highest_stable_timestep (src/mjclass.cpp:4745-4850) */
  int32_t calibrate_gauges;    /* in: auto_calibrate_gauges -> calibrate_simulated_sensors  (src/mjclass.cpp:4643-4676) */
  int32_t set_sim_steps;       /* in: auto_sim_steps        -> sim_steps_per_action = ceil(time_for_action / timestep)   */
  int32_t sim_steps_per_action;      /* out */
  double highest_stable_timestep;    /* out: before the safety factor (0 when set_timestep was off)                      */
  double timestep;                   /* out: mujoco_timestep in force afterwards                                         */
  double wrist_raw_value_offset;     /* out: wrist_sensor_Z.raw_value_offset                                             */
  double bend_gauge_normalise;       /* out: bending_gauge.normalise (gauge reading under the saturation tip load)       */
  double sim_gauge_raw_to_N_factor;  /* out */
  double reset_qpos[8];              /* out: calibrate_reset equilibrium, base + 7 motor joints (src/myfunctions.cpp:1450-1516) */
  double yield_load;                 /* out: luke::calc_yield_point_load (src/myfunctions.cpp:3587-3606)                 */
  int32_t launches, pad_;            /* out: scratch simulations run                                                      */
} GmjCalib;

typedef struct GmjStats {
  int64_t substeps;          /* substeps executed since create                    */
  int64_t unstable_resets;   /* BADQACC-equivalent events (src/myfunctions.cpp:4233) */
  int64_t contacts;          /* sum of ncon over substeps                         */
  int64_t rows;              /* sum of constraint rows over substeps              */
  int64_t solver_iters;      /* sum of active-set pivoting rounds                 */
  int64_t contact_overflow;  /* contacts dropped because > GMJ_MAXCON             */
  int64_t relaxed_solves;    /* constraint solves that needed 3+ pivoting rounds: accepted under the widened feasibility tolerance */
  int64_t capped_solves;     /* constraint solves that hit the round cap with an infeasible active set    */
} GmjStats;

#ifdef __cplusplus
}
#endif
#endif
