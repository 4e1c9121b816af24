/* gmj.h -- C ABI of libgripper_b200.so: the batched, B200-resident replacement for the
 * hot path of the reference's pybind module (rl/env/mjpy/bind.so, class MjClass, src/bind.cpp).
 *
 * Every entry point returns 0 on success, non-zero on error (gmj_last_error() gives the text).
 * All pointers named *_dev are DEVICE pointers on the handle's GPU, all pointers named *_host are
 * host pointers; the library owns every piece of simulation state, the caller owns in/out buffers.
 * No entry point synchronises the host except gmj_*_host(), gmj_stats(), gmj_set_settings() and gmj_destroy(): those
 * call cudaDeviceSynchronize() first, so they are ordered after work queued on ANY stream of the device (also streams
 * created cudaStreamNonBlocking). Every entry point leaves the caller's current CUDA device as it found it. Handles on
 * several devices may live in one process (one handle per GPU).
 *
 * Reference interface each entry replaces (path:line in lukebeddow/gripper-mujoco):
 *   gmj_create / gmj_destroy     MjClass::MjClass / load / init        src/mjclass.cpp:22-79,377-409 ; src/bind.cpp:43-52
 *   gmj_set_settings             MjClass::set (mj.set.<field> r/w)      src/bind.cpp:195-204,246-365
 *   gmj_set_objects              luke::get_objects / ObjectHandler::init src/objecthandler.cpp:22-133
 *   gmj_reset                    MjClass::reset                         src/mjclass.cpp:434-486 ; src/bind.cpp:55
 *   gmj_reset_done               MjEnv.step done / truncation -> MjEnv.reset   rl/env/MjEnv.py:616-637,2222-2263
 *   gmj_spawn                    MjClass::spawn_object(idx,x,y,rot)     src/mjclass.cpp:2323-2420 ; src/bind.cpp:88-95
 *   gmj_step                     set_action + action_step + get_observation + reward + is_done
 *                                src/mjclass.cpp:1483-1698,1700-1959,3000-3049 ; src/bind.cpp:60-66,119-123
 *   gmj_set_action               MjClass::set_action                    src/mjclass.cpp:1528-1630
 *   gmj_action_step              MjClass::action_step                   src/mjclass.cpp:1483-1508
 *   gmj_substeps                 MjClass::step (n times)                src/mjclass.cpp:504-530 ; src/bind.cpp:57
 *   gmj_observe                  get_observation / reward / is_done / get_event_state
 *   gmj_get_state_host           data->qpos / qvel reads (parity dumps)
 *   gmj_probe_host               read_armadillo_gauge / sim_sensors_SI_ latest / calibrate_reset
 *   gmj_event_track_host         MjClass::get_test_report (TestReport.cnt)  src/mjclass.cpp:3800-3809 ; src/bind.cpp:155
 *   gmj_stats                    (new) solver + stability counters, mirrors BADQACC src/myfunctions.cpp:4233
 */
#ifndef GMJ_H_
#define GMJ_H_

#include <stdint.h>
#include "gmj_types.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct GmjHandle GmjHandle;

/* precision: 0 = FP32 physics (product), 1 = FP64 physics (validation build of the same kernels) */
int gmj_create(const GmjModel* model, const GmjSettings* settings, int n_envs, int device,
               int precision, uint32_t env_id_offset, GmjHandle** out);
int gmj_destroy(GmjHandle* h);
int gmj_set_settings(GmjHandle* h, const GmjSettings* settings);
int gmj_set_objects(GmjHandle* h, const GmjObject* objects_host, int n_objects);

/* mask_dev: uint8[n_envs] (NULL = all): MjClass::reset of the selected envs. Each env keeps its own episode counter
 * (it keys the counter RNG: draw counter = episode << 12). stream: cudaStream_t as void*. */
int gmj_reset(GmjHandle* h, const uint8_t* mask_dev, void* stream);
/* MjEnv.step's end-of-episode rule on the device (rl/env/MjEnv.py:616-637, 2222-2263): reset every env whose done flag
 * (done_dev uint8[n_envs], as written by gmj_step) is set OR whose episode has reached step_cap action steps
 * (truncation); mask_out_dev uint8[n_envs] (may be NULL) receives which envs were reset, for the gmj_spawn that follows. */
int gmj_reset_done(GmjHandle* h, const uint8_t* done_dev, int step_cap, uint8_t* mask_out_dev, void* stream);
/* idx_dev int32[n_envs], xyr_dev float[n_envs][3] (x, y, z-rotation): spawn_object, placed unconditionally.
 * xyr_dev == NULL: spawn_into_scene -- the reference's search (src/mjclass.cpp:2475-2654): candidate (x, y) on a 2 mm
 * grid within +-xy_range and rotations on a pi/30 grid within +-rot_range, both lists shuffled (counter RNG) and walked
 * pairwise; the first pose whose footprint clears the three fingertip boxes by 1 mm spawns. idx_dev == NULL draws the
 * object per env as well. ok_dev uint8[n_envs] or NULL: 1 where an object was placed, 0 where no candidate was free
 * (spawn_into_scene's `false`: the env has no object, the caller retries / falls back as rl/env/MjEnv.py:1216-1244).
 * mask_dev as above; ok_dev is written for masked-in envs only.
 * tries >= 1: that many spawn_into_scene attempts per env (fresh shuffles each), the loop of MjEnv._spawn_object.
 * fallback_xy_mm >= 0: an env still without a free pose gets MjEnv's "old method" -- a pose on the millimetre grid
 * within +-fallback_xy_mm with a base angle of 0 / 60 / 120 degrees +-fallback_rot_deg, no collision test (ok = 2);
 * fallback_xy_mm < 0: no fallback (ok = 0, the env has no object). */
int gmj_spawn(GmjHandle* h, const int32_t* idx_dev, const float* xyr_dev, const uint8_t* mask_dev,
              double xy_range, double rot_range, int tries, int fallback_xy_mm, int fallback_rot_deg,
              uint8_t* ok_dev, void* stream);

/* fused hot path: actions_dev int32[n_envs] (discrete action index), frac_dev float[n_envs] or NULL.
 * Outputs (any may be NULL): obs_dev float[n_envs][n_obs], reward_dev float[n_envs],
 * done_dev uint8[n_envs], events_dev float[n_envs][57]. */
int gmj_step(GmjHandle* h, const int32_t* actions_dev, const float* frac_dev, float* obs_dev,
             float* reward_dev, uint8_t* done_dev, float* events_dev, void* stream);
int gmj_set_action(GmjHandle* h, const int32_t* actions_dev, const float* frac_dev, void* stream);
int gmj_action_step(GmjHandle* h, void* stream);
int gmj_substeps(GmjHandle* h, int n, int sensors, void* stream);
/* observation / reward (side effect on the cumulative reward exactly like MjClass::reward, call once
 * per step) / done / event vector of the CURRENT state without stepping */
int gmj_observe(GmjHandle* h, float* obs_dev, float* reward_dev, uint8_t* done_dev, float* events_dev,
                void* stream);

/* MjEnv.reset() ends with `return self._next_observation()` (rl/env/MjEnv.py:2263): the first observation of the new
 * episode, for the envs in mask_dev only (uint8[n_envs], e.g. the mask gmj_reset_done returned, after gmj_spawn); the
 * other rows of obs_dev are left as they are. No side effects on the env state. */
int gmj_observe_masked(GmjHandle* h, const uint8_t* mask_dev, float* obs_dev, void* stream);

int gmj_n_obs(GmjHandle* h);
int gmj_n_actions(GmjHandle* h);
int gmj_n_envs(GmjHandle* h);
/* copies qpos[n_envs][nq] / qvel[n_envs][nv] (reference joint order, double) to host buffers */
int gmj_get_state_host(GmjHandle* h, double* qpos_host, double* qvel_host);
/* per-env probe, double[n_envs][GMJ_PROBE_LEN] to host: see enum below */
enum { GP_time = 0, GP_wrist_z, GP_gauge0, GP_gauge1, GP_gauge2, GP_si_gauge0, GP_si_gauge1, GP_si_gauge2,
       GP_si_palm, GP_si_wrist, GP_ncon, GP_nrow, GP_niter, GP_unstable, GP_weld_on, GP_base_z,
       GP_tend_x, GP_tend_y, GP_tend_z, GP_tend_th, GP_tnext_x, GP_tnext_y, GP_tnext_z, GP_tnext_th,
       GP_cumulative_reward, GP_has_object, GP_obj_index, GP_episode, GP_last_reward, GP_last_done,
       GP_last_nrow, GP_base_x, GP_base_y, GP_base_yaw,
       /* latest readings of the state sensors, normalised to [-1, 1] (sim_sensors_.read_*_position(),
        * src/mjclass.cpp:3160-3178 un-normalises them in get_state_metres) */
       GP_read_motor_x, GP_read_motor_y, GP_read_motor_z, GP_read_base_x, GP_read_base_y, GP_read_base_z,
       GMJ_PROBE_LEN = 40 };
int gmj_probe_host(GmjHandle* h, double* probe_host);
/* Allocation-free variant of gmj_probe_host. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (tight serving loops, harnesses that forbid allocation; NOT
 * usable while a stream is being captured: it synchronises and copies to pageable host memory). The device scratch is a static 5 MiB array inside the library (held by every GPU that loads it, until exit),
 * so n_envs <= GMJ_NOALLOC_MAX_ENVS; larger handles are rejected with an error. It calls no cudaMalloc / cudaFree. All
 * calls on a device share the array: calls that overlap on the GPU race; results equal gmj_probe_host's. */
#define GMJ_NOALLOC_MAX_ENVS 16384
int gmj_probe_host_noalloc(GmjHandle* h, double* probe_host);
/* MjClass::get_test_report / env_.cnt (src/mjclass.cpp:3800-3809, src/mjclass.h:575-647): per env and event (15 binary
 * then 21 linear, X-macro order) the counters row, abs -> row_abs_host int32[n_envs][36][2], last_value ->
 * last_host float[n_envs][36], and the cumulative reward -> cum_reward_host float[n_envs]. Host-synchronising. */
int gmj_event_track_host(GmjHandle* h, int32_t* row_abs_host, float* last_host, float* cum_reward_host);
int gmj_stats(GmjHandle* h, GmjStats* out);
/* The auto-calibration MjClass::configure_settings runs on the first reset (src/mjclass.cpp:225-314), for hosts that do
 * not use the Python facade: luke::calibrate_reset, then -- as io->set_timestep / calibrate_gauges / set_sim_steps ask --
 * find_highest_stable_timestep, calibrate_simulated_sensors (wrist offset, bending-gauge normalisation under the
 * saturation tip load) and the sim-steps count. Every probe simulation is ONE scratch env created on `device`, stepped
 * with gmj_substeps and destroyed; `settings` is updated in place, create the production handle with it afterwards.
 * Host-synchronising; takes ~0.4 s for 5-segment fingers on a B200. Same arithmetic as gripper_mujoco_b200/calibration.py
 * (tests/test_gpu_parity.py::test_gmj_calibrate_equals_the_python_calibration). */
int gmj_calibrate(const GmjModel* model, GmjSettings* settings, int device, int precision, GmjCalib* io);
/* luke::set_all_constraints / apply_segment_force's weld locking (src/myfunctions.cpp:1157-1170,1642-1727): switch the
 * welds in mask_on (bits 0-2 prismatic, 3-5 revolute, 6 palm) on with their anchors at the current joint values, the
 * others off, in every env. Used by the gauge calibration (tip load with the motor joints locked). */
int gmj_set_welds(GmjHandle* h, int mask_on, void* stream);
const char* gmj_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
