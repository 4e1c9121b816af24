// ref_tail_shim.cpp -- drives the reference's OWN MjClass (src/mjclass.{h,cpp}, compiled in place from
// /root/reference/src against the stand-in headers of oracle/stub/, see oracle/Makefile target `ref`) over SCRIPTED
// physics observables. TEST INFRASTRUCTURE: it pins the oracle's non-physics tail -- configure_settings (action table,
// sampler selection, update_sensor_settings), set_action, monitor_sensors, sense_gripper_state, update_env,
// update_events, calc_rewards / reward, is_done, get_observation and its seven samplers, spawn_object /
// spawn_into_scene's Box2d test -- to what the reference's compiled code does, and generates tests/golden/tail_*.npz.
// No reference source is copied: mjclass.cpp, gripper.cpp and the headers are compiled where they lie.
//
// What is NOT the reference here: the `luke::` functions below. In the reference they live in src/myfunctions.cpp
// and read MuJoCo's mjData; MuJoCo is absent, so each is replaced by a stand-in that returns the scripted value of
// the current substep (forces, gauges, wrist force, object pose) or, for the motor / base target, forwards to the
// reference's real luke::Target / luke::Gripper (customtypes.h, gripper.cpp) exactly as the cited lines do.
#include <stdexcept>
#include <vector>
#include <string>
#include <cstring>
#include <cmath>
#include "mjclass.h"
#include "../include/gmj_types.h"

#define REF_ROWLEN 46

namespace fake {
luke::Target target;                 // target_ (src/myfunctions.cpp:470)
const double* script = nullptr;      // rows of REF_ROWLEN doubles, one per substep
int nrows = 0, cursor = 0;
double cur[REF_ROWLEN];
double finger_length = 235e-3, EI = 1.0, clearance = 10e-3, hook_length = 35e-3, hook_angle_deg = 75, finger_width = 28e-3;
double lim[8] = { -30e-3, 30e-3, -0.5, 0.5, -0.5, 0.5, -1.5707963, 1.5707963 };   // z, x, y, yaw (min, max)
int N = 5;
bool live = false;
double z_rest = 0, bbox[3] = { 0, 0, 0 };
void update_base_limits() {          // src/myfunctions.cpp:2286-2308
  target.base_min.z = lim[0]; target.base_max.z = lim[1]; target.base_min.x = lim[2]; target.base_max.x = lim[3];
  target.base_min.y = lim[4]; target.base_max.y = lim[5]; target.base_min.yaw = lim[6]; target.base_max.yaw = lim[7];
  target.base_min.roll = -0.5; target.base_max.roll = 0.5; target.base_min.pitch = -0.5; target.base_max.pitch = 0.5;
}
}  // namespace fake

namespace luke {
// ---- string helpers of src/myfunctions.cpp:137-160 (ActionSetting::update_action_function picks the mover by name)
bool str_starts_with(std::string ref_str, std::string check_str) {
  return check_str.size() <= ref_str.size() && ref_str.substr(0, check_str.size()) == check_str;
}
bool str_ends_with(std::string ref_str, std::string check_str) {
  return check_str.size() <= ref_str.size() && ref_str.substr(ref_str.size() - check_str.size(), check_str.size()) == check_str;
}
// colours: MjClass::init -> set_neat_colours writes geom_rgba only (out of scope, no effect on the path)
void set_finger_colour(mjModel*, std::vector<float>, int) {}
void set_ground_colour(mjModel*, std::vector<float>) {}
void set_object_colour(mjModel*, std::vector<float>) {}
void set_main_body_colour(mjModel*, std::vector<float>) {}
void set_everything_colour(mjModel*, std::vector<float>) {}
void set_all_objects_colour(mjModel*, std::vector<float>) {}
// ---- no-op / scripted stand-ins for the MuJoCo-facing half of src/myfunctions.cpp
void init(mjModel*, mjData*) {}
void init_J(mjModel*, mjData*) {}
void set_finger_stiffness_using_model(mjModel*) {}
void before_step(mjModel*, mjData*) {}
void after_step(mjModel*, mjData*) {}       // stepper / weld toggling act on the physics only (pinned separately: luke::Gripper)
void resolve_segment_forces(mjModel*, mjData*) {}
void step(mjModel* m, mjData* d) {          // mj_step advances the clock by opt.timestep; the observables come from the script
  d->time += m->opt.timestep;
  if (fake::script && fake::cursor < fake::nrows) {
    std::memcpy(fake::cur, fake::script + (size_t)fake::cursor * REF_ROWLEN, sizeof(fake::cur));
    fake::cursor++;
  }
  d->userdata[2] = fake::cur[39];
}
void reset(mjModel*, mjData* d) {           // src/myfunctions.cpp:544-570 (target, clock, limits)
  fake::target.reset();
  fake::update_base_limits();
  d->time = 0;
  fake::live = false;
  std::memset(fake::cur, 0, sizeof(fake::cur));
  d->userdata[2] = 0;
}
std::vector<std::string> get_objects() { return std::vector<std::string>{ "scripted_object" }; }
std::vector<double> get_base_min() { auto& b = fake::target.base_min; return { b.x, b.y, b.z, b.roll, b.pitch, b.yaw }; }
std::vector<double> get_base_max() { auto& b = fake::target.base_max; return { b.x, b.y, b.z, b.roll, b.pitch, b.yaw }; }
std::vector<gfloat> get_gauge_data(const mjModel*, mjData*) {
  return { (gfloat)fake::cur[0], (gfloat)fake::cur[1], (gfloat)fake::cur[2] };
}
Forces_faster get_object_forces_faster(const mjModel*, mjData*) {
  Forces_faster F;
  if (!fake::live) { F.empty = true; return F; }
  F.obj.resize(1);
  const double* c = fake::cur;
  for (int k = 0; k < 3; k++) {
    F.all.finger1_local[k] = c[3 + k]; F.all.finger2_local[k] = c[6 + k]; F.all.finger3_local[k] = c[9 + k];
    F.all.palm_local[k] = c[12 + k];
    F.obj[0].finger1_local[k] = c[15 + k]; F.obj[0].finger2_local[k] = c[18 + k]; F.obj[0].finger3_local[k] = c[21 + k];
    F.obj[0].palm_local[k] = c[24 + k]; F.obj[0].ground[k] = c[27 + k];
    F.gnd.finger1_local[k] = c[30 + k]; F.gnd.finger2_local[k] = c[33 + k]; F.gnd.finger3_local[k] = c[36 + k];
  }
  return F;
}
std::vector<QPos> get_object_qpos(mjModel*, mjData*) {
  std::vector<QPos> v;
  if (fake::live) v.push_back(QPos(fake::cur[40], fake::cur[41], fake::cur[42], 0, 0, 0, 1));
  return v;
}
std::vector<std::vector<double>> get_object_XY_relative_to_gripper(mjModel*, mjData*) {   // src/myfunctions.cpp:3875-3896
  std::vector<std::vector<double>> out;
  JointStates s = fake::target.get_target_m();
  if (fake::live) out.push_back({ s.base_x - fake::cur[40], s.base_y - fake::cur[41] });
  return out;
}
bool is_object_live(int) { return fake::live; }
void reset_object(mjModel*, mjData*) { fake::live = false; }
QPos spawn_object(mjModel*, mjData*, int, QPos pose) {       // ObjectHandler::spawn_object: z = keyframe z + 1e-6
  fake::live = true;
  pose.z = fake::z_rest + 1e-6;
  fake::cur[40] = pose.x; fake::cur[41] = pose.y; fake::cur[42] = pose.z;
  return pose;
}
Vec3 get_object_xyz_bounding_box(int) { Vec3 v; v.x = fake::bbox[0]; v.y = fake::bbox[1]; v.z = fake::bbox[2]; return v; }
std::vector<std::vector<double>> get_live_object_bounding_boxes() { return {}; }
int get_N() { return fake::N; }
bool use_segments() { return true; }
bool use_base_xyz() { return false; }
// ---- motor / base target: thin forwards onto the reference's real Target / Gripper, as in src/myfunctions.cpp
JointStates get_target_state() { return fake::target.get_target_m(); }                    // :3324-3329
std::vector<gfloat> get_target_state_vector() {                                           // :3331-3356, nb = 1
  JointStates s = fake::target.get_target_m();
  return { (gfloat)s.gripper_x, (gfloat)s.gripper_y, (gfloat)s.gripper_z, (gfloat)s.base_z };
}
Gripper get_gripper_target() { return fake::target.end; }
bool move_gripper_target_m(double x, double y, double z) {                               // :2422-2431
  fake::target.last_robot = Target::Robot::gripper;
  return fake::target.end.set_xyz_m(fake::target.end.x + x, fake::target.end.y + y, fake::target.end.z + z);
}
bool move_gripper_target_m_rad(double x, double th, double z) {                          // :2433-2442
  fake::target.last_robot = Target::Robot::gripper;
  return fake::target.end.set_xyz_m_rad(fake::target.end.x + x, fake::target.end.th + th, fake::target.end.z + z);
}
static bool clamp_base_xyz() {
  Target& t = fake::target; bool wl = true;
  if (t.base.x > t.base_max.x) { t.base.x = t.base_max.x; wl = false; }
  if (t.base.x < t.base_min.x) { t.base.x = t.base_min.x; wl = false; }
  if (t.base.y > t.base_max.y) { t.base.y = t.base_max.y; wl = false; }
  if (t.base.y < t.base_min.y) { t.base.y = t.base_min.y; wl = false; }
  if (t.base.z > t.base_max.z) { t.base.z = t.base_max.z; wl = false; }
  if (t.base.z < t.base_min.z) { t.base.z = t.base_min.z; wl = false; }
  return wl;
}
bool move_base_target_m(double x, double y, double z) {                                  // :2444-2486
  fake::target.last_robot = Target::Robot::panda;
  fake::target.base.x += x; fake::target.base.y += y; fake::target.base.z += z;
  return clamp_base_xyz();
}
bool move_base_target_rad(double, double, double yaw) {                                  // :2488-2511
  Target& t = fake::target; bool wl = true;
  t.last_robot = Target::Robot::panda;
  t.base.yaw += yaw;
  if (t.base.yaw > t.base_max.yaw) { t.base.yaw = t.base_max.yaw; wl = false; }
  if (t.base.yaw < t.base_min.yaw) { t.base.yaw = t.base_min.yaw; wl = false; }
  return wl;
}
bool lift_base_to_height(double z) {                                                     // :2513-2536
  Target& t = fake::target; bool wl = true;
  t.last_robot = Target::Robot::panda;
  t.base.z = -z;
  if (t.base.z > t.base_max.z) { t.base.z = t.base_max.z; wl = false; }
  if (t.base.z < t.base_min.z) { t.base.z = t.base_min.z; wl = false; }
  return wl;
}
bool set_base_to_Z_position(mjData*, float z) {                                          // :2610-2640 (target part)
  Target& t = fake::target; bool wl = true;
  t.base.z = z;
  if (t.base.z > t.base_max.z) { t.base.z = t.base_max.z; wl = false; }
  if (t.base.z < t.base_min.z) { t.base.z = t.base_min.z; wl = false; }
  return wl;
}
void set_base_XYZ_limits(double x, double y, double z) {                                 // :2310-2324
  fake::lim[2] = -x; fake::lim[3] = x; fake::lim[4] = -y; fake::lim[5] = y; fake::lim[0] = -z; fake::lim[1] = z;
  fake::update_base_limits();
}
void set_base_yaw_limit(double yaw) { fake::lim[6] = -yaw; fake::lim[7] = yaw; fake::update_base_limits(); }
float get_fingertip_z_height() {                                                         // :3608-3620
  float straight_finger_distance = -fake::target.base_min.z - fake::target.base.z;
  float tip_lift = fake::finger_length * (1 - std::cos(fake::target.end.get_th_rad()));
  float height_above_min = straight_finger_distance + tip_lift;
  return height_above_min + fake::target.base_min.z;
}
std::vector<luke::Vec3> get_fingerend_and_palm_xyz(std::vector<double> finger_forces_SI) {   // :3622-3688
  std::vector<luke::Vec3> pos(4);
  Target& t = fake::target;
  double base_x = t.base.x, base_y = t.base.y, base_z_rot = t.base.yaw;
  double fing_x = t.end.get_x_m(), fing_th = t.end.get_th_rad(), palm_z = t.end.get_z_m();
  constexpr double fingerend_to_palm_Z = 165e-3;
  constexpr double PI_23 = M_PI * (2.0 / 3.0);
  constexpr double angles[3] = { 0.0, PI_23, 2 * PI_23 };
  double untilted_fingerend_height = fake::clearance - t.base.z;
  double tip_lift = fake::finger_length * (1 - std::cos(fing_th));
  double tilted_fingerend_height = untilted_fingerend_height + tip_lift;
  for (int i = 0; i < 3; i++) {
    double tilt_fing_x = fing_x - fake::finger_length * std::sin(fing_th);
    double deflection = finger_forces_SI[i] * std::pow(fake::finger_length, 3) / (3 * fake::EI);
    double deflection_x = deflection * std::cos(fing_th);
    double deflection_z = deflection * std::sin(fing_th);
    double final_fing_x = tilt_fing_x + deflection_x;
    luke::Vec3 tippos;
    tippos.x = -final_fing_x * sin(angles[i] + base_z_rot) + base_x;
    tippos.y = -final_fing_x * cos(angles[i] + base_z_rot) + base_y;
    tippos.z = tilted_fingerend_height - deflection_z;
    pos[i] = tippos;
  }
  luke::Vec3 palmpos;
  palmpos.x = base_x; palmpos.y = base_y; palmpos.z = untilted_fingerend_height + fingerend_to_palm_Z - palm_z;
  pos[3] = palmpos;
  return pos;
}
std::vector<luke::Vec3> get_finger_hook_locations() {                                    // :3717-3761
  std::vector<luke::Vec3> pos(6);
  Target& t = fake::target;
  double fing_x = t.end.get_x_m();
  double hook_th = fake::hook_angle_deg * (M_PI / 180.000);
  constexpr double PI_23 = M_PI * (2.0 / 3.0);
  constexpr double angles[3] = { 0.0, PI_23, 2 * PI_23 };
  for (int i = 0; i < 3; i++) {
    double hook_x = 0.5 * fake::hook_length * sin(hook_th);
    luke::Vec3 tippos, tipinfo;
    tippos.x = -(fing_x - hook_x) * sin(angles[i]) + t.base.x;
    tippos.y = -(fing_x - hook_x) * cos(angles[i]) + t.base.y;
    tippos.z = 0.0;
    tipinfo.x = fake::hook_length; tipinfo.y = fake::finger_width; tipinfo.z = angles[i];
    pos[i] = tippos; pos[3 + i] = tipinfo;
  }
  return pos;
}
}  // namespace luke

// MuJoCo entry points the compiled reference code can reach on the driven paths
void mj_forward(const mjModel*, mjData*) {}
void mj_deleteData(mjData*) {}
void mj_deleteModel(mjModel*) {}

// ------------------------------------------------------------------------------------------ settings transfer
static void apply_settings(MjType::Settings& s, const GmjSettings* g) {
  const double* G = g->general;
  s.debug = false;
  s.auto_set_timestep = false; s.auto_calibrate_gauges = false; s.auto_sim_steps = false; s.auto_exceed_lateral_lim = false;
  s.mujoco_timestep = G[GS_mujoco_timestep]; s.curve_validation = false; s.tip_force_applied = 0;
  s.random_seed = (uint)G[GS_random_seed];
  s.time_for_action = G[GS_time_for_action]; s.saturation_yield_factor = G[GS_saturation_yield_factor];
  s.sensor_sample_mode = (int)G[GS_sensor_sample_mode]; s.state_sample_mode = (int)G[GS_state_sample_mode];
  s.sensor_n_prev_steps = (int)G[GS_sensor_n_prev_steps]; s.state_n_prev_steps = (int)G[GS_state_n_prev_steps];
  s.sensor_noise_mag = G[GS_sensor_noise_mag]; s.sensor_noise_mu = G[GS_sensor_noise_mu]; s.sensor_noise_std = G[GS_sensor_noise_std];
  s.state_noise_mag = G[GS_state_noise_mag]; s.state_noise_mu = G[GS_state_noise_mu]; s.state_noise_std = G[GS_state_noise_std];
  s.base_position_noise = G[GS_base_position_noise]; s.palm_scale_factor = G[GS_palm_scale_factor];
  s.oob_distance = G[GS_oob_distance]; s.lift_height = G[GS_lift_height]; s.gripper_target_height = G[GS_gripper_target_height];
  s.stable_finger_force = G[GS_stable_finger_force]; s.stable_palm_force = G[GS_stable_palm_force];
  s.stable_finger_force_lim = G[GS_stable_finger_force_lim]; s.stable_palm_force_lim = G[GS_stable_palm_force_lim];
  s.XY_distance_threshold = G[GS_XY_distance_threshold];
  s.cap_reward = G[GS_cap_reward] != 0; s.quit_if_cap_exceeded = G[GS_quit_if_cap_exceeded] != 0;
  s.reward_cap_lower_bound = G[GS_reward_cap_lower_bound]; s.reward_cap_upper_bound = G[GS_reward_cap_upper_bound];
  s.continous_actions = G[GS_continous_actions] != 0; s.use_termination_action = G[GS_use_termination_action] != 0;
  s.termination_threshold = G[GS_termination_threshold]; s.lift_on_termination = G[GS_lift_on_termination] != 0;
  s.sim_steps_per_action = (int)G[GS_sim_steps_per_action]; s.fingertip_min_mm = G[GS_fingertip_min_mm];
  s.render_on_step = false;
  int k = 0;
#define SS(NAME, USED, NORM, RATE) { const double* v = g->sensor[k++]; auto& z = s.NAME; \
    z.in_use = v[GSF_in_use] != 0; z.normalise = (float)v[GSF_normalise]; z.read_rate = (float)v[GSF_read_rate]; \
    z.use_normalisation = v[GSF_use_normalisation] != 0; z.use_noise = v[GSF_use_noise] != 0; \
    z.raw_value_offset = (float)v[GSF_raw_value_offset]; z.noise_mag = (float)v[GSF_noise_mag]; \
    z.noise_mu = (float)v[GSF_noise_mu]; z.noise_std = (float)v[GSF_noise_std]; z.noise_overriden = v[GSF_noise_overriden] != 0; }
  LUKE_MJSETTINGS_SENSOR
#undef SS
  k = 0;
#define AA(NAME, USED, VALUE, SIGN) { const double* v = g->action[k++]; s.NAME.in_use = v[GAF_in_use] != 0; \
    s.NAME.value = v[GAF_value]; s.NAME.sign = (int)v[GAF_sign]; }
  LUKE_MJSETTINGS_ACTION
#undef AA
  k = 0;
#define BR(NAME, REWARD, DONE, TRIGGER) { const double* v = g->binary[k++]; s.NAME.reward = (float)v[GBF_reward]; \
    s.NAME.done = (int)v[GBF_done]; s.NAME.trigger = (int)v[GBF_trigger]; }
  LUKE_MJSETTINGS_BINARY_REWARD
#undef BR
  k = 0;
#define LR(NAME, REWARD, DONE, TRIGGER, MIN, MAX, OVERSHOOT) { const double* v = g->linear[k++]; s.NAME.reward = (float)v[GLF_reward]; \
    s.NAME.done = (int)v[GLF_done]; s.NAME.trigger = (int)v[GLF_trigger]; s.NAME.min = (float)v[GLF_min]; \
    s.NAME.max = (float)v[GLF_max]; s.NAME.overshoot = (float)v[GLF_overshoot]; }
  LUKE_MJSETTINGS_LINEAR_REWARD
#undef LR
}

struct RefMj {
  mjModel model; mjData data; double userdata[8];
  MjClass* mj;
};

extern "C" {

int ref_mj_rowlen(void) { return REF_ROWLEN; }

// geom: finger_length, EI, fingertip_clearance, hook_length, hook_angle_deg, finger_width, N
void* ref_mj_new(const GmjSettings* g, const double* geom) {
  RefMj* r = new RefMj();
  std::memset(&r->model, 0, sizeof(r->model)); std::memset(&r->data, 0, sizeof(r->data));
  std::memset(r->userdata, 0, sizeof(r->userdata));
  r->data.userdata = r->userdata;
  fake::finger_length = geom[0]; fake::EI = geom[1]; fake::clearance = geom[2]; fake::hook_length = geom[3];
  fake::hook_angle_deg = geom[4]; fake::finger_width = geom[5]; fake::N = (int)geom[6];
  const double* G = g->general;
  fake::lim[0] = G[GS_base_z_min]; fake::lim[1] = G[GS_base_z_max]; fake::lim[2] = G[GS_base_x_min]; fake::lim[3] = G[GS_base_x_max];
  fake::lim[4] = G[GS_base_y_min]; fake::lim[5] = G[GS_base_y_max]; fake::lim[6] = G[GS_base_yaw_min]; fake::lim[7] = G[GS_base_yaw_max];
  fake::script = nullptr; fake::nrows = 0; fake::cursor = 0;
  MjType::Settings s;
  apply_settings(s, g);
  // quirk Q2: state_sample_mode == scaled_change_sq assigns sampleFcnPtr and leaves stateFcnPtr as it was. A fresh
  // object has no earlier value (uninitialised member: undefined behaviour), so the defined path is taken -- the
  // object starts with the header default (sign) as every bind.so user's does, then the mode is switched.
  const int want_state_mode = s.state_sample_mode;
  if (want_state_mode == MjType::Sample::scaled_change_sq) s.state_sample_mode = MjType::Sample::sign;
  r->mj = new MjClass(s);
  r->mj->sim_gauge_raw_to_N_factor = G[GS_sim_gauge_raw_to_N_factor];
  r->mj->model = &r->model; r->mj->data = &r->data;
  r->mj->init();                       // luke::init -> reset -> configure_settings (the reference's own)
  if (want_state_mode != r->mj->s_.state_sample_mode) { r->mj->s_.state_sample_mode = want_state_mode; r->mj->reset(); }
  r->mj->env_.grp.peak_finger_lateral_force = 0;     // quirk Q4: no initialiser in the reference; defined as 0
  r->mj->env_.grp.peak_finger_axial_force = 0;
  return r;
}
void ref_mj_free(void* p) { RefMj* r = (RefMj*)p; r->mj->model = NULL; r->mj->data = NULL; delete r->mj; delete r; }
void ref_mj_reset(void* p) {
  RefMj* r = (RefMj*)p;
  r->mj->reset();
  r->mj->env_.grp.peak_finger_lateral_force = 0; r->mj->env_.grp.peak_finger_axial_force = 0;
}
void ref_mj_spawn(void* p, double x, double y, double zrot, double z_rest, const double* bbox) {
  fake::z_rest = z_rest; for (int k = 0; k < 3; k++) fake::bbox[k] = bbox[k];
  ((RefMj*)p)->mj->spawn_object(0, x, y, zrot);
}
// the reference's own spawn test: spawn_into_scene with zero ranges tries exactly the one pose (x, y, rot)
int ref_mj_try_spawn(void* p, double x, double y, double zrot, double z_rest, const double* bbox) {
  RefMj* r = (RefMj*)p;
  fake::z_rest = z_rest; for (int k = 0; k < 3; k++) fake::bbox[k] = bbox[k];
  r->mj->reset_object();
  return r->mj->spawn_into_scene(0, x, y, zrot, 0.0, 0.0, 0.0) ? 1 : 0;
}
void ref_mj_set_script(void*, const double* rows, int nrows) { fake::script = rows; fake::nrows = nrows; fake::cursor = 0; }
int ref_mj_script_cursor(void*) { return fake::cursor; }
int ref_mj_n_actions(void* p) { return ((RefMj*)p)->mj->n_actions; }
int ref_mj_n_obs(void* p) { return (int)((RefMj*)p)->mj->get_observation().size(); }
// one MjEnv.step: _take_action (set_action + action_step), _next_observation, _is_done, _reward
// (rl/env/MjEnv.py:2170-2220). continuous != 0: MjClass::set_continous_action (clamps the fraction first).
int ref_mj_step(void* p, int action, float fraction, int continuous, float* obs, float* reward, int* done, float* events57,
                float* target4) {
  MjClass* mj = ((RefMj*)p)->mj;
  std::vector<float> t = continuous ? mj->set_continous_action(action, fraction) : mj->set_action(action, fraction);
  for (size_t i = 0; i < t.size() && i < 4; i++) target4[i] = t[i];
  mj->action_step();
  std::vector<luke::gfloat> o = mj->get_observation();
  for (size_t i = 0; i < o.size(); i++) obs[i] = o[i];
  *done = mj->is_done() ? 1 : 0;
  *reward = mj->reward();
  std::vector<float> ev = mj->get_event_state();
  for (size_t i = 0; i < ev.size() && i < 57; i++) events57[i] = ev[i];
  return (int)o.size();
}
// sensor bookkeeping of Settings::update_sensor_settings (quirk Q11): prev_steps, readings_per_step per X-macro sensor
void ref_mj_sensor_layout(void* p, int* prev_steps, int* readings) {
  MjType::Settings& s = ((RefMj*)p)->mj->s_;
  int k = 0;
#define SS(NAME, USED, NORM, RATE) prev_steps[k] = s.NAME.prev_steps; readings[k] = s.NAME.readings_per_step; k++;
  LUKE_MJSETTINGS_SENSOR
#undef SS
}
// free helpers of the reference (src/mjclass.cpp:4866-4904, src/customtypes.h:35-172)
float ref_linear_reward(float v, float mn, float mx, float ov) { return linear_reward(v, mn, mx, ov); }
float ref_normalise_between(float v, float mn, float mx) { return normalise_between(v, mn, mx); }
int ref_box2d_overlap(double cx1, double cy1, double w1, double h1, double th1, double cx2, double cy2, double w2, double h2,
                      double th2, double gap) {
  luke::Box2d a, b;
  a.initCentre(cx1, cy1, w1, h1); a.rotate(th1);
  b.initCentre(cx2, cy2, w2, h2); b.rotate(th2);
  return a.overlapsWith(b, gap) ? 1 : 0;
}
// one sampler call on a window filled oldest-first with vals[0..n)
int ref_sample(int mode, int prev_steps, int readings_per_step, const float* vals, int n, float* out) {
  MjType::Sensor s(true, 1, 10);
  s.prev_steps = prev_steps; s.readings_per_step = readings_per_step; s.total_readings = 1 + prev_steps * readings_per_step;
  luke::SlidingWindow<luke::gfloat> w(1000);
  for (int i = 0; i < n; i++) w.add(vals[i]);
  std::vector<luke::gfloat> o;
  switch (mode) {
    case 0: o = s.raw_sample(w); break; case 1: o = s.change_sample(w); break; case 2: o = s.average_sample(w); break;
    case 3: o = s.median_sample(w); break; case 4: o = s.sign_sample(w); break; case 5: o = s.scaled_change_sample(w); break;
    default: o = s.scaled_change_sq_sample(w); break;
  }
  for (size_t i = 0; i < o.size(); i++) out[i] = o[i];
  return (int)o.size();
}

}  // extern "C"
