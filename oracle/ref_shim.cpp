// ref_shim.cpp -- C wrappers around the reference's OWN standalone classes, compiled against the
// sources where they lie (/root/reference/src/gripper.{h,cpp}, slidingwindow.h; see oracle/Makefile
// target `ref`). TEST INFRASTRUCTURE: used to pin the oracle's restatement of the motor kinematics
// and of the sensor ring buffer, and to generate tests/golden/*. No reference source is copied here.
#include <stdexcept>
#include <vector>
#include "gripper.h"
#include "slidingwindow.h"

extern "C" {

void* ref_gripper_new() { return new luke::Gripper(); }
void ref_gripper_free(void* g) { delete (luke::Gripper*)g; }
void ref_gripper_reset(void* g) { ((luke::Gripper*)g)->reset(); }
int ref_gripper_set_xyz_m(void* g, double x, double y, double z) { return ((luke::Gripper*)g)->set_xyz_m(x, y, z); }
int ref_gripper_set_xyz_m_rad(void* g, double x, double th, double z) { return ((luke::Gripper*)g)->set_xyz_m_rad(x, th, z); }
int ref_gripper_step_to(void* g, void* target, int num) { return ((luke::Gripper*)g)->step_to(*(luke::Gripper*)target, num); }
// out: x y z th stepx stepy stepz th_rad(get) prismatic revolute palm
void ref_gripper_get(void* gp, double* out) {
  luke::Gripper* g = (luke::Gripper*)gp;
  out[0] = g->x; out[1] = g->y; out[2] = g->z; out[3] = g->th;
  out[4] = g->step.x; out[5] = g->step.y; out[6] = g->step.z;
  out[7] = g->get_th_rad(); out[8] = g->get_prismatic_joint(); out[9] = g->get_revolute_joint(); out[10] = g->get_palm_joint();
}

void* ref_win_new(int n) { return new luke::SlidingWindow<float>(n); }
void ref_win_free(void* w) { delete (luke::SlidingWindow<float>*)w; }
void ref_win_reset(void* w) { ((luke::SlidingWindow<float>*)w)->reset(); }
void ref_win_add(void* w, float v) { ((luke::SlidingWindow<float>*)w)->add(v); }
float ref_win_read(void* w, int n) { return ((luke::SlidingWindow<float>*)w)->read_element(n); }
float ref_win_read_latest(void* w) { return ((luke::SlidingWindow<float>*)w)->read_element(); }

}  // extern "C"
