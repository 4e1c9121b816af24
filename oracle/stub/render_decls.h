/* Pre-included (-include) when the reference's src/mjclass.cpp is compiled for oracle/_ref with -DRENDERING_H_:
 * the reference's rendering.h needs GLFW/OpenGL (unavailable on this box), and its own LUKE_CLUSTER branch does not
 * compile, so the include guard is pre-defined and the render:: entry points mjclass.cpp mentions are declared here.
 * They are never called (abort stubs, oracle/gen_ref_stubs.py). */
#pragma once
#include "customtypes.h"
class MjClass;
namespace render {
void init_camera(MjClass&); void init_window(MjClass&); void resize_camera_window(int, int); void create_camera_window(int, int);
void reload_for_rendering(MjClass&); bool render_window(); bool render_camera(); bool render_camera_with_seg_mask();
void render_rgbd_feed(); void finish_window(); void finish_camera(); luke::RGBD read_rgbd(); luke::RGBD read_mask();
}
