/* Stand-in for DeepMind MuJoCo's <mujoco.h> (MuJoCo 2.1.5 is a binary dependency of the reference and is not
 * installed here). TEST INFRASTRUCTURE: it declares only the handful of names the reference's src/mjclass.{h,cpp},
 * src/customtypes.h and src/mynum.h mention, so that those files compile IN PLACE from /root/reference/src into
 * oracle/_ref/ (oracle/Makefile, target `ref`). Nothing here simulates anything: the physics observables the
 * compiled reference code consumes are scripted by oracle/ref_tail_shim.cpp. */
#pragma once
#include <stdint.h>
#include <cstdint>
typedef double mjtNum;
typedef unsigned char mjtByte;
struct mjOption { mjtNum timestep; };
struct mjModel { int nq, nv; mjOption opt; };
struct mjData { mjtNum time; mjtNum* qpos; mjtNum* qvel; mjtNum* userdata; };
void mju_zero(mjtNum*, int);
void mju_printMat(const mjtNum*, int, int);
void mju_transpose(mjtNum*, const mjtNum*, int, int);
void mju_mulMatMat(mjtNum*, const mjtNum*, const mjtNum*, int, int, int);
void mj_forward(const mjModel*, mjData*);
void mj_deleteData(mjData*);
void mj_deleteModel(mjModel*);
mjModel* mj_loadXML(const char*, const void*, char*, int);
mjData* mj_makeData(const mjModel*);
void mju_error_s(const char*, const char*);
