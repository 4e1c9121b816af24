/* empty stand-in for MuJoCo's mjxmacro.h (see mujoco.h in this directory) */
