// gmj_inst.cu -- one explicit instantiation of a step kernel + launcher per translation unit:
//   nvcc ... -DINST_T=float -DINST_N=5 [-DINST_WIDE] -c gmj_inst.cu
#include "gmj_step.cuh"
#ifdef INST_WIDE
template int launch_step_wide_TN<INST_T, INST_N>(GmjHandle*, int, int, const int32_t*, const float*, float*, float*, uint8_t*, float*, cudaStream_t);
#else
template int launch_step_TN<INST_T, INST_N>(GmjHandle*, int, int, const int32_t*, const float*, float*, float*, uint8_t*, float*, cudaStream_t);
#endif
