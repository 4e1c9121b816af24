// fma_peak.cu -- FP32 FMA-pipe microbenchmark (measurement tooling for bench.py's roofline: the
// driver's MEASURED_PEAKS.json has no non-tensor FP32 entry, SURVEY.md 8(d) asks for a measured one).
// extern "C" gmj_fma_peak(device, iters, *tflops): best of 5 timed launches, 16 independent FMA
// chains per thread, 148*8 blocks of 256 threads.
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) k_fma(float* out, int iters, float a, float b) {
  float x[16];
#pragma unroll
  for (int k = 0; k < 16; k++) x[k] = threadIdx.x * 1e-3f + k;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = fmaf(x[k], a, b);
  }
  float s = 0;
#pragma unroll
  for (int k = 0; k < 16; k++) s += x[k];
  if (s == 12345.678f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;   // never true: keeps the chain live
}

extern "C" int gmj_fma_peak(int device, int iters, double* tflops) {
  if (cudaSetDevice(device) != cudaSuccess) return 1;
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  int blocks = sms * 8, threads = 256;
  float* out = nullptr;
  if (cudaMalloc(&out, sizeof(float) * blocks * threads) != cudaSuccess) return 1;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best = 0;
  for (int rep = 0; rep < 6; rep++) {
    cudaEventRecord(e0);
    k_fma<<<blocks, threads>>>(out, iters, 0.999f, 1e-3f);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return 1; }
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    double tf = 2.0 * 16.0 * (double)iters * blocks * threads / (ms * 1e-3) * 1e-12;
    if (rep > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
  *tflops = best;
  return 0;
}
