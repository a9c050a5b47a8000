// rt_peak.cu — measurement helper, not on the render path: an FFMA-only micro-kernel that
// measures the FP32 roofline denominator (MEASURED_PEAKS.json has no FP32 entry).
// 8 independent accumulator chains per thread, 1024 threads/CTA, 2 CTAs/SM.
#include "rt_device.h"

namespace rt {

constexpr int kPeakIters = 4096;
constexpr int kPeakChains = 8;

__global__ void __launch_bounds__(1024) fp32_peak_kernel(float* out, float a, float b) {
  float acc[kPeakChains];
#pragma unroll
  for (int j = 0; j < kPeakChains; ++j) acc[j] = (float)(threadIdx.x + j);
#pragma unroll 1
  for (int i = 0; i < kPeakIters; ++i) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int j = 0; j < kPeakChains; ++j) acc[j] = fmaf(acc[j], a, b);
    }
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < kPeakChains; ++j) s += acc[j];
  if (s == 123.456f) out[0] = s;       // never true; keeps the chains alive
}

cudaError_t measure_fp32_peak(int sm_count, cudaStream_t stream, double* tflops, double* ms_out) {
  float* d = nullptr;
  cudaError_t e = cudaMalloc(&d, 4);
  if (e != cudaSuccess) return e;
  cudaEvent_t t0, t1;
  cudaEventCreate(&t0); cudaEventCreate(&t1);
  const int ctas = sm_count * 2 * 8;
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(t0, stream);
    fp32_peak_kernel<<<ctas, 1024, 0, stream>>>(d, 0.999f, 0.001f);
    cudaEventRecord(t1, stream);
    e = cudaEventSynchronize(t1);
    if (e != cudaSuccess) break;
    float ms = 0.f;
    cudaEventElapsedTime(&ms, t0, t1);
    if (rep > 0 && ms < best) best = ms;
  }
  cudaEventDestroy(t0); cudaEventDestroy(t1);
  cudaFree(d);
  if (e != cudaSuccess) return e;
  const double flops = 2.0 * (double)ctas * 1024.0 * kPeakIters * 4.0 * kPeakChains;
  *tflops = flops / (best * 1e-3) / 1e12;
  *ms_out = best;
  return cudaGetLastError();
}

}  // namespace rt
