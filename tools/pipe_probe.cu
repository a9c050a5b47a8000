// pipe_probe.cu — dev tool (not on the render path): measures the per-SM issue rate of the
// instruction classes the object loop is made of, so the kernel's inner-loop budget can be
// stated in cycles. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_probe pipe_probe.cu
// Output: one line per probe, warp-instructions per clock per SM at the measured SM clock.
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kIters = 2048;
constexpr int kChains = 8;

#define PROBE(name, DECL, BODY, FIN)                                           \
  __global__ void __launch_bounds__(1024) name(float* out, float a, float b) { \
    DECL;                                                                      \
    _Pragma("unroll 1") for (int i = 0; i < kIters; ++i) {                     \
      _Pragma("unroll") for (int r = 0; r < 4; ++r) {                          \
        _Pragma("unroll") for (int j = 0; j < kChains; ++j) { BODY; }          \
      }                                                                        \
    }                                                                          \
    float s = 0.f;                                                             \
    _Pragma("unroll") for (int j = 0; j < kChains; ++j) { FIN; }               \
    if (s == 123.456f) out[0] = s;                                             \
  }

#define DECL_F float acc[kChains]; for (int j = 0; j < kChains; ++j) acc[j] = (float)(threadIdx.x + j)
#define DECL_U unsigned long long acc[kChains]; for (int j = 0; j < kChains; ++j) acc[j] = ((unsigned long long)__float_as_uint((float)(threadIdx.x + j)) << 32) | __float_as_uint(1.0f + j)

PROBE(k_ffma, DECL_F, acc[j] = fmaf(acc[j], a, b), s += acc[j])
PROBE(k_fmul, DECL_F, acc[j] = __fmul_rn(acc[j], a), s += acc[j])
PROBE(k_fadd, DECL_F, acc[j] = __fadd_rn(acc[j], b), s += acc[j])
PROBE(k_fmnmx, DECL_F, acc[j] = fminf(acc[j], a + (float)i), s += acc[j])
PROBE(k_fmnmx3, DECL_F, acc[j] = fminf(fminf(acc[j], a), b + (float)i), s += acc[j])
PROBE(k_fsel, DECL_F, acc[j] = (acc[j] > a) ? b : acc[j] , s += acc[j])
PROBE(k_mufu_rcp, DECL_F, acc[j] = __frcp_rn(acc[j]), s += acc[j])
PROBE(k_mufu_rcp_approx, DECL_F, asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(acc[j])), s += acc[j])
PROBE(k_mufu_rsq_approx, DECL_F, asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(acc[j])), s += acc[j])
PROBE(k_div_ieee, DECL_F, acc[j] = __fdiv_rn(a, acc[j]), s += acc[j])
PROBE(k_sqrt_ieee, DECL_F, acc[j] = __fsqrt_rn(acc[j]), s += acc[j])

__device__ __forceinline__ unsigned long long pk(float x, float y) { return ((unsigned long long)__float_as_uint(y) << 32) | __float_as_uint(x); }
__device__ __forceinline__ unsigned long long pk2(float lo, float hi) { return pk(fminf(lo, hi), fmaxf(lo, hi)); }
PROBE(k_fadd2, DECL_U; const unsigned long long bb = pk(b, b), asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(acc[j]) : "l"(bb)), s += __uint_as_float((unsigned)acc[j]))
PROBE(k_fmul2, DECL_U; const unsigned long long aa = pk(a, a), asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(acc[j]) : "l"(aa)), s += __uint_as_float((unsigned)acc[j]))
PROBE(k_ffma2, DECL_U; const unsigned long long aa = pk(a, a); const unsigned long long bb = pk(b, b), asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[j]) : "l"(aa), "l"(bb)), s += __uint_as_float((unsigned)acc[j]))

// mixed: one FMA-pipe op + one ALU-pipe op per chain step (do the pipes overlap?)
PROBE(k_mix_fmul_fmnmx, DECL_F, acc[j] = fminf(__fmul_rn(acc[j], a), b + (float)i), s += acc[j])
// mixed packed: FMUL2 + 2 FMNMX
PROBE(k_mix_fmul2_fmnmx, DECL_U; const unsigned long long aa = pk(a, a),
      { asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(acc[j]) : "l"(aa));
        float lo = __uint_as_float((unsigned)acc[j]); float hi = __uint_as_float((unsigned)(acc[j] >> 32));
        acc[j] = pk2(lo, hi); }, s += __uint_as_float((unsigned)acc[j]))

// broadcast LDS.128 (every lane reads the same address, as the object loop does)
__global__ void __launch_bounds__(1024) k_lds128(float* out, float a, float b) {
  __shared__ float4 sm[256];
  if (threadIdx.x < 256) sm[threadIdx.x] = make_float4(a, b, a, b);
  __syncthreads();
  float4 acc = make_float4(0, 0, 0, 0);
#pragma unroll 1
  for (int i = 0; i < kIters; ++i) {
#pragma unroll
    for (int r = 0; r < 32; ++r) {
      float4 v = sm[(i + r * 7) & 255];
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;   // 4 FADD per LDS.128 (kept below the FP32 rate)
    }
  }
  if (acc.x + acc.y + acc.z + acc.w == 123.456f) out[0] = acc.x;
}

template <typename K> double run(K kern, const char* name, int sms, double inst_per_thread, double mhz) {
  float* d; cudaMalloc(&d, 4);
  cudaEvent_t t0, t1; cudaEventCreate(&t0); cudaEventCreate(&t1);
  const int ctas = sms * 2 * 4;
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(t0); kern<<<ctas, 1024>>>(d, 0.999f, 0.001f); cudaEventRecord(t1);
    cudaEventSynchronize(t1);
    float ms; cudaEventElapsedTime(&ms, t0, t1);
    if (rep > 0 && ms < best) best = ms;
  }
  const double warp_inst = inst_per_thread * ctas * 32.0;   // 1024 threads = 32 warps
  const double per_clk_sm = warp_inst / (best * 1e-3) / (mhz * 1e6) / sms;
  printf("%-22s %8.3f ms  %6.3f warp-inst/clk/SM (at %.0f MHz)  err=%s\n", name, best, per_clk_sm, mhz, cudaGetErrorString(cudaGetLastError()));
  cudaFree(d);
  return per_clk_sm;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double mhz = khz / 1000.0; const int sms = p.multiProcessorCount;
  printf("%s, %d SMs, clockRate %.0f MHz (rates assume this clock)\n", p.name, sms, mhz);
  const double n = (double)kIters * 4 * kChains;
  run(k_ffma, "FFMA", sms, n, mhz);
  run(k_fmul, "FMUL", sms, n, mhz);
  run(k_fadd, "FADD", sms, n, mhz);
  run(k_fmnmx, "FMNMX", sms, n, mhz);
  run(k_fmnmx3, "FMNMX3", sms, n, mhz);
  run(k_fsel, "FSETP+FSEL (2 inst)", sms, 2 * n, mhz);
  run(k_fadd2, "FADD2 (f32x2)", sms, n, mhz);
  run(k_fmul2, "FMUL2 (f32x2)", sms, n, mhz);
  run(k_ffma2, "FFMA2 (f32x2)", sms, n, mhz);
  run(k_mix_fmul_fmnmx, "FMUL+FMNMX (2 inst)", sms, 2 * n, mhz);
  run(k_mix_fmul2_fmnmx, "FMUL2+2FMNMX (3 inst)", sms, 3 * n, mhz);
  run(k_mufu_rcp_approx, "MUFU.RCP", sms, n, mhz);
  run(k_mufu_rsq_approx, "MUFU.RSQ", sms, n, mhz);
  run(k_mufu_rcp, "rcp IEEE (seq)", sms, n, mhz);
  run(k_div_ieee, "div IEEE (seq)", sms, n, mhz);
  run(k_sqrt_ieee, "sqrt IEEE (seq)", sms, n, mhz);
  run(k_lds128, "LDS.128 bcast", sms, (double)kIters * 32, mhz);
  return 0;
}
