#!/usr/bin/env python3
"""Dev tool: generates + builds tools/icache_probe.cu — how the B200 SM behaves when warps walk
through a code footprint of F KB at different phases (the situation of the ray-tracing
megakernel on small scenes). Each "block" is 1 KB of code (64 FFMA on 8 chains); a warp runs
blocks (start + k) % M in order through a switch, where start differs per warp.
Prints warp-inst/clk/SM for footprints 4..256 KB at 4, 8 and 16 warps per SMSP."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
M_LIST = [4, 8, 16, 24, 32, 48, 64, 96, 128, 192, 256]
MAXM = max(M_LIST)

src = ['#include <cstdio>', '#include <cuda_runtime.h>',
       '#define B8 a0=fmaf(a0,x,y);a1=fmaf(a1,x,y);a2=fmaf(a2,x,y);a3=fmaf(a3,x,y);a4=fmaf(a4,x,y);a5=fmaf(a5,x,y);a6=fmaf(a6,x,y);a7=fmaf(a7,x,y);',
       '#define B64 B8 B8 B8 B8 B8 B8 B8 B8']
for M in M_LIST:
    src.append(f'__global__ void __launch_bounds__(512) k{M}(float* out, float x, float y, int iters, int stride) {{')
    src.append('  float a0=threadIdx.x,a1=1,a2=2,a3=3,a4=4,a5=5,a6=6,a7=7;')
    src.append(f'  int b = ((threadIdx.x >> 5) * stride + blockIdx.x * 7) % {M};')
    src.append('  for (int it = 0; it < iters; ++it) {')
    src.append('    switch (b) {')
    for c in range(M):
        # distinct constants per block so the compiler cannot merge the cases
        src.append(f'      case {c}: {{ B64 a0 += {c}.0f; }} break;')
    src.append('    }')
    src.append(f'    b = b + 1 == {M} ? 0 : b + 1;')
    src.append('  }')
    src.append('  float s=a0+a1+a2+a3+a4+a5+a6+a7; if (s == 123.456f) out[0] = s;')
    src.append('}')
src.append('''
template <typename K> void run(K k, int M, int warps_per_smsp, int stride, int sms, double mhz) {
  float* d; cudaMalloc(&d, 4);
  cudaEvent_t t0, t1; cudaEventCreate(&t0); cudaEventCreate(&t1);
  const int threads = 512, ctas_per_sm = warps_per_smsp * 4 * 32 / threads;
  const int ctas = sms * ctas_per_sm, iters = 4096;
  float best = 1e30f;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(t0); k<<<ctas, threads>>>(d, 0.999f, 0.001f, iters, stride); cudaEventRecord(t1);
    cudaEventSynchronize(t1);
    float ms; cudaEventElapsedTime(&ms, t0, t1); if (rep > 0 && ms < best) best = ms;
  }
  const double winst = (double)ctas * (threads / 32) * iters * 65.0;
  printf("footprint %3d KB  warps/SMSP %2d  stride %2d : %6.3f warp-inst/clk/SM  (%s)\\n", M, warps_per_smsp, stride,
         winst / (best * 1e-3) / (mhz * 1e6) / sms, cudaGetErrorString(cudaGetLastError()));
  cudaFree(d);
}
int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double mhz = khz / 1000.0; const int sms = p.multiProcessorCount;
  printf("%s %d SMs %.0f MHz; stride 0 = all warps in phase, stride 5 = warps spread over the footprint\\n", p.name, sms, mhz);
  for (int w : {4, 8, 16}) for (int stride : {0, 5}) {''')
for M in M_LIST:
    src.append(f'    run(k{M}, {M}, w, stride, sms, mhz);')
src.append('  }\n  return 0;\n}')
cu = os.path.join(HERE, 'icache_probe.cu')
open(cu, 'w').write('\n'.join(src) + '\n')
r = subprocess.run(['nvcc', '-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-o', os.path.join(HERE, 'icache_probe'), cu])
sys.exit(r.returncode)
