// rt_kernel_fast.cu — the "fast" build of the ray-tracing kernel (see rt_kernel.cuh).
// Compiled with -fmad=false; SFU approximations and explicit fmaf only in colour-only shading code.
#define RT_PARITY 0
#define RT_VARIANT_NS fast
#include "rt_kernel.cuh"

namespace rt {
cudaError_t launch_render_fast(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched) {
  return fast::launch_render(kp, sm_count, stream, warps_launched);
}
cudaError_t upload_tables_fast(const float* halton2, const float* halton3) {
  return fast::upload_tables(halton2, halton3);
}
}  // namespace rt
