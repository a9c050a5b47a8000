// rt_kernel_fast.cu — the "fast" build of the ray-tracing kernel (see rt_kernel.cuh).
// Compiled with -fmad=false; SFU approximations and explicit fmaf only in colour-only shading code.
#define RT_PARITY 0
#define RT_VARIANT_NS fast
#include "rt_kernel.cuh"
#include "rt_packet.cuh"
#include "rt_wavefront.cuh"

namespace rt {
cudaError_t launch_render_fast(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched) {
  return fast::launch_render(kp, sm_count, stream, warps_launched);
}
int packet_block_tiles_y() { return fast::pk::kBlockTilesY; }
cudaError_t launch_packet_fast(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched) {
  return fast::pk::launch_packet(kp, sm_count, stream, warps_launched);
}
cudaError_t upload_tables_fast(const float* halton2, const float* halton3, const int* probe_order) {
  cudaError_t e = fast::upload_tables(halton2, halton3);
  if (e != cudaSuccess) return e;
  return fast::pk::upload_probe_order(halton3, probe_order);
}
cudaError_t launch_wf_bounce_fast(const KernelParams& kp, const WfParams& wp, int sm_count, cudaStream_t stream) {
  return fast::pk::launch_wf_bounce(kp, wp, sm_count, stream);
}
cudaError_t launch_wf_shadow_fast(const KernelParams& kp, const WfParams& wp, int sm_count, cudaStream_t stream) {
  return fast::pk::launch_wf_shadow(kp, wp, sm_count, stream);
}
size_t packet_smem_bytes(int pk_float4s) {
  return (size_t)pk_float4s * sizeof(float4) + (size_t)fast::pk::kScratchBytes;
}
bool wavefront_supported(const KernelParams& kp) {
  return kp.sc.pk_float4s > 0 && kp.sc.num_lights <= kWfMaxLights && kp.fp.max_ray_depth <= kWfMaxDepth &&
         kp.fp.width <= 65535 && kp.fp.height <= 65535;
}
}  // namespace rt
