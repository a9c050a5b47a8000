// rt_kernel_parity.cu — the "parity" build of the ray-tracing kernel (see rt_kernel.cuh).
// Compiled with -fmad=false: no FMA contraction, IEEE division and square root everywhere.
#define RT_PARITY 1
#define RT_VARIANT_NS parity
#include "rt_kernel.cuh"
#include "rt_packet.cuh"
#include "rt_wavefront.cuh"

namespace rt {
cudaError_t launch_render_parity(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched) {
  return parity::launch_render(kp, sm_count, stream, warps_launched);
}
cudaError_t launch_packet_parity(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched) {
  return parity::pk::launch_packet(kp, sm_count, stream, warps_launched);
}
cudaError_t upload_tables_parity(const float* halton2, const float* halton3, const int* probe_order) {
  cudaError_t e = parity::upload_tables(halton2, halton3);
  if (e != cudaSuccess) return e;
  return parity::pk::upload_probe_order(halton3, probe_order);
}
cudaError_t launch_wf_bounce_parity(const KernelParams& kp, const WfParams& wp, int sm_count, cudaStream_t stream) {
  return parity::pk::launch_wf_bounce(kp, wp, sm_count, stream);
}
cudaError_t launch_wf_shadow_parity(const KernelParams& kp, const WfParams& wp, int sm_count, cudaStream_t stream) {
  return parity::pk::launch_wf_shadow(kp, wp, sm_count, stream);
}
}  // namespace rt
