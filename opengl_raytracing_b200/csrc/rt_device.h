// rt_device.h — types shared by the host side of the C ABI (rt_abi.cu) and the sm_100a
// kernels (rt_kernel.cuh). Not a public header.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <mutex>
#include <string>
#include "../../include/rt_abi.h"

namespace rt {

// Packet-kernel sphere clusters: every kPkClusterSize consecutive sphere slots share one
// cluster record; built from kPkClusterMinSpheres spheres on (rt_abi.cu stage_scene_packet).
// 8 since the warp-cooperative scan (rt_packet.cuh coop_cluster_scan): the warp shares the first level, so twice as many,
// tighter clusters cost it two more steps and save every lane member tests (B200, regroup: C3 10.20 / 10.03 / 10.06 ms and
// C5 4.68 / 4.55 / 6.76 ms at 16 / 8 / 4 spheres per cluster, shared sphere list 128; 32: 10.33 / 4.92). The per-lane walk
// of round 1 preferred 16.
#ifndef RT_PK_CLUSTER
#define RT_PK_CLUSTER 8
#endif
constexpr int kPkClusterSize = RT_PK_CLUSTER;
constexpr int kPkClusterMinSpheres = 64;

// Device-resident scene in the layout the kernel stages into shared memory. Built on the
// host by stage_scene() from the std430 Object[]/Light[] bytes (rt_upload_scene); every
// derived value is computed with the same FP32 operations the shader would use per call,
// so hoisting it is value-exact.
//
//   isect[3*i+0] = (bounds.min.xyz, type bits)         intersectAABB operands (glsl:91-103)
//   isect[3*i+1] = (bounds.max.xyz, plane slot bits)
//   isect[3*i+2] = sphere: (position.xyz, radius*radius)   (glsl:106-109)
//                  plane : (position.xyz, 0)
//   plane[3*k+0] = (normal.xyz,  size.x/2)              intersectPlane operands (glsl:121-146)
//   plane[3*k+1] = (right.xyz,   size.y/2)              frame of glsl:130-138, hoisted
//   plane[3*k+2] = (forward.xyz, 0)
//   material[4*i..] = (albedo.rgb, metallic) (roughness, diffuseStrength, ior, transparency)
//                     (subsurfaceColor.rgb, subsurfaceScatter) (scatterDistance, 0, 0, 0)
//   lights[6*j..]   = the raw 96-byte Light record
struct DeviceScene {
  const float4* isect;
  const float4* plane;
  const float4* material;
  const float4* lights;
  int num_objects;
  int num_planes;
  int num_lights;
  int scene_in_smem;         // 1: isect/plane/lights fit the shared-memory budget
  const uint2* sky;          // 6 * size * size RGBA16F texels (manual FP32 bilinear path)
  unsigned long long sky_tex;  // cudaTextureObject_t cubemap (RT_FLAG_SKYBOX_HW)
  int sky_size;
  const unsigned char* noise;  // R8
  int noise_w, noise_h;
  // packet-kernel scene blob (rt_packet.cuh): cluster records (64 B), sphere records (64 B), plane records (128 B),
  // sphere_index (4 B each), padding to 16 B, raw light records (96 B). Copied verbatim into
  // shared memory by every CTA. pk_float4s == 0: the packet kernel cannot take this scene.
  const float4* pk_blob;
  int pk_float4s;
  int pk_num_spheres, pk_num_planes;
  int pk_num_clusters;       // 0: no cluster level (fewer than kPkClusterMinSpheres spheres)
  int pk_lights_offset;      // bytes from the start of the blob
  int pk_material_offset;    // bytes from the start of the blob to the material records (small scenes), -1 = not in the blob
  int pk_big;                // 1: large-scene kernel (independent warps, two-level boxes)
  int pk_coherent;           // 1: CTA-coherent phases (small scenes), 0: independent warps
  const float4* pk_grid;     // light grids of large scenes (rt_abi.cu build_light_grids), nullptr = none; global memory
  int pk_grid_off_f4, pk_grid_ent_f4;   // float4 offsets of the cell-offset (u32) and entry (u16) arrays inside pk_grid
  int pk_fan_ok;             // 1: every sphere's box contains its sphere (analytic PCSS probe classification allowed)
  int pk_dense;              // large-scene kernel only: 1 = the 80-register build (3 CTAs / 24 warps per SM), 0 = 128 registers (2 / 16)
};

struct KernelParams {
  rt_frame_params fp;
  float tan_fov;             // tan(radians(fov) * 0.5), evaluated on the host (uniform)
  float aspect;
  DeviceScene sc;
  float4* color;
  float4* position;
  void* normal;
  int* object_id;
  unsigned* tile_counter;    // monotonically increasing work counter (never reset)
  unsigned tile_base;        // value of the counter when this launch started
  unsigned tiles_x, tiles_y;
  unsigned blocks_x, blocks_y;   // 32x8-pixel CTA blocks of the coherent mode (4x2 tiles)
  unsigned long long* counters;   // RT_CNT_COUNT reference-literal op counters
};

size_t scene_smem_bytes(int num_objects, int num_planes, int num_lights);

// Resident CTAs per SM of `kernel` at `threads` threads and `smem` bytes of dynamic shared memory, with the
// kernel's dynamic shared-memory limit raised to `smem` — cached per (kernel, device, smem): the two runtime
// calls behind it cost tens of microseconds of host time, which is a visible share of a 0.1 - 0.4 ms frame and of
// every launch of the wavefront pipeline. One cache per kernel instantiation (function-local static of the template).
template <typename K>
inline cudaError_t cached_occupancy(K kernel, int threads, size_t smem, int* per_sm) {
  // (K is the function-pointer TYPE, shared by every kernel of the same signature: the kernel's address is part of the key)
  struct Entry { const void* fn; int device; size_t smem; int per_sm; };
  static Entry cache[32];
  static int used = 0;
  static std::mutex lock;                      // contexts of different devices may launch from different threads
  std::lock_guard<std::mutex> guard(lock);
  int device = 0;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess) return e;
  for (int i = 0; i < used; ++i)
    if (cache[i].fn == (const void*)kernel && cache[i].device == device && cache[i].smem == smem) { *per_sm = cache[i].per_sm; return cudaSuccess; }
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int n = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, smem);
  if (e != cudaSuccess) return e;
  if (n < 1) n = 1;
  // the attribute is a per-device maximum: a cached smaller size stays valid only while no larger one replaced it,
  // so a new size evicts every entry of this device (scenes change rarely; sizes alternate almost never)
  int w = 0;
  for (int i = 0; i < used; ++i) if (cache[i].device != device || cache[i].fn != (const void*)kernel) cache[w++] = cache[i];
  used = w;
  if (used == 32) used = 31;
  cache[used++] = Entry{(const void*)kernel, device, smem, n};
  *per_sm = n;
  return cudaSuccess;
}
// ---- wavefront pipeline (rt_wavefront.cuh): the same pass as three kinds of launches over compacted
// queues in global memory instead of one persistent megakernel. Per sample and per chunk of tile rows:
//   bounce(0) shadow(0) bounce(1) shadow(1) ... bounce(D-1) shadow(D-1) bounce(D)
// bounce(d): shades the hits of depth d-1 from their (BRDF term, shadow factor) records, applies Russian
//            roulette, builds the next ray, traces it (closest hit), finishes the pixels whose path ended, and
//            pushes the surviving hits — compacted with warp ballots — into the hit queue of depth d together
//            with one (BRDF term) record per light and one shadow job per light that needs rays;
// shadow(d): one lane per shadow job (shading point x light), light by light: the cone-culled packets of
//            rt_packet.cuh, writing the job's shadow factor next to its BRDF term.
// Every kernel has its own register budget; every lane of a warp has work. Queue planes are SoA float4.
constexpr int kWfMaxDepth = 16, kWfMaxLights = 32;
constexpr int kWfCounters = 64 + 2 * (kWfMaxDepth + 1) * kWfMaxLights;
__host__ __device__ inline int wf_c_work(int d) { return d; }                                   // work cursor of bounce(d)
__host__ __device__ inline int wf_c_nq(int d) { return 32 + d; }                                // hits pushed by bounce(d)
__host__ __device__ inline int wf_c_jobn(int d, int l) { return 64 + d * kWfMaxLights + l; }    // shadow jobs of (d, light)
__host__ __device__ inline int wf_c_jobcur(int d, int l) { return 64 + (kWfMaxDepth + 1 + d) * kWfMaxLights + l; }
struct WfParams {
  float4* path[2];           // ping-pong, 3 planes each: (rd, pixel bits) (throughput, noise_r) (colour, 0); [k * cap + slot]
  float4* hit[2];            // ping-pong, 2 planes each: (P, object index bits) (N, contributing-light mask bits)
  float4* brdf[2];           // ping-pong, num_lights planes each: (BRDF term of the light, shadow factor)
  unsigned* jobq;            // num_lights planes: hit-queue slot of each shadow job
  float4* acc;               // full-frame running colour sum of a multi-sample render (num_samples > 1), else null
  unsigned* counters;        // kWfCounters, zeroed before every (chunk, sample)
  unsigned cap;              // slots per plane
  int depth;                 // this launch: bounce(depth) / shadow(depth)
  int sample;                // 0 .. num_samples-1
  unsigned tile_row0, tile_rows;   // the chunk: tile rows [tile_row0, tile_row0 + tile_rows) of the launch window
};
bool wavefront_supported(const KernelParams& kp);
size_t packet_smem_bytes(int pk_float4s);   // dynamic shared memory of the packet / wavefront kernels for a blob of that size
// the persistent megakernel with in-CTA ray regrouping (rt_regroup.cuh)
// Four translation units (rt_regroup_{fast,parity}_{small,big}.cu, see rt_regroup_tu.inc): the small- and the large-scene
// kernel run different CTA sizes, and each unit has its own copies of the __constant__ tables.
constexpr int kRgMaxLights = 16;
#define RT_RG_DECL(tu) \
  cudaError_t launch_regroup_##tu(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* advance); \
  size_t regroup_smem_bytes_##tu(int pk_float4s, int num_lights); \
  cudaError_t upload_tables_##tu(const float* halton2, const float* halton3, const int* probe_order);
RT_RG_DECL(fast_small) RT_RG_DECL(fast_big) RT_RG_DECL(parity_small) RT_RG_DECL(parity_big)
#undef RT_RG_DECL
inline size_t regroup_smem_bytes(int pk_float4s, int num_lights, bool big) {     // (the parity units use the same sizes)
  return big ? regroup_smem_bytes_fast_big(pk_float4s, num_lights) : regroup_smem_bytes_fast_small(pk_float4s, num_lights);
}
inline int regroup_max_lights() { return kRgMaxLights; }
inline cudaError_t launch_regroup_fast(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* advance) {
  return kp.sc.pk_big ? launch_regroup_fast_big(kp, sm_count, stream, advance) : launch_regroup_fast_small(kp, sm_count, stream, advance);
}
inline cudaError_t launch_regroup_parity(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* advance) {
  return kp.sc.pk_big ? launch_regroup_parity_big(kp, sm_count, stream, advance) : launch_regroup_parity_small(kp, sm_count, stream, advance);
}
cudaError_t launch_wf_bounce_fast(const KernelParams& kp, const WfParams& wp, int sm_count, cudaStream_t stream);
cudaError_t launch_wf_bounce_parity(const KernelParams& kp, const WfParams& wp, int sm_count, cudaStream_t stream);
cudaError_t launch_wf_shadow_fast(const KernelParams& kp, const WfParams& wp, int sm_count, cudaStream_t stream);
cudaError_t launch_wf_shadow_parity(const KernelParams& kp, const WfParams& wp, int sm_count, cudaStream_t stream);

// One launcher per (variant) translation unit: rt_kernel_fast.cu is compiled with FMA
// contraction, rt_kernel_parity.cu with -fmad=false and IEEE-only arithmetic.
cudaError_t launch_render_fast(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched);
cudaError_t launch_render_parity(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched);

int packet_block_tiles_y();   // CTA block height in tiles (coherent mode), fixed at build time
cudaError_t launch_packet_fast(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched);
cudaError_t launch_packet_parity(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched);

// halton2/halton3: 64 entries each; probe_order: the 16 PCSS probe indices sorted by offset
cudaError_t upload_tables_fast(const float* halton2, const float* halton3, const int* probe_order);
cudaError_t upload_tables_parity(const float* halton2, const float* halton3, const int* probe_order);

// Hooks between rt_abi.cu (owner of rt_context) and rt_post.cu (the f1-f3 stages keep their
// per-context state behind one opaque slot; rt_destroy releases it through post_release).
int ctx_device(rt_context* c);
rt_status ctx_fail(rt_context* c, rt_status s, const std::string& msg);
void** ctx_post_slot(rt_context* c);
void ctx_count_launch(rt_context* c);   // rt_launch_count bookkeeping for kernels launched outside rt_abi.cu
void post_release(void* state);

}  // namespace rt
