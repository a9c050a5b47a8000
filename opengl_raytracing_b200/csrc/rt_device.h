// rt_device.h — types shared by the host side of the C ABI (rt_abi.cu) and the sm_100a
// kernels (rt_kernel.cuh). Not a public header.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include "../../include/rt_abi.h"

namespace rt {

// Packet-kernel sphere clusters: every kPkClusterSize consecutive sphere slots share one
// cluster record; built from kPkClusterMinSpheres spheres on (rt_abi.cu stage_scene_packet).
#ifndef RT_PK_CLUSTER
#define RT_PK_CLUSTER 16
#endif
constexpr int kPkClusterSize = RT_PK_CLUSTER;   // (dev A/B: -DRT_PK_CLUSTER=8)
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
void post_release(void* state);

}  // namespace rt
