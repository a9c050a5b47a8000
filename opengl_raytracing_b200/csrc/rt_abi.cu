// rt_abi.cu — host side of the C ABI (include/rt_abi.h): context, scene staging, texture
// setup, launch, host-buffer (e2e) path. The reference-side call sites each entry point
// replaces are cited in the header. No CPU rendering path exists here.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <cuda_fp16.h>
#include "rt_device.h"

namespace rt {
size_t scene_smem_bytes(int num_objects, int num_planes, int num_lights) {
  return sizeof(float4) * ((size_t)3 * num_objects + (size_t)3 * num_planes + (size_t)6 * num_lights);
}
cudaError_t measure_fp32_peak(int sm_count, cudaStream_t stream, double* tflops, double* ms);   // rt_peak.cu
}  // namespace rt

struct rt_context {
  int device = 0;
  int sm_count = 0;
  int max_smem_optin = 0;
  std::string err;
  // scene blob (device) + pinned staging (host)
  float4* d_scene = nullptr; size_t d_scene_cap = 0;
  float4* h_stage = nullptr; size_t h_stage_cap = 0;
  cudaEvent_t stage_done = nullptr; bool stage_pending = false;
  int n_obj = 0, n_planes = 0, n_lights = 0;
  size_t off_plane = 0, off_material = 0, off_lights = 0;   // in float4 units
  // textures
  uint2* d_sky = nullptr; cudaArray_t sky_array = nullptr; cudaTextureObject_t sky_tex = 0; int sky_size = 0;
  unsigned char* d_noise = nullptr; int noise_w = 0, noise_h = 0;
  // work distribution + counters
  unsigned* d_tile_counter = nullptr; unsigned tile_base = 0;
  unsigned long long* d_counters = nullptr;
  uint64_t launches = 0;
  // e2e path
  cudaStream_t stream = nullptr;
  float4* o_color = nullptr; float4* o_position = nullptr; void* o_normal = nullptr; int* o_id = nullptr;
  size_t o_pixels = 0;
  bool tables_uploaded = false;
};

namespace {

rt_status fail(rt_context* c, rt_status s, const std::string& msg) {
  if (c) c->err = msg;
  return s;
}
rt_status cuda_fail(rt_context* c, cudaError_t e, const char* what) {
  rt_status s = (e == cudaErrorMemoryAllocation) ? RT_ERR_OOM : RT_ERR_CUDA;
  return fail(c, s, std::string(what) + ": " + cudaGetErrorString(e));
}
#define RT_CUDA(ctx, call)                                         \
  do {                                                             \
    cudaError_t e__ = (call);                                      \
    if (e__ != cudaSuccess) return cuda_fail((ctx), e__, #call);   \
  } while (0)

// ---- FP32 helpers evaluated exactly as the shader would (this TU is compiled with
// -ffp-contract=off for the host compiler, so no FMA forms).
struct h3 { float x, y, z; };
inline h3 hcross(h3 a, h3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline float hdot(h3 a, h3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline h3 hnormalize(h3 a) { float inv = 1.0f / std::sqrt(hdot(a, a)); return {a.x * inv, a.y * inv, a.z * inv}; }

float halton_host(int index, int base) {           // glsl:278-288, IEEE FP32
  float result = 0.0f, f = 1.0f / (float)base;
  for (int i = index; i > 0; i /= base) { result = result + f * (float)(i % base); f = f / (float)base; }
  return result;
}

// Object[]/Light[] (std430 bytes) -> the kernel's SoA blob. Layout in rt_device.h.
void stage_scene(const rt_object* obj, int n, const rt_light* lights, int m, int n_planes, float4* out,
                 size_t off_plane, size_t off_material, size_t off_lights) {
  float4* isect = out; float4* plane = out + off_plane; float4* mat = out + off_material; float4* lt = out + off_lights;
  int slot = 0;
  for (int i = 0; i < n; ++i) {
    const rt_object& o = obj[i];
    int this_slot = 0;
    if (o.type == 1) {
      this_slot = slot++;
      const h3 nrm = {o.normal[0], o.normal[1], o.normal[2]};
      h3 right, forward;                             // frame of glsl:130-138, hoisted per object
      if (std::fabs(nrm.y) > 0.9f) right = hnormalize(hcross(nrm, h3{0.f, 0.f, 1.f}));
      else right = hnormalize(hcross(nrm, h3{0.f, 1.f, 0.f}));
      forward = hnormalize(hcross(right, nrm));
      plane[3 * this_slot + 0] = make_float4(nrm.x, nrm.y, nrm.z, o.size[0] / 2.0f);
      plane[3 * this_slot + 1] = make_float4(right.x, right.y, right.z, o.size[1] / 2.0f);
      plane[3 * this_slot + 2] = make_float4(forward.x, forward.y, forward.z, 0.f);
    }
    float type_bits, slot_bits;
    std::memcpy(&type_bits, &o.type, 4);
    std::memcpy(&slot_bits, &this_slot, 4);
    isect[3 * i + 0] = make_float4(o.bounds.min[0], o.bounds.min[1], o.bounds.min[2], type_bits);
    isect[3 * i + 1] = make_float4(o.bounds.max[0], o.bounds.max[1], o.bounds.max[2], slot_bits);
    isect[3 * i + 2] = make_float4(o.position[0], o.position[1], o.position[2], o.type == 0 ? o.radius * o.radius : 0.f);
    const rt_material& a = o.material;
    mat[4 * i + 0] = make_float4(a.albedo[0], a.albedo[1], a.albedo[2], a.metallic);
    mat[4 * i + 1] = make_float4(a.roughness, a.diffuseStrength, a.ior, a.transparency);
    mat[4 * i + 2] = make_float4(a.subsurfaceColor[0], a.subsurfaceColor[1], a.subsurfaceColor[2], a.subsurfaceScatter);
    mat[4 * i + 3] = make_float4(a.scatterDistance, 0.f, 0.f, 0.f);
  }
  (void)n_planes;
  if (m > 0) std::memcpy(lt, lights, (size_t)m * sizeof(rt_light));
}

rt_status ensure_outputs(rt_context* c, size_t pixels) {
  if (pixels <= c->o_pixels) return RT_OK;
  cudaFree(c->o_color); cudaFree(c->o_position); cudaFree(c->o_normal); cudaFree(c->o_id);
  c->o_color = c->o_position = nullptr; c->o_normal = nullptr; c->o_id = nullptr; c->o_pixels = 0;
  RT_CUDA(c, cudaMalloc(&c->o_color, pixels * sizeof(float4)));
  RT_CUDA(c, cudaMalloc(&c->o_position, pixels * sizeof(float4)));
  RT_CUDA(c, cudaMalloc(&c->o_normal, pixels * sizeof(float4)));
  RT_CUDA(c, cudaMalloc(&c->o_id, pixels * sizeof(int)));
  c->o_pixels = pixels;
  return RT_OK;
}

}  // namespace

extern "C" {

int rt_abi_version(void) { return RT_ABI_VERSION; }

const char* rt_status_string(rt_status s) {
  switch (s) {
    case RT_OK: return "RT_OK";
    case RT_ERR_INVALID_ARG: return "RT_ERR_INVALID_ARG";
    case RT_ERR_CUDA: return "RT_ERR_CUDA";
    case RT_ERR_NCCL: return "RT_ERR_NCCL";
    case RT_ERR_OOM: return "RT_ERR_OOM";
    case RT_ERR_UNAVAILABLE: return "RT_ERR_UNAVAILABLE";
  }
  return "RT_ERR_?";
}

const char* rt_last_error(rt_context* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

rt_status rt_create(rt_context** out, int cuda_device) {
  if (!out) return RT_ERR_INVALID_ARG;
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) return RT_ERR_UNAVAILABLE;   // no GPU: there is no CPU fallback
  if (cuda_device < 0 || cuda_device >= count) return RT_ERR_INVALID_ARG;
  rt_context* c = new rt_context();
  c->device = cuda_device;
  auto bail = [&](cudaError_t err) { rt_status s = cuda_fail(c, err, "rt_create"); fprintf(stderr, "rt_create: %s\n", c->err.c_str()); delete c; return s; };
  if ((e = cudaSetDevice(cuda_device)) != cudaSuccess) return bail(e);
  if ((e = cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, cuda_device)) != cudaSuccess) return bail(e);
  if ((e = cudaDeviceGetAttribute(&c->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, cuda_device)) != cudaSuccess) return bail(e);
  if ((e = cudaMalloc(&c->d_tile_counter, sizeof(unsigned))) != cudaSuccess) return bail(e);
  if ((e = cudaMemset(c->d_tile_counter, 0, sizeof(unsigned))) != cudaSuccess) return bail(e);
  if ((e = cudaMalloc(&c->d_counters, sizeof(unsigned long long) * RT_CNT_COUNT)) != cudaSuccess) return bail(e);
  if ((e = cudaMemset(c->d_counters, 0, sizeof(unsigned long long) * RT_CNT_COUNT)) != cudaSuccess) return bail(e);
  if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e);
  if ((e = cudaEventCreateWithFlags(&c->stage_done, cudaEventDisableTiming)) != cudaSuccess) return bail(e);
  float h2[64], h3t[64];
  for (int i = 0; i < 64; ++i) { h2[i] = halton_host(i, 2); h3t[i] = halton_host(i, 3); }
  if ((e = rt::upload_tables_fast(h2, h3t)) != cudaSuccess) return bail(e);
  if ((e = rt::upload_tables_parity(h2, h3t)) != cudaSuccess) return bail(e);
  *out = c;
  return RT_OK;
}

void rt_destroy(rt_context* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  if (c->sky_tex) cudaDestroyTextureObject(c->sky_tex);
  if (c->sky_array) cudaFreeArray(c->sky_array);
  cudaFree(c->d_sky); cudaFree(c->d_noise); cudaFree(c->d_scene);
  cudaFree(c->d_tile_counter); cudaFree(c->d_counters);
  cudaFree(c->o_color); cudaFree(c->o_position); cudaFree(c->o_normal); cudaFree(c->o_id);
  if (c->h_stage) cudaFreeHost(c->h_stage);
  if (c->stage_done) cudaEventDestroy(c->stage_done);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

rt_status rt_upload_scene(rt_context* c, const void* objects, int n, const void* lights, int m,
                          void* cuda_stream, int async) {
  if (!c) return RT_ERR_INVALID_ARG;
  if (n < 0 || m < 0 || (n > 0 && !objects) || (m > 0 && !lights)) return fail(c, RT_ERR_INVALID_ARG, "rt_upload_scene: bad counts or null arrays");
  RT_CUDA(c, cudaSetDevice(c->device));
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  const rt_object* obj = (const rt_object*)objects;
  int planes = 0;
  for (int i = 0; i < n; ++i) planes += (obj[i].type == 1);
  const size_t off_plane = (size_t)3 * n, off_material = off_plane + (size_t)3 * planes;
  const size_t off_lights = off_material + (size_t)4 * n, total = off_lights + (size_t)6 * m;
  const size_t need = (total > 0 ? total : 1) * sizeof(float4);
  if (c->stage_pending) { RT_CUDA(c, cudaEventSynchronize(c->stage_done)); c->stage_pending = false; }
  if (need > c->h_stage_cap) {
    if (c->h_stage) cudaFreeHost(c->h_stage);
    c->h_stage = nullptr; c->h_stage_cap = 0;
    RT_CUDA(c, cudaHostAlloc((void**)&c->h_stage, need * 2, cudaHostAllocDefault));
    c->h_stage_cap = need * 2;
  }
  if (need > c->d_scene_cap) {
    RT_CUDA(c, cudaStreamSynchronize(stream));
    cudaFree(c->d_scene); c->d_scene = nullptr; c->d_scene_cap = 0;
    RT_CUDA(c, cudaMalloc((void**)&c->d_scene, need * 2));
    c->d_scene_cap = need * 2;
  }
  stage_scene(obj, n, (const rt_light*)lights, m, planes, c->h_stage, off_plane, off_material, off_lights);
  RT_CUDA(c, cudaMemcpyAsync(c->d_scene, c->h_stage, total * sizeof(float4), cudaMemcpyHostToDevice, stream));
  c->n_obj = n; c->n_planes = planes; c->n_lights = m;
  c->off_plane = off_plane; c->off_material = off_material; c->off_lights = off_lights;
  if (async) {
    RT_CUDA(c, cudaEventRecord(c->stage_done, stream));
    c->stage_pending = true;
  } else {
    RT_CUDA(c, cudaStreamSynchronize(stream));
  }
  return RT_OK;
}

rt_status rt_set_skybox(rt_context* c, const void* faces, int face_size, int channels, int is_fp16) {
  if (!c) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  RT_CUDA(c, cudaDeviceSynchronize());
  if (c->sky_tex) { cudaDestroyTextureObject(c->sky_tex); c->sky_tex = 0; }
  if (c->sky_array) { cudaFreeArray(c->sky_array); c->sky_array = nullptr; }
  cudaFree(c->d_sky); c->d_sky = nullptr; c->sky_size = 0;
  if (!faces) return RT_OK;
  if (face_size <= 0 || face_size > 16384 || (channels != 3 && channels != 4)) return fail(c, RT_ERR_INVALID_ARG, "rt_set_skybox: face_size/channels");
  const size_t texels = (size_t)6 * face_size * face_size;
  std::vector<__half> rgba(texels * 4);
  for (size_t t = 0; t < texels; ++t) {
    for (int ch = 0; ch < 4; ++ch) {
      __half v = __float2half_rn(1.0f);
      if (ch < channels) v = is_fp16 ? ((const __half*)faces)[t * channels + ch] : __float2half_rn(((const float*)faces)[t * channels + ch]);
      rgba[t * 4 + ch] = v;
    }
  }
  RT_CUDA(c, cudaMalloc((void**)&c->d_sky, texels * sizeof(uint2)));
  RT_CUDA(c, cudaMemcpy(c->d_sky, rgba.data(), texels * sizeof(uint2), cudaMemcpyHostToDevice));
  c->sky_size = face_size;
  // Cubemap texture object for the RT_FLAG_SKYBOX_HW miss path (pure CUDA, no graphics API).
  cudaChannelFormatDesc desc = cudaCreateChannelDescHalf4();
  cudaError_t e = cudaMalloc3DArray(&c->sky_array, &desc, make_cudaExtent(face_size, face_size, 6), cudaArrayCubemap);
  if (e == cudaSuccess) {
    cudaMemcpy3DParms cp = {};
    cp.srcPtr = make_cudaPitchedPtr(rgba.data(), (size_t)face_size * sizeof(uint2), face_size, face_size);
    cp.dstArray = c->sky_array;
    cp.extent = make_cudaExtent(face_size, face_size, 6);
    cp.kind = cudaMemcpyHostToDevice;
    e = cudaMemcpy3D(&cp);
  }
  if (e == cudaSuccess) {
    cudaResourceDesc rd = {};
    rd.resType = cudaResourceTypeArray; rd.res.array.array = c->sky_array;
    cudaTextureDesc td = {};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 1;
    e = cudaCreateTextureObject(&c->sky_tex, &rd, &td, nullptr);
  }
  if (e != cudaSuccess) {   // the exact (manual) miss path does not need the texture object
    cudaGetLastError();
    c->sky_tex = 0;
    c->err = std::string("cubemap texture object unavailable: ") + cudaGetErrorString(e);
  }
  return RT_OK;
}

rt_status rt_set_noise(rt_context* c, const unsigned char* r8, int w, int h) {
  if (!c) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  RT_CUDA(c, cudaDeviceSynchronize());
  cudaFree(c->d_noise); c->d_noise = nullptr; c->noise_w = c->noise_h = 0;
  if (!r8) return RT_OK;
  if (w <= 0 || h <= 0) return fail(c, RT_ERR_INVALID_ARG, "rt_set_noise: size");
  RT_CUDA(c, cudaMalloc((void**)&c->d_noise, (size_t)w * h));
  RT_CUDA(c, cudaMemcpy(c->d_noise, r8, (size_t)w * h, cudaMemcpyHostToDevice));
  c->noise_w = w; c->noise_h = h;
  return RT_OK;
}

rt_status rt_render(rt_context* c, const rt_frame_params* p, float* d_color, float* d_position, void* d_normal,
                    int32_t* d_object_id, void* cuda_stream) {
  if (!c || !p) return RT_ERR_INVALID_ARG;
  if (p->width <= 0 || p->height <= 0 || p->row_begin < 0 || p->row_end > p->height || p->row_begin > p->row_end)
    return fail(c, RT_ERR_INVALID_ARG, "rt_render: image size / row window");
  if (p->num_objects < 0 || p->num_objects > c->n_obj || p->num_lights < 0 || p->num_lights > c->n_lights)
    return fail(c, RT_ERR_INVALID_ARG, "rt_render: num_objects/num_lights exceed the uploaded scene");
  if (p->max_ray_depth < 0 || p->num_samples < 0) return fail(c, RT_ERR_INVALID_ARG, "rt_render: depth/samples");
  if (p->use_skybox && !c->d_sky) return fail(c, RT_ERR_INVALID_ARG, "rt_render: use_skybox without rt_set_skybox");
  if (p->row_begin == p->row_end) return RT_OK;
  RT_CUDA(c, cudaSetDevice(c->device));
  rt::KernelParams kp;
  std::memset(&kp, 0, sizeof(kp));
  kp.fp = *p;
  kp.tan_fov = std::tan(p->fov_deg * 0.017453292519943295f * 0.5f);   // tan(radians(fov) * 0.5), glsl:209
  kp.aspect = (float)p->width / (float)p->height;                      // glsl:208
  rt::DeviceScene& sc = kp.sc;
  sc.isect = c->d_scene; sc.plane = c->d_scene + c->off_plane;
  sc.material = c->d_scene + c->off_material; sc.lights = c->d_scene + c->off_lights;
  sc.num_objects = p->num_objects; sc.num_planes = c->n_planes; sc.num_lights = p->num_lights;
  const size_t smem = rt::scene_smem_bytes(sc.num_objects, sc.num_planes, sc.num_lights);
  sc.scene_in_smem = smem <= (size_t)c->max_smem_optin - 2048 ? 1 : 0;
  sc.sky = c->d_sky; sc.sky_tex = (unsigned long long)c->sky_tex; sc.sky_size = c->sky_size;
  sc.noise = c->d_noise; sc.noise_w = c->noise_w; sc.noise_h = c->noise_h;
  kp.color = (float4*)d_color; kp.position = (float4*)d_position; kp.normal = d_normal; kp.object_id = d_object_id;
  kp.tile_counter = c->d_tile_counter; kp.tile_base = c->tile_base;
  kp.tiles_x = (unsigned)((p->width + 7) / 8);
  kp.tiles_y = (unsigned)((p->row_end - p->row_begin + 3) / 4);
  kp.counters = c->d_counters;
  unsigned warps = 0;
  cudaError_t e = (p->flags & RT_FLAG_PARITY_MATH) ? rt::launch_render_parity(kp, c->sm_count, (cudaStream_t)cuda_stream, &warps)
                                                   : rt::launch_render_fast(kp, c->sm_count, (cudaStream_t)cuda_stream, &warps);
  if (e != cudaSuccess) return cuda_fail(c, e, "rt_render launch");
  c->tile_base += kp.tiles_x * kp.tiles_y + warps;     // every warp over-fetches exactly once
  c->launches += 1;
  return RT_OK;
}

rt_status rt_render_host(rt_context* c, const rt_frame_params* p, const void* objects, int n, const void* lights, int m,
                         float* h_color, float* h_position, void* h_normal, int32_t* h_object_id) {
  if (!c || !p) return RT_ERR_INVALID_ARG;
  if (p->width <= 0 || p->height <= 0) return fail(c, RT_ERR_INVALID_ARG, "rt_render_host: image size");
  RT_CUDA(c, cudaSetDevice(c->device));
  rt_status st;
  if (objects || lights) {                               // ssbo.update() + lightSSBO.update(), every frame
    st = rt_upload_scene(c, objects, n, lights, m, c->stream, 1);
    if (st != RT_OK) return st;
  }
  const size_t pixels = (size_t)p->width * p->height;
  st = ensure_outputs(c, pixels);
  if (st != RT_OK) return st;
  st = rt_render(c, p, h_color ? (float*)c->o_color : nullptr, h_position ? (float*)c->o_position : nullptr,
                 h_normal ? c->o_normal : nullptr, h_object_id ? c->o_id : nullptr, c->stream);
  if (st != RT_OK) return st;
  const size_t first = (size_t)p->row_begin * p->width, count = (size_t)(p->row_end - p->row_begin) * p->width;
  const size_t nbytes = (p->flags & RT_FLAG_NORMAL_FP32) ? sizeof(float4) : sizeof(uint2);
  if (h_color) RT_CUDA(c, cudaMemcpyAsync(h_color + first * 4, c->o_color + first, count * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
  if (h_position) RT_CUDA(c, cudaMemcpyAsync(h_position + first * 4, c->o_position + first, count * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
  if (h_normal) RT_CUDA(c, cudaMemcpyAsync((char*)h_normal + first * nbytes, (char*)c->o_normal + first * nbytes, count * nbytes, cudaMemcpyDeviceToHost, c->stream));
  if (h_object_id) RT_CUDA(c, cudaMemcpyAsync(h_object_id + first, c->o_id + first, count * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  RT_CUDA(c, cudaStreamSynchronize(c->stream));
  return RT_OK;
}

rt_status rt_synchronize(rt_context* c) {
  if (!c) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  RT_CUDA(c, cudaDeviceSynchronize());
  return RT_OK;
}

rt_status rt_render_gl(rt_context* c, const rt_frame_params*, unsigned, unsigned, unsigned) {
  return fail(c, RT_ERR_UNAVAILABLE, "rt_render_gl: no OpenGL/EGL stack on this machine (CUDA-GL interop unavailable on this box)");
}

rt_status rt_get_counters(rt_context* c, uint64_t* out, int reset) {
  if (!c || !out) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  RT_CUDA(c, cudaDeviceSynchronize());
  RT_CUDA(c, cudaMemcpy(out, c->d_counters, sizeof(uint64_t) * RT_CNT_COUNT, cudaMemcpyDeviceToHost));
  if (reset) RT_CUDA(c, cudaMemset(c->d_counters, 0, sizeof(uint64_t) * RT_CNT_COUNT));
  return RT_OK;
}

rt_status rt_host_alloc(void** out, size_t bytes) {
  if (!out) return RT_ERR_INVALID_ARG;
  cudaError_t e = cudaHostAlloc(out, bytes, cudaHostAllocDefault);
  return e == cudaSuccess ? RT_OK : (e == cudaErrorMemoryAllocation ? RT_ERR_OOM : RT_ERR_CUDA);
}

rt_status rt_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? RT_OK : RT_ERR_CUDA; }

rt_status rt_measure_fp32_peak(rt_context* c, double* tflops, int* sm_count, double* ms) {
  if (!c || !tflops) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  double t = 0, m = 0;
  cudaError_t e = rt::measure_fp32_peak(c->sm_count, c->stream, &t, &m);
  if (e != cudaSuccess) return cuda_fail(c, e, "rt_measure_fp32_peak");
  c->launches += 4;
  *tflops = t;
  if (sm_count) *sm_count = c->sm_count;
  if (ms) *ms = m;
  return RT_OK;
}

uint64_t rt_launch_count(rt_context* c) { return c ? c->launches : 0; }

}  // extern "C"
