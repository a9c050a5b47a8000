// rt_frame.cu — include/rt_frame.h: one whole iteration of ForwardShadingPipline::Render
// (/root/reference/src/ForwardShadingPipeline.cpp:139-272) with every intermediate image resident
// in HBM. Built on the library's own C ABI (rt_render, rt_post_chain, rt_upload_scene): this file
// adds the image ownership of ForwardShadingPipline::Init (:51-137), the history ping-pong of
// :141/:236/:249 and the presentation step (default-framebuffer conversion + read-back).
#include <nvtx3/nvToolsExt.h>
#include <cuda_runtime.h>
#include <cstring>
#include <string>

#include "../../include/rt_frame.h"
#include "rt_device.h"

struct rt_frame {
  rt_context* ctx = nullptr;
  int width = 0, height = 0;
  float* color = nullptr;       // outputTex     RGBA32F (:57-65)
  float* position = nullptr;    // gPositionTex  RGBA32F (:115-120)
  void* normal = nullptr;       // gNormalTex    RGBA16F (:121-126)
  float* display = nullptr;     // the window's framebuffer before its 8-bit conversion
  float* history[2] = {nullptr, nullptr};   // historyTex[2] RGBA32F (:86-101)
  unsigned char* display8 = nullptr;        // RGBA8 staging of rt_frame_read_display
  int last_history = 0;
  cudaStream_t stream = nullptr;            // rt_frame_render_host
  uint64_t launches = 0;
};

namespace {

#define FRAME_CUDA(f, call)                                                                                  \
  do {                                                                                                       \
    cudaError_t e_ = (call);                                                                                 \
    if (e_ != cudaSuccess)                                                                                   \
      return rt::ctx_fail((f)->ctx, e_ == cudaErrorMemoryAllocation ? RT_ERR_OOM : RT_ERR_CUDA,             \
                          std::string(#call) + ": " + cudaGetErrorString(e_));                              \
  } while (0)

// Float colour -> 8-bit UNORM as a GL default framebuffer stores it (OpenGL 4.3 section 2.3.4.2 / 17.3.9:
// clamp to [0,1], scale by 255, round to nearest). One thread per pixel: 16 B in, 4 B out.
__global__ void __launch_bounds__(256) display_rgba8_kernel(const float4* __restrict__ src, uchar4* __restrict__ dst, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 v = src[i];
  auto q = [](float x) -> unsigned char {
    x = fminf(fmaxf(x, 0.0f), 1.0f);                     // fmaxf(NaN, 0) = 0
    return (unsigned char)__float2uint_rn(x * 255.0f);
  };
  dst[i] = make_uchar4(q(v.x), q(v.y), q(v.z), q(v.w));
}

void release(rt_frame* f) {
  if (!f) return;
  cudaSetDevice(rt::ctx_device(f->ctx));
  cudaFree(f->color); cudaFree(f->position); cudaFree(f->normal); cudaFree(f->display);
  cudaFree(f->history[0]); cudaFree(f->history[1]); cudaFree(f->display8);
  if (f->stream) cudaStreamDestroy(f->stream);
  delete f;
}

rt_status check_params(rt_frame* f, const rt_frame_params* rp, const rt_post_params* pp, const char* who) {
  if (!rp || !pp) return rt::ctx_fail(f->ctx, RT_ERR_INVALID_ARG, std::string(who) + ": null params");
  if (rp->width != f->width || rp->height != f->height || pp->width != f->width || pp->height != f->height)
    return rt::ctx_fail(f->ctx, RT_ERR_INVALID_ARG, std::string(who) + ": params size differs from the frame's");
  if (rp->row_begin != 0 || rp->row_end != f->height || rp->row_band != 0)
    return rt::ctx_fail(f->ctx, RT_ERR_INVALID_ARG, std::string(who) + ": a frame renders every row (the post chain reads the whole image)");
  if (rp->flags & RT_FLAG_NORMAL_FP32)
    return rt::ctx_fail(f->ctx, RT_ERR_INVALID_ARG, std::string(who) + ": gNormalTex is RGBA16F (RT_FLAG_NORMAL_FP32 not allowed)");
  return RT_OK;
}

}  // namespace

extern "C" {

rt_status rt_frame_create(rt_context* ctx, int width, int height, rt_frame** out) {
  if (!ctx || !out) return RT_ERR_INVALID_ARG;
  *out = nullptr;
  if (width <= 0 || height <= 0) return rt::ctx_fail(ctx, RT_ERR_INVALID_ARG, "rt_frame_create: image size");
  rt_frame* f = new rt_frame;
  f->ctx = ctx; f->width = width; f->height = height;
  const size_t px = (size_t)width * height;
  cudaError_t e = cudaSetDevice(rt::ctx_device(ctx));
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->color, px * 16);
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->position, px * 16);
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->normal, px * 8);
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->display, px * 16);
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->history[0], px * 16);
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->history[1], px * 16);
  if (e == cudaSuccess) e = cudaMalloc((void**)&f->display8, px * 4);
  if (e == cudaSuccess) e = cudaMemset(f->history[0], 0, px * 16);
  if (e == cudaSuccess) e = cudaMemset(f->history[1], 0, px * 16);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) {
    const rt_status st = rt::ctx_fail(ctx, e == cudaErrorMemoryAllocation ? RT_ERR_OOM : RT_ERR_CUDA,
                                      std::string("rt_frame_create: ") + cudaGetErrorString(e));
    cudaGetLastError();
    release(f);
    return st;
  }
  *out = f;
  return RT_OK;
}

void rt_frame_destroy(rt_frame* f) { release(f); }

rt_status rt_frame_render(rt_frame* f, const rt_frame_params* rp, const rt_post_params* pp, void* cuda_stream) {
  nvtxRangePushA("rt_frame_render");
  struct Pop { ~Pop() { nvtxRangePop(); } } pop;
  if (!f) return RT_ERR_INVALID_ARG;
  rt_status st = check_params(f, rp, pp, "rt_frame_render");
  if (st != RT_OK) return st;
  st = rt_render(f->ctx, rp, f->color, f->position, f->normal, nullptr, cuda_stream);      // :153-180
  if (st != RT_OK) return st;
  const int cur = ((rp->frame_count % 2) + 2) % 2;                                           // :141
  st = rt_post_chain(f->ctx, pp, f->color, f->normal, f->history[1 - cur], f->history[cur], f->display, cuda_stream);   // :185-254
  if (st == RT_OK && pp->taa_enabled) f->last_history = cur;
  return st;
}

rt_status rt_frame_images(rt_frame* f, float** d_color, float** d_position, void** d_normal, float** d_display, float** d_history) {
  if (!f) return RT_ERR_INVALID_ARG;
  if (d_color) *d_color = f->color;
  if (d_position) *d_position = f->position;
  if (d_normal) *d_normal = f->normal;
  if (d_display) *d_display = f->display;
  if (d_history) *d_history = f->history[f->last_history];
  return RT_OK;
}

rt_status rt_frame_read_display(rt_frame* f, void* host_dst, int format, void* cuda_stream) {
  if (!f) return RT_ERR_INVALID_ARG;
  if (!host_dst) return rt::ctx_fail(f->ctx, RT_ERR_INVALID_ARG, "rt_frame_read_display: null destination");
  const cudaStream_t s = (cudaStream_t)cuda_stream;
  const size_t px = (size_t)f->width * f->height;
  FRAME_CUDA(f, cudaSetDevice(rt::ctx_device(f->ctx)));
  if (format == RT_DISPLAY_RGBA32F) {
    FRAME_CUDA(f, cudaMemcpyAsync(host_dst, f->display, px * 16, cudaMemcpyDeviceToHost, s));
  } else if (format == RT_DISPLAY_RGBA8) {
    display_rgba8_kernel<<<(unsigned)((px + 255) / 256), 256, 0, s>>>(reinterpret_cast<const float4*>(f->display),
                                                                       reinterpret_cast<uchar4*>(f->display8), px);
    FRAME_CUDA(f, cudaGetLastError());
    f->launches += 1;
    FRAME_CUDA(f, cudaMemcpyAsync(host_dst, f->display8, px * 4, cudaMemcpyDeviceToHost, s));
  } else {
    return rt::ctx_fail(f->ctx, RT_ERR_INVALID_ARG, "rt_frame_read_display: unknown format");
  }
  FRAME_CUDA(f, cudaStreamSynchronize(s));
  return RT_OK;
}

rt_status rt_frame_render_host(rt_frame* f, const rt_frame_params* rp, const rt_post_params* pp, const void* objects, int n,
                               const void* lights, int m, void* host_display, int format) {
  if (!f) return RT_ERR_INVALID_ARG;
  rt_status st = check_params(f, rp, pp, "rt_frame_render_host");
  if (st != RT_OK) return st;
  if (objects || lights) {                               // ssbo.update() + lightSSBO.update(), every frame
    st = rt_upload_scene(f->ctx, objects, n, lights, m, f->stream, 1);
    if (st != RT_OK) return st;
  }
  st = rt_frame_render(f, rp, pp, f->stream);
  if (st != RT_OK) return st;
  return rt_frame_read_display(f, host_display, format, f->stream);
}

uint64_t rt_frame_launch_count(rt_frame* f) { return f ? f->launches : 0; }

}  // extern "C"
