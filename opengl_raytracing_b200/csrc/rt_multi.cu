// rt_multi.cu — multi-GPU entry points (include/rt_multi.h): N rt_contexts driven by one process, and the
// peer pull-reduce kernel of the sample-range sharding — the path's ONE real exchange step —
// written over peer memory (NVLink loads / stores) instead of an NCCL call.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <nvtx3/nvToolsExt.h>
#include "rt_device.h"
#include "../../include/rt_multi.h"

namespace {

struct PeerArgs {
  const float4* partial[RT_MAX_PEERS];
  float4* result;
  unsigned* flags[RT_MAX_PEERS];
  unsigned long long begin, end;     // this rank's slice of the frame, in pixels
  float scale;
  unsigned epoch;
  int rank, world, use_flags;
};

__device__ __forceinline__ unsigned long long global_ns() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) { unsigned v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ float4 ld_volatile_f4(const float4* p) {
  float4 v; asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory"); return v;
}
constexpr unsigned long long kWaitNs = 10ull * 1000ull * 1000ull * 1000ull;   // bounded waits: a missing peer must not hang the GPU

// until *p has reached `epoch` (wrap-safe); false on timeout
__device__ bool wait_epoch(const unsigned* p, unsigned epoch) {
  const unsigned long long t0 = global_ns();
  while ((int)(ld_acquire_sys(p) - epoch) < 0) {
    __nanosleep(200);
    if (global_ns() - t0 > kWaitNs) return false;
  }
  return true;
}

// Rank r's share of the reduce: out[px] = (sum over ranks, IN RANK ORDER, of partial[rank][px].rgb) * scale, alpha 1,
// for px in its slice. use_flags: the cross-process protocol of rt_peer_reduce (arrive / wait / done through flag
// words in peer memory); without, the launch is already ordered behind every render by CUDA events (rt_multi).
// WORLD is a template parameter so that the per-rank registers are exactly WORLD float4 per pixel in flight; PX pixels
// per thread and iteration, every load issued before the first use: PX * WORLD 16-byte loads in flight per thread are
// what hides the NVLink round trip (the kernel moves (WORLD + 1) x 16 B per pixel and is bound by latency x bytes in
// flight, not by arithmetic).
template <int WORLD, int PX>
__global__ void __launch_bounds__(256) peer_reduce_kernel(const __grid_constant__ PeerArgs a) {
  __shared__ int s_bad;
  const int world = a.world;                                        // <= WORLD: the instantiation is the next power of two
  if (a.use_flags) {
    // everything this rank enqueued before the kernel (its render) is complete and visible: tell every peer
    if (blockIdx.x == 0 && threadIdx.x < world) st_release_sys(a.flags[threadIdx.x] + a.rank, a.epoch);
    if (threadIdx.x == 0) s_bad = 0;
    __syncthreads();
    if (threadIdx.x < world && !wait_epoch(a.flags[a.rank] + threadIdx.x, a.epoch)) s_bad = 1;
    __syncthreads();
    if (s_bad) { if (threadIdx.x == 0) a.flags[a.rank][33] = 1u; return; }
  }
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  for (unsigned long long px0 = a.begin + (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; px0 < a.end; px0 += PX * stride) {
    float4 v[PX][WORLD];
#pragma unroll
    for (int k = 0; k < PX; ++k) {
      const unsigned long long px = px0 + k * stride;
#pragma unroll
      for (int r = 0; r < WORLD; ++r) if (r < world && px < a.end) v[k][r] = ld_volatile_f4(a.partial[r] + px);
    }
#pragma unroll
    for (int k = 0; k < PX; ++k) {
      const unsigned long long px = px0 + k * stride;
      if (px >= a.end) break;
      float x = v[k][0].x, y = v[k][0].y, z = v[k][0].z;
#pragma unroll
      for (int r = 1; r < WORLD; ++r) if (r < world) { x = x + v[k][r].x; y = y + v[k][r].y; z = z + v[k][r].z; }   // rank order: deterministic
      a.result[px] = make_float4(x * a.scale, y * a.scale, z * a.scale, 1.0f);
    }
  }
  if (a.use_flags) {
    // bar.sync orders every thread's stores before thread 0; its system-scope fence is cumulative: ONE fence per CTA
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      unsigned* counter = a.flags[a.rank] + 32;
      if (atomicAdd(counter, 1u) == gridDim.x - 1) {           // the last CTA of this rank: the slice is in rank 0's frame
        *counter = 0u;
        __threadfence_system();
        st_release_sys(a.flags[0] + 16 + a.rank, a.epoch);
        if (a.rank == 0)
          for (int r = 0; r < world; ++r)
            if (!wait_epoch(a.flags[0] + 16 + r, a.epoch)) a.flags[0][33] = 1u;
      }
    }
  }
}

// dev knobs (read once): RT_PR_CTAS = CTAs per SM of the reduce grid, RT_PR_PX = pixels per thread and iteration (1, 2, 4)
int knob(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }

template <int WORLD>
cudaError_t launch_peer_reduce_w(const PeerArgs& a, unsigned ctas, int px, cudaStream_t stream) {
  if (px >= 4) peer_reduce_kernel<WORLD, 4><<<ctas, 256, 0, stream>>>(a);
  else if (px == 2) peer_reduce_kernel<WORLD, 2><<<ctas, 256, 0, stream>>>(a);
  else peer_reduce_kernel<WORLD, 1><<<ctas, 256, 0, stream>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_peer_reduce(const PeerArgs& a, int sm_count, cudaStream_t stream) {
  if (a.end <= a.begin && !a.use_flags) return cudaSuccess;
  static const int per_sm = knob("RT_PR_CTAS", 4), ppt = knob("RT_PR_PX", 2);
  const unsigned long long px = a.end > a.begin ? a.end - a.begin : 0;
  unsigned ctas = (unsigned)((px + 256ull * ppt - 1) / (256ull * ppt));
  // No CTA waits for another CTA of its own grid — only for the peers' flags, which each peer's CTA 0 (always in its
  // first wave) sets — so the grid need not be co-resident.
  const unsigned cap = (unsigned)(sm_count > 0 ? sm_count : 148) * (unsigned)(per_sm > 0 ? per_sm : 4);
  if (ctas > cap) ctas = cap;
  if (ctas < 1) ctas = 1;
  if (a.world <= 2) return launch_peer_reduce_w<2>(a, ctas, ppt, stream);
  if (a.world <= 4) return launch_peer_reduce_w<4>(a, ctas, ppt, stream);
  if (a.world <= 8) return launch_peer_reduce_w<8>(a, ctas, ppt, stream);
  return launch_peer_reduce_w<RT_MAX_PEERS>(a, ctas, ppt < 2 ? ppt : 2, stream);
}

void slice_of(unsigned long long pixels, int rank, int world, unsigned long long& b, unsigned long long& e) {
  const unsigned long long per = ((pixels + world - 1) / world + 31) / 32 * 32;   // whole 512-B runs
  b = per * rank < pixels ? per * rank : pixels;
  e = b + per < pixels ? b + per : pixels;
}

}  // namespace

struct rt_multi {
  int n = 0;
  int dev[RT_MAX_PEERS] = {};
  int sm_count[RT_MAX_PEERS] = {};
  rt_context* ctx[RT_MAX_PEERS] = {};
  cudaStream_t stream[RT_MAX_PEERS] = {};
  cudaEvent_t render_done[RT_MAX_PEERS] = {}, reduce_done[RT_MAX_PEERS] = {};
  float4* partial[RT_MAX_PEERS] = {};
  size_t partial_px = 0;
  float4* o_color = nullptr; float4* o_position = nullptr; void* o_normal = nullptr; int* o_id = nullptr; size_t o_px = 0;
  std::string err;
  uint64_t reduce_launches = 0;
};

namespace {
rt_status mfail(rt_multi* m, rt_status s, const std::string& msg) { if (m) m->err = msg; return s; }
rt_status mcuda(rt_multi* m, cudaError_t e, const char* what) {
  return mfail(m, e == cudaErrorMemoryAllocation ? RT_ERR_OOM : RT_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define RTM_CUDA(m, call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return mcuda((m), e__, #call); } while (0)
#define RTM_CTX(m, i, call) do { rt_status s__ = (call); if (s__ != RT_OK) return mfail((m), s__, std::string(#call) + ": " + rt_last_error((m)->ctx[i])); } while (0)
}  // namespace

extern "C" {

rt_status rt_multi_create(rt_multi** out, const int* devices, int n) {
  if (!out) return RT_ERR_INVALID_ARG;
  *out = nullptr;
  if (!devices || n < 1 || n > RT_MAX_PEERS) return RT_ERR_INVALID_ARG;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return RT_ERR_UNAVAILABLE;
  for (int i = 0; i < n; ++i) if (devices[i] < 0 || devices[i] >= count) return RT_ERR_INVALID_ARG;
  rt_multi* m = new rt_multi();
  m->n = n;
  auto bail = [&](rt_status s, const char* what) { fprintf(stderr, "rt_multi_create: %s\n", what); rt_multi_destroy(m); return s; };
  for (int i = 0; i < n; ++i) {
    m->dev[i] = devices[i];
    rt_status st = rt_create(&m->ctx[i], devices[i]);
    if (st != RT_OK) return bail(st, "rt_create failed");
    if (cudaSetDevice(devices[i]) != cudaSuccess) return bail(RT_ERR_CUDA, "cudaSetDevice");
    cudaDeviceGetAttribute(&m->sm_count[i], cudaDevAttrMultiProcessorCount, devices[i]);
    if (cudaStreamCreateWithFlags(&m->stream[i], cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&m->render_done[i], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&m->reduce_done[i], cudaEventDisableTiming) != cudaSuccess) return bail(RT_ERR_CUDA, "stream / event creation");
  }
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      if (devices[i] == devices[j]) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, devices[i], devices[j]);
      if (!can) return bail(RT_ERR_UNAVAILABLE, "the devices cannot access each other's memory (no NVLink / PCIe peer path)");
      cudaSetDevice(devices[i]);
      cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return bail(RT_ERR_CUDA, cudaGetErrorString(e));
      cudaGetLastError();
    }
  *out = m;
  return RT_OK;
}

void rt_multi_destroy(rt_multi* m) {
  if (!m) return;
  for (int i = 0; i < m->n; ++i) {
    cudaSetDevice(m->dev[i]);
    cudaDeviceSynchronize();
  }
  for (int i = 0; i < m->n; ++i) {
    cudaSetDevice(m->dev[i]);
    cudaFree(m->partial[i]);
    if (m->stream[i]) cudaStreamDestroy(m->stream[i]);
    if (m->render_done[i]) cudaEventDestroy(m->render_done[i]);
    if (m->reduce_done[i]) cudaEventDestroy(m->reduce_done[i]);
    if (m->ctx[i]) rt_destroy(m->ctx[i]);
  }
  if (m->n > 0) {
    cudaSetDevice(m->dev[0]);
    cudaFree(m->o_color); cudaFree(m->o_position); cudaFree(m->o_normal); cudaFree(m->o_id);
  }
  delete m;
}

const char* rt_multi_last_error(rt_multi* m) { return m ? m->err.c_str() : "null rt_multi"; }
int rt_multi_size(rt_multi* m) { return m ? m->n : 0; }
rt_context* rt_multi_context(rt_multi* m, int i) { return (m && i >= 0 && i < m->n) ? m->ctx[i] : nullptr; }

rt_status rt_multi_upload_scene(rt_multi* m, const void* objects, int n, const void* lights, int nl) {
  if (!m) return RT_ERR_INVALID_ARG;
  for (int i = 0; i < m->n; ++i) RTM_CTX(m, i, rt_upload_scene(m->ctx[i], objects, n, lights, nl, m->stream[i], 1));
  return RT_OK;
}
rt_status rt_multi_set_skybox(rt_multi* m, const void* faces, int face_size, int channels, int is_fp16) {
  if (!m) return RT_ERR_INVALID_ARG;
  for (int i = 0; i < m->n; ++i) RTM_CTX(m, i, rt_set_skybox(m->ctx[i], faces, face_size, channels, is_fp16));
  return RT_OK;
}
rt_status rt_multi_set_noise(rt_multi* m, const unsigned char* r8, int w, int h) {
  if (!m) return RT_ERR_INVALID_ARG;
  for (int i = 0; i < m->n; ++i) RTM_CTX(m, i, rt_set_noise(m->ctx[i], r8, w, h));
  return RT_OK;
}

rt_status rt_multi_render(rt_multi* m, const rt_frame_params* p, int sharding, float* d_color, float* d_position, void* d_normal,
                          int32_t* d_object_id) {
  if (!m || !p) return RT_ERR_INVALID_ARG;
  if (sharding != RT_MULTI_ROWS && sharding != RT_MULTI_SAMPLES) return mfail(m, RT_ERR_INVALID_ARG, "rt_multi_render: sharding");
  if (p->width <= 0 || p->height <= 0) return mfail(m, RT_ERR_INVALID_ARG, "rt_multi_render: image size");
  if (p->row_band != 0 || p->row_begin != 0 || p->row_end != p->height)
    return mfail(m, RT_ERR_INVALID_ARG, "rt_multi_render: params must describe the whole frame (the sharding sets the row window)");
  if (p->flags & (RT_FLAG_ACCUMULATE | RT_FLAG_SUM_SAMPLES | RT_FLAG_COUNTERS))
    return mfail(m, RT_ERR_INVALID_ARG, "rt_multi_render: ACCUMULATE / SUM_SAMPLES / COUNTERS are not accepted");
  nvtxRangePushA("rt_multi_render");
  struct Pop { ~Pop() { nvtxRangePop(); } } pop;
  const int n = m->n;
  const int nsamp = p->num_samples > 0 ? p->num_samples : 1;
  // the previous frame's reduce kernels read every partial: nothing of this frame may start before they are done
  for (int g = 0; g < n; ++g) {
    RTM_CUDA(m, cudaSetDevice(m->dev[g]));
    for (int h = 0; h < n; ++h) RTM_CUDA(m, cudaStreamWaitEvent(m->stream[g], m->reduce_done[h], 0));
  }
  if (sharding == RT_MULTI_ROWS || n == 1) {
    for (int g = 0; g < n; ++g) {
      RTM_CUDA(m, cudaSetDevice(m->dev[g]));
      rt_frame_params q = *p;
      if (n > 1) { q.row_band = 32; q.row_period = n; q.row_phase = g; }      // band k of the frame belongs to device k % n
      RTM_CTX(m, g, rt_render(m->ctx[g], &q, d_color, d_position, d_normal, d_object_id, m->stream[g]));
      RTM_CUDA(m, cudaEventRecord(m->render_done[g], m->stream[g]));
    }
    RTM_CUDA(m, cudaSetDevice(m->dev[0]));
    for (int g = 1; g < n; ++g) RTM_CUDA(m, cudaStreamWaitEvent(m->stream[0], m->render_done[g], 0));
    return RT_OK;
  }
  // ---- sample ranges
  const size_t px = (size_t)p->width * p->height;
  if (px > m->partial_px) {
    for (int g = 0; g < n; ++g) {
      RTM_CUDA(m, cudaSetDevice(m->dev[g]));
      RTM_CUDA(m, cudaDeviceSynchronize());
      cudaFree(m->partial[g]); m->partial[g] = nullptr;
      RTM_CUDA(m, cudaMalloc((void**)&m->partial[g], px * sizeof(float4)));
    }
    m->partial_px = px;
  }
  const int base = nsamp / n, rem = nsamp % n;
  for (int g = 0; g < n; ++g) {
    RTM_CUDA(m, cudaSetDevice(m->dev[g]));
    const int first = g * base + (g < rem ? g : rem), count = base + (g < rem ? 1 : 0);
    if (count == 0) {
      RTM_CUDA(m, cudaMemsetAsync(m->partial[g], 0, px * sizeof(float4), m->stream[g]));
    } else {
      rt_frame_params q = *p;
      q.frame_count = p->frame_count + first; q.num_samples = count; q.flags = p->flags | RT_FLAG_SUM_SAMPLES;
      // position / normal / id are those of sample 0: device 0 renders them straight into the caller's images
      RTM_CTX(m, g, rt_render(m->ctx[g], &q, (float*)m->partial[g], g == 0 ? d_position : nullptr, g == 0 ? d_normal : nullptr,
                              g == 0 ? d_object_id : nullptr, m->stream[g]));
    }
    RTM_CUDA(m, cudaEventRecord(m->render_done[g], m->stream[g]));
  }
  if (d_color) {
    PeerArgs a;
    std::memset(&a, 0, sizeof(a));
    for (int g = 0; g < n; ++g) a.partial[g] = m->partial[g];
    a.result = (float4*)d_color; a.scale = 1.0f / (float)nsamp; a.world = n; a.use_flags = 0;
    for (int g = 0; g < n; ++g) {
      RTM_CUDA(m, cudaSetDevice(m->dev[g]));
      for (int h = 0; h < n; ++h) if (h != g) RTM_CUDA(m, cudaStreamWaitEvent(m->stream[g], m->render_done[h], 0));
      a.rank = g;
      slice_of(px, g, n, a.begin, a.end);
      RTM_CUDA(m, launch_peer_reduce(a, m->sm_count[g], m->stream[g]));
      RTM_CUDA(m, cudaEventRecord(m->reduce_done[g], m->stream[g]));
      m->reduce_launches += 1;
    }
  }
  RTM_CUDA(m, cudaSetDevice(m->dev[0]));
  for (int g = 1; g < n; ++g) {
    RTM_CUDA(m, cudaStreamWaitEvent(m->stream[0], m->render_done[g], 0));
    RTM_CUDA(m, cudaStreamWaitEvent(m->stream[0], m->reduce_done[g], 0));
  }
  return RT_OK;
}

rt_status rt_multi_synchronize(rt_multi* m) {
  if (!m) return RT_ERR_INVALID_ARG;
  RTM_CUDA(m, cudaSetDevice(m->dev[0]));
  RTM_CUDA(m, cudaStreamSynchronize(m->stream[0]));      // stream 0 waits for every device's events of the frame
  return RT_OK;
}

rt_status rt_multi_render_host(rt_multi* m, const rt_frame_params* p, int sharding, const void* objects, int n, const void* lights, int nl,
                               float* h_color, float* h_position, void* h_normal, int32_t* h_object_id) {
  if (!m || !p) return RT_ERR_INVALID_ARG;
  if (p->width <= 0 || p->height <= 0) return mfail(m, RT_ERR_INVALID_ARG, "rt_multi_render_host: image size");
  if (objects || lights) {
    rt_status st = rt_multi_upload_scene(m, objects, n, lights, nl);
    if (st != RT_OK) return st;
  }
  const size_t px = (size_t)p->width * p->height;
  RTM_CUDA(m, cudaSetDevice(m->dev[0]));
  if (px > m->o_px) {
    RTM_CUDA(m, cudaDeviceSynchronize());
    cudaFree(m->o_color); cudaFree(m->o_position); cudaFree(m->o_normal); cudaFree(m->o_id);
    m->o_color = m->o_position = nullptr; m->o_normal = nullptr; m->o_id = nullptr; m->o_px = 0;
    RTM_CUDA(m, cudaMalloc((void**)&m->o_color, px * sizeof(float4)));
    RTM_CUDA(m, cudaMalloc((void**)&m->o_position, px * sizeof(float4)));
    RTM_CUDA(m, cudaMalloc(&m->o_normal, px * sizeof(float4)));
    RTM_CUDA(m, cudaMalloc((void**)&m->o_id, px * sizeof(int)));
    m->o_px = px;
  }
  rt_status st = rt_multi_render(m, p, sharding, h_color ? (float*)m->o_color : nullptr, h_position ? (float*)m->o_position : nullptr,
                                 h_normal ? m->o_normal : nullptr, h_object_id ? m->o_id : nullptr);
  if (st != RT_OK) return st;
  RTM_CUDA(m, cudaSetDevice(m->dev[0]));
  const size_t nbytes = (p->flags & RT_FLAG_NORMAL_FP32) ? sizeof(float4) : sizeof(uint2);
  if (h_color) RTM_CUDA(m, cudaMemcpyAsync(h_color, m->o_color, px * sizeof(float4), cudaMemcpyDeviceToHost, m->stream[0]));
  if (h_position) RTM_CUDA(m, cudaMemcpyAsync(h_position, m->o_position, px * sizeof(float4), cudaMemcpyDeviceToHost, m->stream[0]));
  if (h_normal) RTM_CUDA(m, cudaMemcpyAsync(h_normal, m->o_normal, px * nbytes, cudaMemcpyDeviceToHost, m->stream[0]));
  if (h_object_id) RTM_CUDA(m, cudaMemcpyAsync(h_object_id, m->o_id, px * sizeof(int), cudaMemcpyDeviceToHost, m->stream[0]));
  RTM_CUDA(m, cudaStreamSynchronize(m->stream[0]));
  return RT_OK;
}

uint64_t rt_multi_launch_count(rt_multi* m) {
  if (!m) return 0;
  uint64_t total = m->reduce_launches;
  for (int i = 0; i < m->n; ++i) total += rt_launch_count(m->ctx[i]);
  return total;
}

rt_status rt_peer_reduce(rt_context* ctx, const rt_peer_reduce_desc* d, void* cuda_stream) {
  if (!ctx || !d) return RT_ERR_INVALID_ARG;
  if (d->world < 1 || d->world > RT_MAX_PEERS || d->rank < 0 || d->rank >= d->world || !d->result || d->epoch == 0)
    return rt::ctx_fail(ctx, RT_ERR_INVALID_ARG, "rt_peer_reduce: rank / world / result / epoch");
  for (int r = 0; r < d->world; ++r)
    if (!d->partial[r] || !d->flags[r]) return rt::ctx_fail(ctx, RT_ERR_INVALID_ARG, "rt_peer_reduce: null partial / flags pointer");
  cudaError_t e = cudaSetDevice(rt::ctx_device(ctx));
  if (e != cudaSuccess) return rt::ctx_fail(ctx, RT_ERR_CUDA, cudaGetErrorString(e));
  PeerArgs a;
  std::memset(&a, 0, sizeof(a));
  for (int r = 0; r < d->world; ++r) { a.partial[r] = (const float4*)d->partial[r]; a.flags[r] = d->flags[r]; }
  a.result = (float4*)d->result; a.scale = d->scale; a.epoch = d->epoch; a.rank = d->rank; a.world = d->world; a.use_flags = 1;
  slice_of(d->pixels, d->rank, d->world, a.begin, a.end);
  int sm = 0;
  cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, rt::ctx_device(ctx));
  nvtxRangePushA("rt_peer_reduce");
  e = launch_peer_reduce(a, sm, (cudaStream_t)cuda_stream);
  nvtxRangePop();
  if (e != cudaSuccess) return rt::ctx_fail(ctx, RT_ERR_CUDA, std::string("rt_peer_reduce launch: ") + cudaGetErrorString(e));
  rt::ctx_count_launch(ctx);
  return RT_OK;
}

rt_status rt_peer_reduce_status(rt_context* ctx, const uint32_t* own_flags) {
  if (!ctx || !own_flags) return RT_ERR_INVALID_ARG;
  unsigned v = 0;
  cudaError_t e = cudaSetDevice(rt::ctx_device(ctx));
  if (e == cudaSuccess) e = cudaMemcpy(&v, own_flags + 33, sizeof(v), cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) return rt::ctx_fail(ctx, RT_ERR_CUDA, cudaGetErrorString(e));
  return v == 0 ? RT_OK : rt::ctx_fail(ctx, RT_ERR_UNAVAILABLE, "rt_peer_reduce: a peer did not arrive within the wait bound");
}

}  // extern "C"
