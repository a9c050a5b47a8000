// rt_post.cu — the stages either side of the ray-tracing pass (include/rt_post.h): AABB
// generation (f1), Radiance RGBE decode + equirectangular -> cubemap (f2), and the post chain
// bloom extract / 9-tap blur / combine / TAA resolve (f3). sm_100a, compiled with -fmad=false:
// every a*b+c below is an unfused multiply and add, as in the shaders' own order of operations.
//
// These are bandwidth-shaped kernels (a few dozen flops per 16-byte texel): one thread per
// texel (or per short run of texels), 128-bit / 64-bit coalesced accesses along x, read-only
// data through the non-coherent path so the 9-tap and 3x3 re-reads are served by L1/L2.
#include <nvtx3/nvToolsExt.h>
#include <cstdint>
#include <string>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include "rt_device.h"
#include "../../include/rt_post.h"

namespace rt_post {

struct PostState {
  void* bloom[2] = {nullptr, nullptr};   // bloomTextures[2], RGBA16F (src/ForwardShadingPipeline.cpp:67-84)
  size_t bloom_pixels = 0;
  uint64_t launches = 0;
  bool pair_attr_set = false;            // opt-in shared memory size of blur_pair_kernel, per device
  cudaStream_t aux = nullptr;            // TAA runs beside the bloom pairs (both only read the colour image)
  cudaEvent_t fork = nullptr, join = nullptr;
};

PostState* state_of(rt_context* c) {
  void** slot = rt::ctx_post_slot(c);
  if (!*slot) *slot = new PostState();
  return static_cast<PostState*>(*slot);
}

#define POST_CUDA(ctx, call)                                                                          \
  do {                                                                                                \
    cudaError_t e__ = (call);                                                                         \
    if (e__ != cudaSuccess)                                                                           \
      return rt::ctx_fail((ctx), e__ == cudaErrorMemoryAllocation ? RT_ERR_OOM : RT_ERR_CUDA,         \
                          std::string(#call) + ": " + cudaGetErrorString(e__));                       \
  } while (0)

// ---------------------------------------------------------------- texel access
struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk(float x, float y, float z) { f3 v; v.x = x; v.y = y; v.z = z; return v; }

// RGBA16F texel -> rgb as FP32 (exact)
__device__ __forceinline__ f3 ld_half_rgb(const uint2* img, size_t at) {
  const uint2 q = __ldg(img + at);
  const float2 rg = __half22float2(*reinterpret_cast<const __half2*>(&q.x));
  const float2 ba = __half22float2(*reinterpret_cast<const __half2*>(&q.y));
  return mk(rg.x, rg.y, ba.x);
}
__device__ __forceinline__ f3 half_rgb(uint2 q) {
  const float2 rg = __half22float2(*reinterpret_cast<const __half2*>(&q.x));
  const float2 ba = __half22float2(*reinterpret_cast<const __half2*>(&q.y));
  return mk(rg.x, rg.y, ba.x);
}
// (r, g, b, 1) -> RGBA16F, round to nearest even
__device__ __forceinline__ uint2 pack_half_rgb1(f3 v) {
  const __half2 rg = __floats2half2_rn(v.x, v.y), ba = __floats2half2_rn(v.z, 1.0f);
  uint2 q;
  q.x = *reinterpret_cast<const unsigned*>(&rg);
  q.y = *reinterpret_cast<const unsigned*>(&ba);
  return q;
}
__device__ __forceinline__ f3 ld_f32_rgb(const float4* img, size_t at) {
  const float4 q = __ldg(img + at);
  return mk(q.x, q.y, q.z);
}
__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
__device__ __forceinline__ int wrap_repeat(int i, int n) { int m = i % n; return m < 0 ? m + n : m; }
// GLSL min/max as the oracle's shim defines them (a NaN operand is ignored) == fminf/fmaxf
__device__ __forceinline__ f3 min3v(f3 a, f3 b) { return mk(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
__device__ __forceinline__ f3 max3v(f3 a, f3 b) { return mk(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }

// GL_LINEAR weight: the fractional position snapped to 8 fractional bits (what GL hardware's
// fixed-point filter does, and what makes a fetch at a texel centre return the texel itself).
__device__ __forceinline__ float snap_weight(float a) { return rintf(a * 256.0f) * (1.0f / 256.0f); }
__device__ __forceinline__ float lerp_tap(float t0, float t1, float a) {
  if (a == 0.0f) return t0;
  if (a == 1.0f) return t1;
  return t0 * (1.0f - a) + t1 * a;
}
__device__ __forceinline__ f3 lerp_tap(f3 t0, f3 t1, float a) {
  if (a == 0.0f) return t0;
  if (a == 1.0f) return t1;
  const float ia = 1.0f - a;
  return mk(t0.x * ia + t1.x * a, t0.y * ia + t1.y * a, t0.z * ia + t1.z * a);
}

// ---------------------------------------------------------------- f1: AABBs
__device__ __forceinline__ f3 cross3(f3 a, f3 b) { return mk(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
__device__ __forceinline__ float dot3(f3 a, f3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ f3 normalize3(f3 a) { const float inv = 1.0f / sqrtf(dot3(a, a)); return mk(a.x * inv, a.y * inv, a.z * inv); }

// One thread per Object record (176 B): reads type / position+radius / normal / size (the first
// 64 bytes = two sectors), writes min.xyz and max.xyz (bytes 144..171).
__global__ void __launch_bounds__(256) aabb_kernel(unsigned char* objects, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  unsigned char* rec = objects + (size_t)i * RT_OBJECT_STRIDE;
  const int type = *reinterpret_cast<const int*>(rec);
  const float4 pr = *reinterpret_cast<const float4*>(rec + 16);       // position, radius
  const f3 pos = mk(pr.x, pr.y, pr.z);
  f3 lo, hi;
  if (type == 0) {
    lo = mk(pos.x - pr.w, pos.y - pr.w, pos.z - pr.w);
    hi = mk(pos.x + pr.w, pos.y + pr.w, pos.z + pr.w);
  } else if (type == 1) {
    const float4 nq = *reinterpret_cast<const float4*>(rec + 32);
    const float2 size = *reinterpret_cast<const float2*>(rec + 48);
    const f3 nrm = mk(nq.x, nq.y, nq.z);
    f3 right, forward;
    // fabsf(y) > 0.9f decides exactly like the reference's double compare: 0.9f < 0.9 < nextafter(0.9f)
    if (fabsf(nrm.y) > 0.9f) { right = mk(1.0f, 0.0f, 0.0f); forward = mk(0.0f, 0.0f, 1.0f); }
    else { right = normalize3(cross3(nrm, mk(0.0f, 1.0f, 0.0f))); forward = normalize3(cross3(right, nrm)); }
    const float hx = size.x / 2.0f, hy = size.y / 2.0f;
    const f3 ax = mk(right.x * hx, right.y * hx, right.z * hx), ay = mk(forward.x * hy, forward.y * hy, forward.z * hy);
    const f3 shift = mk(nrm.x * 0.01f, nrm.y * 0.01f, nrm.z * 0.01f);
    lo = mk(((pos.x - ax.x) - ay.x) + shift.x, ((pos.y - ax.y) - ay.y) + shift.y, ((pos.z - ax.z) - ay.z) + shift.z);
    hi = mk(((pos.x + ax.x) + ay.x) + shift.x, ((pos.y + ax.y) + ay.y) + shift.y, ((pos.z + ax.z) + ay.z) + shift.z);
  } else {
    return;
  }
  float* b = reinterpret_cast<float*>(rec + 144);
  b[0] = lo.x; b[1] = lo.y; b[2] = lo.z;
  b[4] = hi.x; b[5] = hi.y; b[6] = hi.z;
}

// ---------------------------------------------------------------- f2: RGBE decode, equirect -> cube
__global__ void __launch_bounds__(256) rgbe_kernel(const uchar4* __restrict__ rgbe, float* __restrict__ rgb, size_t texels) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= texels) return;
  const uchar4 q = __ldg(rgbe + i);
  float r = 0.0f, g = 0.0f, b = 0.0f;
  if (q.w != 0) {
    const float f = ldexpf(1.0f, (int)q.w - 136);                      // exact power of two
    r = (float)q.x * f; g = (float)q.y * f; b = (float)q.z * f;
  }
  rgb[3 * i + 0] = r; rgb[3 * i + 1] = g; rgb[3 * i + 2] = b;
}

__device__ __forceinline__ f3 ld_equirect_half(const float* img, int w, int x, int y) {
  const float* t = img + ((size_t)y * w + x) * 3;                       // the GL_RGB16F upload rounds to half
  return mk(__half2float(__float2half_rn(__ldg(t))), __half2float(__float2half_rn(__ldg(t + 1))),
            __half2float(__float2half_rn(__ldg(t + 2))));
}

// grid (S/32, S/8, 6), one thread per cube texel.
__global__ void __launch_bounds__(256) equirect_kernel(const float* __restrict__ src, int w, int h, uint2* __restrict__ faces, int S) {
  const int i = blockIdx.x * 32 + threadIdx.x, j = blockIdx.y * 8 + threadIdx.y, face = blockIdx.z;
  if (i >= S || j >= S) return;
  const float s = 2.0f * (((float)i + 0.5f) / (float)S) - 1.0f;
  const float t = 2.0f * (((float)j + 0.5f) / (float)S) - 1.0f;
  f3 p;                                                                 // localPos on the unit cube
  switch (face) {
    case 0: p = mk(1.0f, -t, -s); break;
    case 1: p = mk(-1.0f, -t, s); break;
    case 2: p = mk(s, 1.0f, t); break;
    case 3: p = mk(s, -1.0f, -t); break;
    case 4: p = mk(s, -t, 1.0f); break;
    default: p = mk(-s, -t, -1.0f); break;
  }
  const float inv = 1.0f / sqrtf((p.x * p.x + p.y * p.y) + p.z * p.z);
  const f3 v = mk(p.x * inv, p.y * inv, p.z * inv);
  const float uu = atan2f(v.z, v.x) * 0.1591f + 0.5f;                  // skyboxFs.glsl:8-13
  const float vv = asinf(v.y) * 0.3183f + 0.5f;
  const float u = uu * (float)w - 0.5f, vq = vv * (float)h - 0.5f;     // GL_LINEAR, CLAMP_TO_EDGE
  const float fu = floorf(u), fv = floorf(vq);
  const float a = snap_weight(u - fu), b = snap_weight(vq - fv);
  const int i0 = clampi((int)fu, 0, w - 1), i1 = clampi((int)fu + 1, 0, w - 1);
  const int j0 = clampi((int)fv, 0, h - 1), j1 = clampi((int)fv + 1, 0, h - 1);
  const f3 top = lerp_tap(ld_equirect_half(src, w, i0, j0), ld_equirect_half(src, w, i1, j0), a);
  const f3 bot = lerp_tap(ld_equirect_half(src, w, i0, j1), ld_equirect_half(src, w, i1, j1), a);
  faces[((size_t)face * S + j) * S + i] = pack_half_rgb1(lerp_tap(top, bot, b));
}

// ---------------------------------------------------------------- f3: bloom
// brightness_extractFS.glsl. 2 texels per thread: 2 x LDG.128 in, 1 x STG.128 out.
__global__ void __launch_bounds__(256) extract_kernel(const float4* __restrict__ color, uint4* __restrict__ bright, size_t pairs,
                                                      size_t texels, float threshold) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= pairs) return;
  uint2 out[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const size_t at = 2 * i + k;
    f3 c = mk(0.0f, 0.0f, 0.0f);
    if (at < texels) c = ld_f32_rgb(color, at);
    const float brightness = (c.x * 0.2126f + c.y * 0.7152f) + c.z * 0.0722f;
    out[k] = pack_half_rgb1(brightness > threshold ? c : mk(0.0f, 0.0f, 0.0f));
  }
  if (2 * i + 1 < texels) bright[i] = make_uint4(out[0].x, out[0].y, out[1].x, out[1].y);
  else reinterpret_cast<uint2*>(bright)[2 * i] = out[0];
}

__constant__ float c_blur_w[5] = {0.227027f, 0.1945946f, 0.1216216f, 0.054054f, 0.016216f};

// gaussian_blurFs.glsl, one pass. block (32, 8), one texel per thread; the 9 taps of neighbouring
// threads overlap and are served by L1.
template <bool HORIZONTAL>
__global__ void __launch_bounds__(256) blur_kernel(const uint2* __restrict__ in, uint2* __restrict__ out, int w, int h) {
  const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (x >= w || y >= h) return;
  const size_t row = (size_t)y * w;
  const f3 c = ld_half_rgb(in, row + x);
  f3 r = mk(c.x * c_blur_w[0], c.y * c_blur_w[0], c.z * c_blur_w[0]);
#pragma unroll
  for (int i = 1; i < 5; ++i) {
    const float wi = c_blur_w[i];
    f3 a, b;
    if (HORIZONTAL) {
      a = ld_half_rgb(in, row + clampi(x + i, 0, w - 1));
      b = ld_half_rgb(in, row + clampi(x - i, 0, w - 1));
    } else {
      a = ld_half_rgb(in, (size_t)clampi(y + i, 0, h - 1) * w + x);
      b = ld_half_rgb(in, (size_t)clampi(y - i, 0, h - 1) * w + x);
    }
    r = mk(r.x + a.x * wi, r.y + a.y * wi, r.z + a.z * wi);
    r = mk(r.x + b.x * wi, r.y + b.y * wi, r.z + b.z * wi);
  }
  out[row + x] = pack_half_rgb1(r);
}

// ---------------------------------------------------------------- f3: fused blur pairs
// One kernel = one horizontal + one vertical pass (the reference's passes 2k and 2k+1) over a
// 64 x 56 output tile, optionally with the brightness extract fused in front (FIRST) and the bloom
// combine behind (LAST). The intermediate image of the pair lives in shared memory as RGBA16F —
// the same rounding the reference's RGBA16F render target applies — so the result is bit-identical
// to the two separate passes while the pair reads its input once (1.29 x with the 4-texel halo)
// and writes once. Both passes slide a 9-texel register window along the filter axis: one
// shared-memory load and one conversion per tap-centre instead of nine.
//   A: input region 72 x 64 texels (tile + 4 each side, coordinates clamped to the image: the
//      CLAMP_TO_EDGE taps then simply read their neighbours), row stride 73 texels (odd: the H pass
//      walks rows with one lane per row, conflict-free for 8-byte accesses)
//   B: H-pass result 64 x 64, row stride 65
// 1 (default): one lane per row / column, scalar FMUL + FADD. 0: two rows / columns per lane on packed f32x2 pairs (blur9x2).
// Measured on B200, 1080p chain: scalar 0.1379 ms; packed 0.1586 ms at 80 registers (292 B of spills), 0.1356 ms at 128
// registers (RT_POST_PAIR_MINB=2) — half the FP32 instructions, the same time: the pair kernel is NOT bound by FP32 issue.
// Its 600 CTAs meet 444 resident slots (1.35 waves, paid as 2): the tile shape is what to change next.
#ifndef RT_POST_SCALAR_BLUR
#define RT_POST_SCALAR_BLUR 1
#endif
constexpr int kPairW = 64, kPairH = 56, kPairR = 4;
constexpr int kPairAW = kPairW + 2 * kPairR, kPairAH = kPairH + 2 * kPairR;      // 72 x 64
constexpr int kPairAStride = kPairAW + 1, kPairBStride = kPairW + 1;
constexpr int kPairSeg = 4;                                                       // 4 x 64 = 256 threads
constexpr int kPairHRun = kPairW / kPairSeg, kPairVRun = kPairH / kPairSeg;        // 16, 14
constexpr size_t kPairSmem = (size_t)(kPairAH * kPairAStride + kPairAH * kPairBStride) * sizeof(uint2);

// Packed FP32 pairs (FMUL2 / FFMA2 of sm_100: two IEEE round-to-nearest operations per issue slot). The blur pair is bound
// by FP32 issue (51 multiply / add per output texel and pass, unfused by contract: gaussian_blurFs.glsl:12-22 rounds every
// `texel * weight` before adding it), so both passes run on pairs: a lane filters TWO rows (columns) at once, holding each
// texel's (r, g) as one pair and the two lines' b values as another.
__device__ __forceinline__ float2 pmul2(float2 a, float2 b) {
  float2 r;
  asm("{ .reg .b64 ra, rb, rc; mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5}; mul.rn.f32x2 rc, ra, rb; mov.b64 {%0, %1}, rc; }"
      : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return r;
}
// a + b on pairs, written as fma(a, 1, b): the product a * 1 is exact, so the single rounding is add.rn's. It has to be
// spelled this way because ptxas 12.9 contracts a mul.rn.f32x2 feeding an add.rn.f32x2 into ONE FFMA2 even with
// --fmad=false (it honours the explicit .rn only for scalar f32; SASS inspected) — that would drop the rounding of the
// product. The 1 comes from constant memory so that it stays a run-time value (a literal 1.0 is folded back into an add
// and fused again).
__constant__ float c_one = 1.0f;
__device__ __forceinline__ float2 padd2(float2 a, float2 b) {
  float2 r;
  const float one = c_one;
  asm("{ .reg .b64 ra, rb, rc, ro; mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5}; mov.b64 ro, {%6, %6}; fma.rn.f32x2 rc, ra, ro, rb; mov.b64 {%0, %1}, rc; }"
      : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(one));
  return r;
}
// One tap position of two lines: (r, g) of line 0, (r, g) of line 1, (b of line 0, b of line 1).
struct Tap2 { float2 rg0, rg1, bb; };
__device__ __forceinline__ Tap2 tap2(uint2 q0, uint2 q1) {
  Tap2 t;
  t.rg0 = __half22float2(*reinterpret_cast<const __half2*>(&q0.x));
  t.rg1 = __half22float2(*reinterpret_cast<const __half2*>(&q1.x));
  t.bb = make_float2(__low2float(*reinterpret_cast<const __half2*>(&q0.y)), __low2float(*reinterpret_cast<const __half2*>(&q1.y)));
  return t;
}
// blur9 of both lines: centre, then (+i, -i) for i = 1..4 — the shader's order, per channel.
__device__ __forceinline__ void blur9x2(const Tap2 (&win)[9], f3& out0, f3& out1) {
  const float2 w0 = make_float2(c_blur_w[0], c_blur_w[0]);
  float2 a0 = pmul2(win[4].rg0, w0), a1 = pmul2(win[4].rg1, w0), ab = pmul2(win[4].bb, w0);
#pragma unroll
  for (int i = 1; i < 5; ++i) {
    const float2 wi = make_float2(c_blur_w[i], c_blur_w[i]);
    a0 = padd2(a0, pmul2(win[4 + i].rg0, wi)); a1 = padd2(a1, pmul2(win[4 + i].rg1, wi)); ab = padd2(ab, pmul2(win[4 + i].bb, wi));
    a0 = padd2(a0, pmul2(win[4 - i].rg0, wi)); a1 = padd2(a1, pmul2(win[4 - i].rg1, wi)); ab = padd2(ab, pmul2(win[4 - i].bb, wi));
  }
  out0 = mk(a0.x, a0.y, ab.x);
  out1 = mk(a1.x, a1.y, ab.y);
}

// centre, then (+i, -i) for i = 1..4: the shader's order. win[4] is the centre tap.
__device__ __forceinline__ f3 blur9(const f3 (&win)[9]) {
  f3 r = mk(win[4].x * c_blur_w[0], win[4].y * c_blur_w[0], win[4].z * c_blur_w[0]);
#pragma unroll
  for (int i = 1; i < 5; ++i) {
    const float wi = c_blur_w[i];
    r = mk(r.x + win[4 + i].x * wi, r.y + win[4 + i].y * wi, r.z + win[4 + i].z * wi);
    r = mk(r.x + win[4 - i].x * wi, r.y + win[4 - i].y * wi, r.z + win[4 - i].z * wi);
  }
  return r;
}

template <bool FIRST, bool LAST>
#ifndef RT_POST_PAIR_MINB
#define RT_POST_PAIR_MINB 3
#endif
__global__ void __launch_bounds__(256, RT_POST_PAIR_MINB) blur_pair_kernel(const void* __restrict__ in, const float4* __restrict__ color,
                                                        void* __restrict__ out, int w, int h, float threshold, float strength) {
  extern __shared__ uint2 pair_smem[];
  uint2* A = pair_smem;
  uint2* B = pair_smem + kPairAH * kPairAStride;
  const int x0 = blockIdx.x * kPairW, y0 = blockIdx.y * kPairH;
  // ---- stage the input region: 18 texels per thread. Later pairs copy RGBA16F texels straight
  // into shared memory with cp.async (all 18 in flight at once, no registers); the first pair
  // converts RGBA32F colour through the brightness extract, in register batches of 6.
  constexpr int kStageIters = kPairAW * kPairAH / 256, kStageBatch = 6;
  static_assert(kPairAW * kPairAH % 256 == 0 && kStageIters % kStageBatch == 0, "staging loop shape");
  if (!FIRST) {
#pragma unroll
    for (int it = 0; it < kStageIters; ++it) {
      const int idx = it * 256 + threadIdx.x;
      const int ry = idx / kPairAW, rx = idx - ry * kPairAW;
      const size_t at = (size_t)clampi(y0 - kPairR + ry, 0, h - 1) * w + clampi(x0 - kPairR + rx, 0, w - 1);
      const unsigned dst = (unsigned)__cvta_generic_to_shared(A + ry * kPairAStride + rx);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(static_cast<const uint2*>(in) + at) : "memory");
    }
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
  } else {
#pragma unroll 1
    for (int it = 0; it < kStageIters; it += kStageBatch) {
      float4 cq[kStageBatch];
#pragma unroll
      for (int b = 0; b < kStageBatch; ++b) {
        const int idx = (it + b) * 256 + threadIdx.x;
        const int ry = idx / kPairAW, rx = idx - ry * kPairAW;
        const size_t at = (size_t)clampi(y0 - kPairR + ry, 0, h - 1) * w + clampi(x0 - kPairR + rx, 0, w - 1);
        cq[b] = __ldg(static_cast<const float4*>(in) + at);
      }
#pragma unroll
      for (int b = 0; b < kStageBatch; ++b) {
        const int idx = (it + b) * 256 + threadIdx.x;
        const int ry = idx / kPairAW, rx = idx - ry * kPairAW;
        const f3 c = mk(cq[b].x, cq[b].y, cq[b].z);
        const float brightness = (c.x * 0.2126f + c.y * 0.7152f) + c.z * 0.0722f;
        A[ry * kPairAStride + rx] = pack_half_rgb1(brightness > threshold ? c : mk(0.0f, 0.0f, 0.0f));
      }
    }
  }
  __syncthreads();
#if RT_POST_SCALAR_BLUR
  const int lane_item = threadIdx.x & 63, seg = threadIdx.x >> 6;
  // ---- horizontal pass: one lane per row of A, a run of 16 outputs
  {
    const uint2* row = A + lane_item * kPairAStride + seg * kPairHRun;
    uint2* dst = B + lane_item * kPairBStride + seg * kPairHRun;
    f3 win[9];
#pragma unroll
    for (int k = 0; k < 8; ++k) win[k + 1] = half_rgb(row[k]);
#pragma unroll
    for (int c = 0; c < kPairHRun; ++c) {
#pragma unroll
      for (int k = 0; k < 8; ++k) win[k] = win[k + 1];
      win[8] = half_rgb(row[c + 8]);
      dst[c] = pack_half_rgb1(blur9(win));
    }
  }
  __syncthreads();
  // ---- vertical pass: one lane per column of B, a run of 14 outputs
  {
    const int gx = x0 + lane_item;
    const uint2* col = B + (seg * kPairVRun) * kPairBStride + lane_item;
    f3 win[9];
#pragma unroll
    for (int k = 0; k < 8; ++k) win[k + 1] = half_rgb(col[k * kPairBStride]);
#pragma unroll
    for (int r = 0; r < kPairVRun; ++r) {
#pragma unroll
      for (int k = 0; k < 8; ++k) win[k] = win[k + 1];
      win[8] = half_rgb(col[(r + 8) * kPairBStride]);
      const int gy = y0 + seg * kPairVRun + r;
      const uint2 q = pack_half_rgb1(blur9(win));
      if (gx < w && gy < h) {
        const size_t at = (size_t)gy * w + gx;
        if (LAST) {                                        // bloom_combineFs.glsl on the rounded result
          const f3 s = ld_f32_rgb(color, at), b = half_rgb(q);
          static_cast<float4*>(out)[at] = make_float4(s.x + b.x * strength, s.y + b.y * strength, s.z + b.z * strength, 1.0f);
        } else {
          static_cast<uint2*>(out)[at] = q;
        }
      }
    }
  }
#else
  const int l32 = threadIdx.x & 31, seg8 = threadIdx.x >> 5;
  // ---- horizontal pass: one lane per PAIR of rows of A (l and l + 32: the odd row stride keeps the 8-byte accesses of
  // a half-warp on distinct banks), a run of 8 outputs per row, FP32 work on packed pairs (blur9x2)
  {
    constexpr int kRun = kPairW / 8;                       // 8 segments x 8 outputs
    const uint2* row0 = A + l32 * kPairAStride + seg8 * kRun;
    const uint2* row1 = row0 + 32 * kPairAStride;
    uint2* dst0 = B + l32 * kPairBStride + seg8 * kRun;
    uint2* dst1 = dst0 + 32 * kPairBStride;
    Tap2 win[9];
#pragma unroll
    for (int k = 0; k < 8; ++k) win[k + 1] = tap2(row0[k], row1[k]);
#pragma unroll
    for (int c = 0; c < kRun; ++c) {
#pragma unroll
      for (int k = 0; k < 8; ++k) win[k] = win[k + 1];
      win[8] = tap2(row0[c + 8], row1[c + 8]);
      f3 o0, o1;
      blur9x2(win, o0, o1);
      dst0[c] = pack_half_rgb1(o0);
      dst1[c] = pack_half_rgb1(o1);
    }
  }
  __syncthreads();
  // ---- vertical pass: one lane per PAIR of columns of B (l and l + 32), a run of 7 outputs per column
  {
    constexpr int kRun = kPairH / 8;                       // 8 segments x 7 outputs
    static_assert(kPairH % 8 == 0 && kPairAH == 64 && kPairW == 64, "pair kernel shape");
    const int gx0 = x0 + l32, gx1 = gx0 + 32;
    const uint2* col = B + (seg8 * kRun) * kPairBStride + l32;
    Tap2 win[9];
#pragma unroll
    for (int k = 0; k < 8; ++k) win[k + 1] = tap2(col[k * kPairBStride], col[k * kPairBStride + 32]);
#pragma unroll
    for (int r = 0; r < kRun; ++r) {
#pragma unroll
      for (int k = 0; k < 8; ++k) win[k] = win[k + 1];
      win[8] = tap2(col[(r + 8) * kPairBStride], col[(r + 8) * kPairBStride + 32]);
      const int gy = y0 + seg8 * kRun + r;
      f3 o0, o1;
      blur9x2(win, o0, o1);
      const uint2 q0 = pack_half_rgb1(o0), q1 = pack_half_rgb1(o1);
      if (gy < h) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int gx = half ? gx1 : gx0;
          const uint2 q = half ? q1 : q0;
          if (gx >= w) continue;
          const size_t at = (size_t)gy * w + gx;
          if (LAST) {                                      // bloom_combineFs.glsl on the rounded result
            const f3 s = ld_f32_rgb(color, at), b = half_rgb(q);
            static_cast<float4*>(out)[at] = make_float4(s.x + b.x * strength, s.y + b.y * strength, s.z + b.z * strength, 1.0f);
          } else {
            static_cast<uint2*>(out)[at] = q;
          }
        }
      }
    }
  }
#endif
}

// bloom_combineFs.glsl
__global__ void __launch_bounds__(256) combine_kernel(const float4* __restrict__ color, const uint2* __restrict__ bloom,
                                                      float4* __restrict__ display, size_t texels, float strength) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= texels) return;
  const f3 s = ld_f32_rgb(color, i), b = ld_half_rgb(bloom, i);
  display[i] = make_float4(s.x + b.x * strength, s.y + b.y * strength, s.z + b.z * strength, 1.0f);
}

// ---------------------------------------------------------------- f3: TAA resolve (taaFs.glsl)
__device__ __forceinline__ f3 fetch_or_zero(const float4* img, int w, int h, int x, int y) {
  if (x < 0 || y < 0 || x >= w || y >= h) return mk(0.0f, 0.0f, 0.0f);
  return ld_f32_rgb(img, (size_t)y * w + x);
}

// wrap for coordinates that are almost always already inside [0, n)
__device__ __forceinline__ int wrap_fast(int i, int n) { return ((unsigned)i < (unsigned)n) ? i : wrap_repeat(i, n); }

// block (32, 8): the 34 x 10 colour texels around the block are staged in shared memory once
// (zeros outside the image: the texelFetch rule above), so the 3x3 neighbourhood and — when the
// jitter keeps them inside the staged region, the normal case — the four bilinear texels cost no
// further global loads. Per thread: ~1.3 colour + history + 2 normal loads.
constexpr int kTaaW = 32, kTaaH = 8, kTaaTW = kTaaW + 2, kTaaTH = kTaaH + 2;
__global__ void __launch_bounds__(256) taa_kernel(const float4* __restrict__ color, const float4* __restrict__ history,
                                                  const uint2* __restrict__ normal, float4* __restrict__ out, int w, int h,
                                                  float blend_factor, float jx, float jy) {
  __shared__ float4 tile[kTaaTH * kTaaTW];
  const int x0 = blockIdx.x * kTaaW, y0 = blockIdx.y * kTaaH;
  const int tid = threadIdx.y * kTaaW + threadIdx.x;
  for (int idx = tid; idx < kTaaTW * kTaaTH; idx += 256) {
    const int ty = idx / kTaaTW, tx = idx - ty * kTaaTW;
    const int gx = x0 - 1 + tx, gy = y0 - 1 + ty;
    float4 q = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (gx >= 0 && gy >= 0 && gx < w && gy < h) q = __ldg(color + (size_t)gy * w + gx);
    tile[idx] = q;
  }
  const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
  const bool inside = x < w && y < h;
  const size_t at = (size_t)y * w + x;
  float4 hq = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
  uint2 nq = make_uint2(0u, 0u), nq2 = make_uint2(0u, 0u);
  const float W = (float)w, H = (float)h;
  const float tu = ((float)x + 0.5f) / W, tv = ((float)y + 0.5f) / H;           // TexCoords
  const float ju = tu + jx, jv = tv + jy;                                       // jitteredUV
  if (inside) {                                                                 // issue the independent loads before the barrier
    hq = __ldg(history + at);
    nq = __ldg(normal + at);
    const int nx = wrap_fast((int)floorf(ju * W), w), ny = wrap_fast((int)floorf(jv * H), h);   // GL_NEAREST, REPEAT
    nq2 = __ldg(normal + (size_t)ny * w + nx);
  }
  __syncthreads();
  if (!inside) return;
  // current: GL_LINEAR, REPEAT. The four texels come from the staged region when all of them are
  // inside it (any jitter below one texel), else from global memory.
  f3 current;
  {
    const float u = ju * W - 0.5f, v = jv * H - 0.5f;
    const float fu = floorf(u), fv = floorf(v);
    const float a = snap_weight(u - fu), b = snap_weight(v - fv);
    const int i0 = wrap_fast((int)fu, w), i1 = wrap_fast((int)fu + 1, w);
    const int j0 = wrap_fast((int)fv, h), j1 = wrap_fast((int)fv + 1, h);
    const int ti0 = i0 - x0 + 1, ti1 = i1 - x0 + 1, tj0 = j0 - y0 + 1, tj1 = j1 - y0 + 1;
    f3 t00, t10, t01, t11;
    if ((unsigned)ti0 < (unsigned)kTaaTW && (unsigned)ti1 < (unsigned)kTaaTW && (unsigned)tj0 < (unsigned)kTaaTH && (unsigned)tj1 < (unsigned)kTaaTH) {
      const float4 q00 = tile[tj0 * kTaaTW + ti0], q10 = tile[tj0 * kTaaTW + ti1];
      const float4 q01 = tile[tj1 * kTaaTW + ti0], q11 = tile[tj1 * kTaaTW + ti1];
      t00 = mk(q00.x, q00.y, q00.z); t10 = mk(q10.x, q10.y, q10.z); t01 = mk(q01.x, q01.y, q01.z); t11 = mk(q11.x, q11.y, q11.z);
    } else {
      t00 = ld_f32_rgb(color, (size_t)j0 * w + i0); t10 = ld_f32_rgb(color, (size_t)j0 * w + i1);
      t01 = ld_f32_rgb(color, (size_t)j1 * w + i0); t11 = ld_f32_rgb(color, (size_t)j1 * w + i1);
    }
    current = lerp_tap(lerp_tap(t00, t10, a), lerp_tap(t01, t11, a), b);
  }
  f3 hist = mk(hq.x, hq.y, hq.z);
  f3 lo = current, hi = current;
  {
    const float4* t0 = tile + threadIdx.y * kTaaTW + threadIdx.x;               // tile coords of (x - 1, y - 1)
#pragma unroll
    for (int dx = 0; dx < 3; ++dx)                                              // the shader's order: x outer, y inner
#pragma unroll
      for (int dy = 0; dy < 3; ++dy) {
        const float4 q = t0[dy * kTaaTW + dx];
        const f3 nb = mk(q.x, q.y, q.z);
        lo = min3v(lo, nb);
        hi = max3v(hi, nb);
      }
  }
  float blend = 0.0f;
  {
    const f3 prev = half_rgb(nq), curr = half_rgb(nq2);
    if ((prev.x * curr.x + prev.y * curr.y) + prev.z * curr.z < 0.9f) blend = blend_factor * 0.2f;
  }
  // clipAABB
  const f3 center = mk(0.5f * (hi.x + lo.x), 0.5f * (hi.y + lo.y), 0.5f * (hi.z + lo.z));
  const f3 ext = mk(0.5f * (hi.x - lo.x), 0.5f * (hi.y - lo.y), 0.5f * (hi.z - lo.z));
  f3 clip = mk(hist.x - center.x, hist.y - center.y, hist.z - center.z);
  clip = mk(fminf(fmaxf(clip.x, -ext.x), ext.x), fminf(fmaxf(clip.y, -ext.y), ext.y), fminf(fmaxf(clip.z, -ext.z), ext.z));
  hist = mk(center.x + clip.x, center.y + clip.y, center.z + clip.z);
  const float ia = 1.0f - blend;
  out[at] = make_float4(hist.x * ia + current.x * blend, hist.y * ia + current.y * blend, hist.z * ia + current.z * blend, 1.0f);
}

}  // namespace rt_post
using namespace rt_post;

// ---------------------------------------------------------------- entry points
namespace rt {
void post_release(void* state) {
  PostState* st = static_cast<PostState*>(state);
  if (!st) return;
  cudaFree(st->bloom[0]);
  cudaFree(st->bloom[1]);
  if (st->fork) cudaEventDestroy(st->fork);
  if (st->join) cudaEventDestroy(st->join);
  if (st->aux) cudaStreamDestroy(st->aux);
  delete st;
}
}  // namespace rt

namespace {
rt_status check_size(rt_context* c, const rt_post_params* p, const char* who) {
  if (!p) return rt::ctx_fail(c, RT_ERR_INVALID_ARG, std::string(who) + ": null params");
  if (p->width <= 0 || p->height <= 0) return rt::ctx_fail(c, RT_ERR_INVALID_ARG, std::string(who) + ": image size");
  return RT_OK;
}
rt_status launched(rt_context* c, const char* who) {
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return rt::ctx_fail(c, RT_ERR_CUDA, std::string(who) + ": " + cudaGetErrorString(e));
  ++state_of(c)->launches;
  return RT_OK;
}
inline dim3 grid2d(int w, int h) { return dim3((unsigned)((w + 31) / 32), (unsigned)((h + 7) / 8), 1); }
}  // namespace

extern "C" {

rt_status rt_generate_aabbs_device(rt_context* c, void* d_objects, int n, void* cuda_stream) {
  if (!c) return RT_ERR_INVALID_ARG;
  if (n < 0 || (n > 0 && !d_objects)) return rt::ctx_fail(c, RT_ERR_INVALID_ARG, "rt_generate_aabbs_device: bad count or null array");
  if (n == 0) return RT_OK;
  POST_CUDA(c, cudaSetDevice(rt::ctx_device(c)));
  aabb_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)cuda_stream>>>(static_cast<unsigned char*>(d_objects), n);
  return launched(c, "rt_generate_aabbs_device");
}

rt_status rt_generate_aabbs(rt_context* c, void* objects, int n) {
  if (!c) return RT_ERR_INVALID_ARG;
  if (n < 0 || (n > 0 && !objects)) return rt::ctx_fail(c, RT_ERR_INVALID_ARG, "rt_generate_aabbs: bad count or null array");
  if (n == 0) return RT_OK;
  POST_CUDA(c, cudaSetDevice(rt::ctx_device(c)));
  void* d = nullptr;
  const size_t bytes = (size_t)n * RT_OBJECT_STRIDE;
  POST_CUDA(c, cudaMalloc(&d, bytes));
  rt_status st = RT_OK;
  cudaError_t e = cudaMemcpy(d, objects, bytes, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    st = rt_generate_aabbs_device(c, d, n, nullptr);
    if (st == RT_OK)                                     // only the bounds travel back: 32 of every 176 bytes
      e = cudaMemcpy2D(static_cast<unsigned char*>(objects) + 144, RT_OBJECT_STRIDE, static_cast<unsigned char*>(d) + 144,
                       RT_OBJECT_STRIDE, 32, (size_t)n, cudaMemcpyDeviceToHost);
  }
  cudaFree(d);
  if (st != RT_OK) return st;
  if (e != cudaSuccess) return rt::ctx_fail(c, RT_ERR_CUDA, std::string("rt_generate_aabbs: ") + cudaGetErrorString(e));
  return RT_OK;
}

rt_status rt_decode_rgbe(rt_context* c, const unsigned char* d_rgbe, float* d_rgb32f, size_t texels, void* cuda_stream) {
  if (!c) return RT_ERR_INVALID_ARG;
  if (texels > 0 && (!d_rgbe || !d_rgb32f)) return rt::ctx_fail(c, RT_ERR_INVALID_ARG, "rt_decode_rgbe: null image");
  if (texels == 0) return RT_OK;
  POST_CUDA(c, cudaSetDevice(rt::ctx_device(c)));
  rgbe_kernel<<<(unsigned)((texels + 255) / 256), 256, 0, (cudaStream_t)cuda_stream>>>(reinterpret_cast<const uchar4*>(d_rgbe), d_rgb32f, texels);
  return launched(c, "rt_decode_rgbe");
}

rt_status rt_equirect_to_cubemap(rt_context* c, const float* d_equirect, int w, int h, void* d_faces, int face_size, void* cuda_stream) {
  if (!c) return RT_ERR_INVALID_ARG;
  if (!d_equirect || !d_faces || w <= 0 || h <= 0 || face_size <= 0) return rt::ctx_fail(c, RT_ERR_INVALID_ARG, "rt_equirect_to_cubemap: bad image");
  POST_CUDA(c, cudaSetDevice(rt::ctx_device(c)));
  const dim3 grid((unsigned)((face_size + 31) / 32), (unsigned)((face_size + 7) / 8), 6);
  equirect_kernel<<<grid, dim3(32, 8, 1), 0, (cudaStream_t)cuda_stream>>>(d_equirect, w, h, static_cast<uint2*>(d_faces), face_size);
  return launched(c, "rt_equirect_to_cubemap");
}

rt_status rt_bloom_extract(rt_context* c, const rt_post_params* p, const float* d_color, void* d_bright, void* cuda_stream) {
  if (!c) return RT_ERR_INVALID_ARG;
  rt_status st = check_size(c, p, "rt_bloom_extract");
  if (st != RT_OK) return st;
  if (!d_color || !d_bright) return rt::ctx_fail(c, RT_ERR_INVALID_ARG, "rt_bloom_extract: null image");
  POST_CUDA(c, cudaSetDevice(rt::ctx_device(c)));
  const size_t texels = (size_t)p->width * p->height, pairs = (texels + 1) / 2;
  extract_kernel<<<(unsigned)((pairs + 255) / 256), 256, 0, (cudaStream_t)cuda_stream>>>(
      reinterpret_cast<const float4*>(d_color), static_cast<uint4*>(d_bright), pairs, texels, p->bloom_threshold);
  return launched(c, "rt_bloom_extract");
}

rt_status rt_gaussian_blur(rt_context* c, const rt_post_params* p, const void* d_in, void* d_out, int horizontal, void* cuda_stream) {
  if (!c) return RT_ERR_INVALID_ARG;
  rt_status st = check_size(c, p, "rt_gaussian_blur");
  if (st != RT_OK) return st;
  if (!d_in || !d_out || d_in == d_out) return rt::ctx_fail(c, RT_ERR_INVALID_ARG, "rt_gaussian_blur: null or aliased images");
  POST_CUDA(c, cudaSetDevice(rt::ctx_device(c)));
  const dim3 grid = grid2d(p->width, p->height), block(32, 8, 1);
  if (horizontal) blur_kernel<true><<<grid, block, 0, (cudaStream_t)cuda_stream>>>(static_cast<const uint2*>(d_in), static_cast<uint2*>(d_out), p->width, p->height);
  else blur_kernel<false><<<grid, block, 0, (cudaStream_t)cuda_stream>>>(static_cast<const uint2*>(d_in), static_cast<uint2*>(d_out), p->width, p->height);
  return launched(c, "rt_gaussian_blur");
}

rt_status rt_bloom_combine(rt_context* c, const rt_post_params* p, const float* d_color, const void* d_bloom, float* d_display, void* cuda_stream) {
  if (!c) return RT_ERR_INVALID_ARG;
  rt_status st = check_size(c, p, "rt_bloom_combine");
  if (st != RT_OK) return st;
  if (!d_color || !d_bloom || !d_display) return rt::ctx_fail(c, RT_ERR_INVALID_ARG, "rt_bloom_combine: null image");
  POST_CUDA(c, cudaSetDevice(rt::ctx_device(c)));
  const size_t texels = (size_t)p->width * p->height;
  combine_kernel<<<(unsigned)((texels + 255) / 256), 256, 0, (cudaStream_t)cuda_stream>>>(
      reinterpret_cast<const float4*>(d_color), static_cast<const uint2*>(d_bloom), reinterpret_cast<float4*>(d_display), texels, p->bloom_strength);
  return launched(c, "rt_bloom_combine");
}

rt_status rt_taa_resolve(rt_context* c, const rt_post_params* p, const float* d_color, const float* d_history_prev,
                         const void* d_normal, float* d_history_out, void* cuda_stream) {
  if (!c) return RT_ERR_INVALID_ARG;
  rt_status st = check_size(c, p, "rt_taa_resolve");
  if (st != RT_OK) return st;
  if (!d_color || !d_history_prev || !d_normal || !d_history_out || d_history_prev == d_history_out)
    return rt::ctx_fail(c, RT_ERR_INVALID_ARG, "rt_taa_resolve: null or aliased images");
  POST_CUDA(c, cudaSetDevice(rt::ctx_device(c)));
  taa_kernel<<<grid2d(p->width, p->height), dim3(32, 8, 1), 0, (cudaStream_t)cuda_stream>>>(
      reinterpret_cast<const float4*>(d_color), reinterpret_cast<const float4*>(d_history_prev), static_cast<const uint2*>(d_normal),
      reinterpret_cast<float4*>(d_history_out), p->width, p->height, p->taa_blend, p->jitter_x, p->jitter_y);
  return launched(c, "rt_taa_resolve");
}

rt_status rt_post_chain(rt_context* c, const rt_post_params* p, const float* d_color, const void* d_normal,
                        const float* d_history_prev, float* d_history_out, float* d_display, void* cuda_stream) {
  if (!c) return RT_ERR_INVALID_ARG;
  nvtxRangePushA("rt_post_chain");
  struct Pop { ~Pop() { nvtxRangePop(); } } pop;
  rt_status st = check_size(c, p, "rt_post_chain");
  if (st != RT_OK) return st;
  if (!d_color || !d_display) return rt::ctx_fail(c, RT_ERR_INVALID_ARG, "rt_post_chain: null image");
  if (p->blur_passes < 0) return rt::ctx_fail(c, RT_ERR_INVALID_ARG, "rt_post_chain: blur_passes");
  POST_CUDA(c, cudaSetDevice(rt::ctx_device(c)));
  PostState* ps = state_of(c);
  const size_t texels = (size_t)p->width * p->height;
  if (ps->bloom_pixels < texels) {                       // InitBloom (:67-84)
    cudaFree(ps->bloom[0]); cudaFree(ps->bloom[1]);
    ps->bloom[0] = ps->bloom[1] = nullptr; ps->bloom_pixels = 0;
    POST_CUDA(c, cudaMalloc(&ps->bloom[0], texels * 8));
    POST_CUDA(c, cudaMalloc(&ps->bloom[1], texels * 8));
    ps->bloom_pixels = texels;
  }
  if (!(p->flags & RT_POST_LITERAL_PASSES) && p->blur_passes >= 2 && (p->blur_passes & 1) == 0) {
    // fused: extract + (H, V) pairs + combine, blur_passes / 2 launches
    const int pairs = p->blur_passes / 2;
    const dim3 grid((unsigned)((p->width + kPairW - 1) / kPairW), (unsigned)((p->height + kPairH - 1) / kPairH), 1);
    const cudaStream_t s = (cudaStream_t)cuda_stream;
    const float4* col = reinterpret_cast<const float4*>(d_color);
    if (!ps->pair_attr_set) {
      POST_CUDA(c, cudaFuncSetAttribute(blur_pair_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPairSmem));
      POST_CUDA(c, cudaFuncSetAttribute(blur_pair_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPairSmem));
      POST_CUDA(c, cudaFuncSetAttribute(blur_pair_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPairSmem));
      POST_CUDA(c, cudaFuncSetAttribute(blur_pair_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPairSmem));
      ps->pair_attr_set = true;
    }
    if (p->taa_enabled) {                                  // fork: TAA on the auxiliary stream
      if (!ps->aux) {
        POST_CUDA(c, cudaStreamCreateWithFlags(&ps->aux, cudaStreamNonBlocking));
        POST_CUDA(c, cudaEventCreateWithFlags(&ps->fork, cudaEventDisableTiming));
        POST_CUDA(c, cudaEventCreateWithFlags(&ps->join, cudaEventDisableTiming));
      }
      POST_CUDA(c, cudaEventRecord(ps->fork, s));
      POST_CUDA(c, cudaStreamWaitEvent(ps->aux, ps->fork, 0));
      st = rt_taa_resolve(c, p, d_color, d_history_prev, d_normal, d_history_out, ps->aux);
      if (st != RT_OK) return st;
      POST_CUDA(c, cudaEventRecord(ps->join, ps->aux));
    }
    for (int k = 0; k < pairs; ++k) {
      const bool first = k == 0, last = k == pairs - 1;
      const void* in = first ? static_cast<const void*>(d_color) : ps->bloom[(k - 1) & 1];
      void* out = last ? static_cast<void*>(d_display) : ps->bloom[k & 1];
      if (first && last) blur_pair_kernel<true, true><<<grid, 256, kPairSmem, s>>>(in, col, out, p->width, p->height, p->bloom_threshold, p->bloom_strength);
      else if (first) blur_pair_kernel<true, false><<<grid, 256, kPairSmem, s>>>(in, col, out, p->width, p->height, p->bloom_threshold, p->bloom_strength);
      else if (last) blur_pair_kernel<false, true><<<grid, 256, kPairSmem, s>>>(in, col, out, p->width, p->height, p->bloom_threshold, p->bloom_strength);
      else blur_pair_kernel<false, false><<<grid, 256, kPairSmem, s>>>(in, col, out, p->width, p->height, p->bloom_threshold, p->bloom_strength);
      st = launched(c, "rt_post_chain (blur pair)");
      if (st != RT_OK) return st;
    }
    if (p->taa_enabled) POST_CUDA(c, cudaStreamWaitEvent(s, ps->join, 0));   // join
    return st;
  }
  st = rt_bloom_extract(c, p, d_color, ps->bloom[0], cuda_stream);                          // :192-201
  if (st != RT_OK) return st;
  bool horizontal = true;                                                                   // :206-216
  for (int i = 0; i < p->blur_passes; ++i) {
    st = rt_gaussian_blur(c, p, ps->bloom[horizontal ? 0 : 1], ps->bloom[horizontal ? 1 : 0], horizontal ? 1 : 0, cuda_stream);
    if (st != RT_OK) return st;
    horizontal = !horizontal;
  }
  st = rt_bloom_combine(c, p, d_color, ps->bloom[horizontal ? 0 : 1], d_display, cuda_stream);   // :220-230
  if (st != RT_OK) return st;
  if (p->taa_enabled) st = rt_taa_resolve(c, p, d_color, d_history_prev, d_normal, d_history_out, cuda_stream);   // :233-254
  return st;
}

uint64_t rt_post_launch_count(rt_context* c) { return c ? state_of(c)->launches : 0; }

}  // extern "C"
