// rt_packet.cuh — the product kernel of librt_b200.so: the reference's per-pixel ray-tracing
// pass (/root/reference/shader/raytracingCs.glsl) restructured around what ncu showed bounds
// it on B200 (profiles/r01_v0_*): the ALU pipe (FMNMX/FSETP issue at half rate), a 170 KB
// instruction footprint, and 124 registers. Same results as the literal kernel
// (rt_kernel.cuh) — bit-identical in the parity build, tests/test_parity_gpu.py — from:
//
//   * sign-select slabs: every box is staged per axis as (lo,hi,hi,lo); a lane reads the
//     8-byte (near,far) pair its ray's sign picks, so min(t0,t1)/max(t0,t1) cost no
//     instruction. (near-o)*inv is the same two IEEE operations as the shader's
//     (boxMin-origin)*invDir, issued as packed FADD2/FMUL2 (f32x2). Rays with a zero,
//     denormal, infinite or NaN direction component take a generic min/max loop instead.
//   * shadow packets: the 4 PCF samples (and 4 of the 16 PCSS probes) of one shading point
//     share their origin and almost always their direction signs, so one lane traces them
//     as a 4-ray packet: box loads and (near-o) are shared, four independent dependency
//     chains hide latency at 16 warps/SM.
//   * exact work skipping: a light whose BRDF term is exactly (0,0,0) cannot change Lo
//     whatever its (finite) shadow factor is, so its shadow rays are not traced; shadow and
//     probe rays stop at the first occluder (any-hit), PCSS stops at the first blocked probe.
//   * spheres and planes are staged in separate arrays (no type branch per object); ties in t
//     are resolved to the lower original index, which is what the shader's in-order strict
//     `<` does.
//   * one copy of each trace loop in the instruction stream (PCSS probes and PCF samples go
//     through the same packet site).
#pragma once
#include "rt_math.cuh"

namespace rt {
namespace RT_VARIANT_NS {
namespace pk {

#ifndef RT_PK_WARPS
#define RT_PK_WARPS 8
#endif
constexpr int kWarpsPerCta = RT_PK_WARPS;
constexpr int kThreads = kWarpsPerCta * 32;
constexpr int kTileW = 8, kTileH = 4;            // one warp = 8x4 pixels
constexpr int kBlockTilesX = 4, kBlockTilesY = kWarpsPerCta / 4;   // coherent mode: CTA block = 32 x (4*warps/4) pixels

// The 16 PCSS probe offsets rv = halton3(i)*2-1 (glsl:408), sorted: consecutive probes lie next
// to each other on the line lightDir + rv*(1,1,1)*2s, so a packet of 4 rarely straddles a sign
// change. The order cannot change the result (only "is any probe blocked" is used, glsl:425).
// Filled by the host with the same FP32 operations.
__constant__ float c_probe_rv[16];

// ---------------------------------------------------------------- packed f32x2 (FADD2 / FMUL2)
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
  float2 r;
  asm("{ .reg .b64 ra, rb, rc; mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5}; add.rn.f32x2 rc, ra, rb; mov.b64 {%0, %1}, rc; }"
      : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return r;
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
  float2 r;
  asm("{ .reg .b64 ra, rb, rc; mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5}; mul.rn.f32x2 rc, ra, rb; mov.b64 {%0, %1}, rc; }"
      : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return r;
}
__device__ __forceinline__ float2 dup2(float v) { return make_float2(v, v); }
__device__ __forceinline__ float fmax3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }   // FMNMX3
__device__ __forceinline__ float fmin3(float a, float b, float c) { return fminf(fminf(a, b), c); }

// 1/x for shadow-ray slabs: IEEE in the parity build, MUFU.RCP in the fast build (a shadow ray
// only yields a boolean; 1 ulp of inv moves the slab boundary by 1 ulp).
#if RT_PARITY
__device__ __forceinline__ float rcp_shadow(float x) { return __frcp_rn(x); }
__device__ __forceinline__ f3 normalize_shadow(f3 a) { return normalize_ieee(a); }
#else
// .ftz: one MUFU instruction instead of MUFU + the denormal range fix-up. Operands here are
// direction components and squared lengths of regular rays (dir_regular() has already sent
// zero / denormal / infinite components down the generic path), so nothing is flushed.
#if RT_PK_NO_FTZ
__device__ __forceinline__ float rcp_shadow(float x) { float r; asm("rcp.approx.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ f3 normalize_shadow(f3 a) { return a * rsqrtf(sdot3(a, a)); }
#else
__device__ __forceinline__ float rcp_shadow(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ f3 normalize_shadow(f3 a) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(sdot3(a, a))); return a * r; }
#endif
#endif

// ---------------------------------------------------------------- shared-memory scene
// Byte layout, built on the host by stage_scene_packet() (rt_abi.cu):
//   sphere record  64 B : x(lo,hi,hi,lo) y(lo,hi,hi,lo) z(lo,hi,hi,lo) (cx,cy,cz,r)
//   plane record  128 B : the same 48-B box, (n.xyz, size.x/2) (right.xyz, size.y/2)
//                         (forward.xyz, 0) (position.xyz, original index bits)
//                         (bounding radius of the rectangle, |normal|, 0, 0)
//   sphere_index   4 B  : original object index of each sphere slot
//   cluster record 64 B : (lo,hi,hi,lo) x3 of the union of the boxes of sphere slots [16c, 16c+16),
//                         (centre, radius) of a sphere containing those spheres. Present from 64
//                         spheres on; the slots are then in Morton order of the centres.
// lo/hi are sorted per axis on the host (the shader's min/max make the test symmetric in them).
constexpr int kSphereBytes = 64, kPlaneBytes = 128;

typedef unsigned saddr;        // 32-bit shared-window address: explicit LDS, one register
struct Scene {
  saddr clu;                  // shared memory: nc cluster records (union box + bounding sphere), may be empty
  saddr sph;                  // shared memory
  saddr pln;                  // shared memory
  saddr sph_index;            // shared memory
  saddr lights;               // shared memory, raw 96-B records
  const float4* material;     // global memory, 4 float4 per ORIGINAL object index
  saddr material_smem;        // the same records inside the shared-memory blob (small scenes), 0 = not staged
  int nc, ns, np, nl;
  float maxd;
  const float4* grid;         // global memory: light grids (rt_abi.cu build_light_grids), nullptr = none
  int grid_off_f4, grid_ent_f4;
  saddr wlist;                // this WARP's shared candidate lists (cooperative cluster scan), 0 = none
};

__device__ __forceinline__ float lds32(saddr p) { float v; asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(p)); return v; }
__device__ __forceinline__ int lds32i(saddr p) { int v; asm("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(p)); return v; }
__device__ __forceinline__ float2 lds64(saddr p) {
  float2 v; asm("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(p)); return v;
}
__device__ __forceinline__ float4 lds128(saddr p) {
  float4 v; asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(p)); return v;
}

// Direction usable by the sign-select loops: every component normal, finite and non-zero
// (then inv is finite and non-zero, so no 0*inf NaN can appear and min/max are decided by sign).
__device__ __forceinline__ bool dir_regular(f3 d, float a) {
  return fmin3(fabsf(d.x), fabsf(d.y), fabsf(d.z)) > 1.0e-30f && a < 1.0e30f;   // a = dot(d,d): NaN/inf fail
}

// intersectSphere (glsl:105-118), a = dot(d,d) hoisted; sp = (centre, r).
__device__ __forceinline__ bool sphere_t(float4 sp, f3 o, f3 d, float a, float& t) {
  const f3 oc = o - xyz(sp);
  const float b = 2.0f * dot3(oc, d);
  const float c = dot3(oc, oc) - sp.w * sp.w;
  const float disc = b * b - 4.0f * a * c;
  if (disc < 0.0f) return false;
  t = div_ieee(-b - sqrt_ieee(disc), 2.0f * a);
  return t > 0.0f;
}

// "near root in (0, lim)" for shadow rays. Parity build: the literal root. Fast build: the
// same predicate without sqrt/div — t>0 <=> b<0 && c>0 (a>0); t<lim <=> sqrt(disc) > -b-2a*lim.
// The fast form takes two per-ray constants hoisted out of the candidate loop: m4a = -4a and
// m2al = -2a*lim (unused in the parity build).
__device__ __forceinline__ bool sphere_any(f3 oc, float cc, f3 d, float a, float lim, float m4a, float m2al) {
#if RT_PARITY
  const float b = 2.0f * dot3(oc, d);
  const float disc = b * b - 4.0f * a * cc;
  if (disc < 0.0f) return false;
  const float t = div_ieee(-b - sqrt_ieee(disc), 2.0f * a);
  return t > 0.0f && t < lim;
#else
  const float b = 2.0f * sdot3(oc, d);
  const float disc = fmaf(b, b, m4a * cc);
  const float q = m2al - b;
  return disc >= 0.0f && b < 0.0f && cc > 0.0f && (q < 0.0f || disc > q * q);
#endif
}

// intersectPlane (glsl:120-153) on a plane record; the frame is hoisted to staging time.
__device__ __forceinline__ bool plane_t(saddr rec, f3 o, f3 d, float& t) {
  const float4 pn = lds128(rec + 48), pp = lds128(rec + 96);
  const f3 n = xyz(pn);
  const float denom = dot3(n, d);
  if (!(fabsf(denom) > 1e-6f)) return false;
  t = div_ieee(dot3(xyz(pp) - o, n), denom);
  if (t < 0.0f) return false;
  const float4 pr = lds128(rec + 64), pf = lds128(rec + 80);
  const f3 local = (o + d * t) - xyz(pp);
  const float x = dot3(local, xyz(pr));
  const float z = dot3(local, xyz(pf));
  if (fabsf(x) > pn.w || fabsf(z) > pr.w) return false;
  return true;
}

struct Hit { int code; float t; };     // code: sphere slot k, or ns + plane slot; -1 = miss

__device__ __forceinline__ int hit_object_index(const Scene& s, int code) {
  return code < s.ns ? lds32i(s.sph_index + 4u * (unsigned)code) : lds32i(s.pln + (unsigned)(code - s.ns) * kPlaneBytes + 108u);
}

// Dev-only statistics build (-DRT_PK_STATS=1): per-thread atomics into kp.counters, read with
// rt_get_counters. 0 jobs, 1 jobs with no candidate, 2 overflow jobs, 3 candidates (list jobs),
// 4 prescan cone tests, 5 packets traced, 6 packet-loop object visits, 7 closest rays,
// 8 cluster slab tests, 9 member slab tests, 10 closest linear fallbacks; cooperative scan, per warp: 11 converged warps,
// 12 scans that produced a list, 13 clusters listed, 14 spheres listed, 15 not coherent, 16 / 17 cluster / sphere list overflow,
// 18 entered (some lane wanted a cluster scan), 19 closest-hit warps that shared their scan
#ifndef RT_PK_STATS
#define RT_PK_STATS 0
#endif
#if RT_PK_STATS
__device__ unsigned long long* g_stats;
#define PK_STAT(i, n) atomicAdd(g_stats + (i), (unsigned long long)(n))
#else
#define PK_STAT(i, n) ((void)0)
#endif

constexpr int kCandCap = 32;          // per-thread scratch list (u16 codes) in shared memory
#ifndef RT_PK_WCLUSTERS
#define RT_PK_WCLUSTERS 128
#endif
#ifndef RT_PK_WSPHERES
#define RT_PK_WSPHERES 128
#endif
// per-warp lists of the cooperative cluster scan (coop_cluster_scan): passing clusters, then passing sphere slots (u16 each)
constexpr int kWarpClusterCap = RT_PK_WCLUSTERS, kWarpSphereCap = RT_PK_WSPHERES;
constexpr int kWarpListBytes = (kWarpClusterCap + kWarpSphereCap) * 2;
// dynamic shared memory behind the scene blob: the per-thread candidate lists, then the per-warp lists
constexpr int kScratchBytes = kCandCap * kThreads * 2 + kWarpsPerCta * kWarpListBytes;
static_assert(kScratchBytes % 16 == 0, "the regrouping kernel's queues follow the scratch as float4");
#ifndef RT_PK_COOP
#define RT_PK_COOP 2                  // 0 off, 1 shadow jobs, 2 + closest-hit rays
#endif
__device__ __forceinline__ void cand_store(saddr cand_base, int i, int code) {
  asm volatile("st.shared.u16 [%0], %1;" :: "r"(cand_base + (unsigned)i * (kThreads * 2u)), "h"((unsigned short)code) : "memory");
}
__device__ __forceinline__ int cand_load(saddr cand_base, int i) {
  unsigned short v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(cand_base + (unsigned)i * (kThreads * 2u)) : "memory");
  return (int)v;
}

struct Cone { f3 L; float s, c2, far; f3 fn; float fe; bool fan; };   // axis, sin bound, cos^2 bound, far = lim*(1+e); fan plane normal (0 = none), PCF slack
struct Cands { saddr base; int n, first, nsph; bool list, prefilter; };   // !list: scan every object (prefilter: with the cone test); list entries [first, n), the first nsph of [0, n) are spheres

// FANSKIP: branch around the fan-plane test when the job has no probes. Worth about 1 % in the
// small-scene kernel's sphere scan (same-box A/B on C2); in the large-scene kernel's long scan loops
// the branch costs more than the 8 instructions it saves (C3 19.3 -> 20.3 ms): straight-line there.
template <bool FANSKIP = false>
__device__ __forceinline__ bool cone_misses_sphere(const Cone& c, f3 o, f3 center, float r) {
  const f3 v = center - o;
  const float along = fmaf(v.z, c.L.z, fmaf(v.y, c.L.y, v.x * c.L.x));
  const float d2 = fmaf(v.z, v.z, fmaf(v.y, v.y, v.x * v.x));
  const float rr = fabsf(r) * 1.001f;                       // a negative radius is the same sphere (glsl:108 squares it)
  const float q = fmaf(along, c.s, rr);                     // r' + along*sin
  const float lhs = fmaf(-along, along, d2) * c.c2;         // perp^2 * cos^2
  const float rhs = fmaf(d2, 4.0e-6f, q * q);
  if (q < 0.0f || lhs > rhs || along - rr > c.far) return true;      // any NaN: not culled
  // PCSS fan: the 16 probes all lie in the plane through o spanned by L and (1,1,1) (glsl:409 adds
  // the same scalar to x, y and z), the PCF samples within fe of L: a sphere farther than that from
  // the plane is out of reach. fn = 0 (no probes, or L || (1,1,1)) disables the test.
  if (FANSKIP && !c.fan) return false;                      // no probes in this job (or L || (1,1,1)): fn = 0, nothing to test
  const float w = fmaf(v.z, c.fn.z, fmaf(v.y, c.fn.y, v.x * c.fn.x));
  return fabsf(w) > fmaf(sqrtf(d2) * 1.001f + rr, c.fe, rr);
}

// ---- warp-cooperative cluster scan. The 32 rays (closest-hit) or shadow jobs a converged warp holds are usually neighbours
// (the pixels of one tile; consecutive hits of one pixel block under one light): instead of 32 lanes each testing every
// cluster and every member of the passing clusters on its own (C3: ~127 cone tests per wide-cone shadow job, 23 % of all
// executed instructions; C5: 155 slab tests per primary ray), the warp scans ONCE against a cone that contains all of them
// — one entry per lane and step — into a shared list, and every lane then runs its own exact tests on that short list.
// The union: apex = the mean origin om, axis = the normalised mean of the axes Lm, sin = max_i(sin angle(L_i, Lm) + s_i),
// far = max_i far_i + ro, and every radius grown by ro = max_i |o_i - om| (a ray point o_i + t d is within ro of om + t d,
// which lies in the union cone). Only the lanes with `want` contribute; the others just help scanning. A lane's final set is
// a subset of the shared one chosen by the SAME exact tests as before, so it is again a conservative superset of what its
// rays can touch: identical images (shadow packets decide with the exact per-ray tests, an any-hit does not depend on the
// order; the closest hit's tie rule is by original index, so its visiting order is free too).
// Returns the list length, or -1 — every lane alike — when the wanting lanes are not coherent enough (sin >= 0.7, a
// non-finite input, a list that overflows): the callers then scan per lane. All 32 lanes must call it together.
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ int lds_u16(saddr p) { unsigned short v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(p) : "memory"); return (int)v; }
__device__ __forceinline__ void sts_u16(saddr p, int v) { asm volatile("st.shared.u16 [%0], %1;" :: "r"(p), "h"((unsigned short)v) : "memory"); }

__device__ __noinline__ int coop_cluster_scan(saddr clu, saddr sph, int nc, int ns, saddr wl, bool want, f3 L, float cs, float cfar, f3 o) {
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  __syncwarp();                                                                      // every reader of the previous list is done
  const float probe = ((o.x + o.y) + o.z) + ((L.x + L.y) + L.z) + cs;                // NaN / inf anywhere -> not finite
  if (__any_sync(0xffffffffu, want && (!(fabsf(probe) < 3.0e38f) || cfar != cfar))) return -1;   // (a NaN far never culls: not unionable)
  if (!want) { o = mk3(0.0f, 0.0f, 0.0f); L = o; cs = 0.0f; cfar = 0.0f; }
  const float inv_n = 1.0f / (float)__popc(__ballot_sync(0xffffffffu, want));
  const f3 om = mk3(warp_sum(o.x) * inv_n, warp_sum(o.y) * inv_n, warp_sum(o.z) * inv_n);
  const f3 Ls = mk3(warp_sum(L.x), warp_sum(L.y), warp_sum(L.z));
  const float ll = fmaf(Ls.z, Ls.z, fmaf(Ls.y, Ls.y, Ls.x * Ls.x));
  if (!(ll > 0.25f)) { if (lane == 0) PK_STAT(15, 1); return -1; }                   // axes pointing apart: no common cone
  const f3 Lm = Ls * rsqrtf(ll);
  const f3 dv = want ? o - om : mk3(0.0f, 0.0f, 0.0f);
  const float ro = warp_max(sqrtf(fmaf(dv.z, dv.z, fmaf(dv.y, dv.y, dv.x * dv.x)))) * 1.001f + 1.0e-6f * (1.0f + fabsf(om.x) + fabsf(om.y) + fabsf(om.z));
  const f3 cx = cross3(L, Lm);                                                       // |L| in (0.995, 1.005) (callers), |Lm| = 1 up to rounding
  const float S = warp_max(want ? sqrtf(fmaf(cx.z, cx.z, fmaf(cx.y, cx.y, cx.x * cx.x))) * 1.01f + cs : 0.0f) * 1.01f + 1.0e-5f;
  const bool behind = want && !(fmaf(L.z, Lm.z, fmaf(L.y, Lm.y, L.x * Lm.x)) > 0.0f);   // sin alone cannot tell an axis from its opposite
  if (!(S < 0.7f) || __any_sync(0xffffffffu, behind)) { if (lane == 0) PK_STAT(15, 1); return -1; }
  Cone u;
  u.L = Lm; u.s = S; u.c2 = (1.0f - S * S) * 0.999f; u.far = warp_max(cfar) + ro * 1.001f;
  u.fn = mk3(0.0f, 0.0f, 0.0f); u.fe = 0.0f; u.fan = false;
  // level 1: cluster bounding spheres
  int n1 = 0;
#pragma unroll 1
  for (int c0 = 0; c0 < nc; c0 += 32) {
    const int c = c0 + lane;
    bool pass = false;
    if (c < nc) {
      const float4 q = lds128(clu + (unsigned)c * kSphereBytes + 48u);
      pass = !cone_misses_sphere(u, om, xyz(q), fabsf(q.w) + ro);
    }
    const unsigned b = __ballot_sync(0xffffffffu, pass);
    const int at = n1 + __popc(b & lt);
    if (pass && at < kWarpClusterCap) sts_u16(wl + 2u * (unsigned)at, c);
    n1 += __popc(b);
  }
  if (n1 > kWarpClusterCap) { if (lane == 0) PK_STAT(16, 1); return -1; }
  __syncwarp();
  // level 2: the members of the passing clusters
  int n2 = 0;
  const int tot = n1 * kPkClusterSize;
#pragma unroll 1
  for (int i0 = 0; i0 < tot; i0 += 32) {
    const int i = i0 + lane;
    bool pass = false;
    int k = 0;
    if (i < tot) {
      k = lds_u16(wl + 2u * (unsigned)(i / kPkClusterSize)) * kPkClusterSize + (i % kPkClusterSize);
      if (k < ns) {
        const float4 q = lds128(sph + (unsigned)k * kSphereBytes + 48u);
        pass = !cone_misses_sphere(u, om, xyz(q), fabsf(q.w) + ro);
      }
    }
    const unsigned b = __ballot_sync(0xffffffffu, pass);
    const int at = n2 + __popc(b & lt);
    if (pass && at < kWarpSphereCap) sts_u16(wl + 2u * (unsigned)(kWarpClusterCap + at), k);
    n2 += __popc(b);
  }
  if (n2 > kWarpSphereCap) { if (lane == 0) PK_STAT(17, 1); return -1; }
  __syncwarp();
  if (lane == 0) { PK_STAT(12, 1); PK_STAT(13, n1); PK_STAT(14, n2); }
  return n2;
}

// Closest hit of one ray over the whole scene, the shader's tie rule (glsl:155-196): among equal
// t the lower ORIGINAL index wins (what the in-order strict `<` of :178 does), so the visiting
// order is free. GENERIC: explicit min/max per axis (any direction, incl. NaN and 0), no clusters.
//
// Two-level boxes: with clusters, the ray first runs the SAME slab arithmetic against the union
// box of each 16 sphere slots. IEEE rounding is monotone, so (lo_u - o)*inv <= (lo_m - o)*inv for
// every member box m of the union u (and >= for hi): a ray that fails the union box fails every
// member box — the cluster test is exactly conservative, no tolerance involved. Passing clusters
// go to the lane's scratch list, then each lane walks the members of its own clusters.
template <bool GENERIC, bool BIG>
__device__ __forceinline__ Hit trace_one(const Scene& s, saddr cand_base, f3 o, f3 d, f3 inv, float a, bool coop = false) {
  const unsigned offx = (!GENERIC && d.x < 0.0f) ? 8u : 0u;      // the (near,far) pair this ray's sign selects
  const unsigned offy = (!GENERIC && d.y < 0.0f) ? 24u : 16u;
  const unsigned offz = (!GENERIC && d.z < 0.0f) ? 40u : 32u;
  const float nox = -o.x, noy = -o.y, noz = -o.z;
  Hit h; h.code = -1; h.t = s.maxd;
  auto slab = [&](saddr ax, saddr ay, saddr az) -> bool {
    const float2 X = mul2(add2(lds64(ax), dup2(nox)), dup2(inv.x));
    const float2 Y = mul2(add2(lds64(ay), dup2(noy)), dup2(inv.y));
    const float2 Z = mul2(add2(lds64(az), dup2(noz)), dup2(inv.z));
    float tmin, tmax;
    if (GENERIC) {
      tmin = fmax3(fminf(X.x, X.y), fminf(Y.x, Y.y), fminf(Z.x, Z.y));
      tmax = fmin3(fmaxf(X.x, X.y), fmaxf(Y.x, Y.y), fmaxf(Z.x, Z.y));
    } else {
      tmin = fmax3(X.x, Y.x, Z.x);
      tmax = fmin3(X.y, Y.y, Z.y);
    }
    return tmax >= tmin && tmin < s.maxd && tmax > 0.0f;      // glsl:102
  };
  auto sphere = [&](int k) {
    float t;
    if (!sphere_t(lds128(s.sph + (unsigned)k * kSphereBytes + 48u), o, d, a, t)) return;   // t > 0 checked in sphere_t
    if (t < h.t || (t == h.t && h.code >= 0 && hit_object_index(s, k) < hit_object_index(s, h.code))) {
      h.code = k; h.t = t;
    }
  };
  bool linear = true;
  PK_STAT(7, 1);
  // coop (warp-uniform; all 32 lanes are here together): one scan of the clusters for the whole warp against the cone that
  // holds its 32 rays, then this ray's own slab + sphere tests on the short shared list. A ray that touches a sphere lies in
  // the union cone and touches the bounding sphere the scan tests (with the margins of cone_misses_sphere), so the list holds
  // every sphere the ray can hit.
  int nshared = -1;
  if (BIG && !GENERIC && RT_PK_COOP && coop) {
    if ((threadIdx.x & 31) == 0) PK_STAT(19, 1);
    nshared = coop_cluster_scan(s.clu, s.sph, s.nc, s.ns, s.wlist, true, d * rsqrtf(a), 2.0e-5f, 3.0e38f, o);
  }
  if (BIG && !GENERIC && nshared >= 0) {
    linear = false;
#pragma unroll 2
    for (int i = 0; i < nshared; ++i) {
      const int k = lds_u16(s.wlist + 2u * (unsigned)(kWarpClusterCap + i));
      const saddr rec = s.sph + (unsigned)k * kSphereBytes;
      PK_STAT(9, 1);
      if (slab(rec + offx, rec + offy, rec + offz)) sphere(k);
    }
  } else if (BIG && !GENERIC && s.nc > 0) {
    int n = 0;
    PK_STAT(8, s.nc);
    saddr ax = s.clu + offx, ay = s.clu + offy, az = s.clu + offz;
#pragma unroll 2
    for (int c = 0; c < s.nc; ++c, ax += kSphereBytes, ay += kSphereBytes, az += kSphereBytes) {
      if (!slab(ax, ay, az)) continue;
      if (n < kCandCap) cand_store(cand_base, n, c);
      ++n;
    }
    if (n > kCandCap) PK_STAT(10, 1);
    if (n <= kCandCap) {
      linear = false;
#pragma unroll 1
      for (int i = 0; i < n; ++i) {
        const int first = cand_load(cand_base, i) * kPkClusterSize;
        const int last = min(first + kPkClusterSize, s.ns);
        PK_STAT(9, last - first);
        const saddr rec = s.sph + (unsigned)first * kSphereBytes;
        ax = rec + offx; ay = rec + offy; az = rec + offz;
#pragma unroll 2
        for (int k = first; k < last; ++k, ax += kSphereBytes, ay += kSphereBytes, az += kSphereBytes)
          if (slab(ax, ay, az)) sphere(k);
      }
    }
  }
  if (linear) {
    saddr ax = s.sph + offx, ay = s.sph + offy, az = s.sph + offz;
    int k = 0;
    for (; k + 2 <= s.ns; k += 2, ax += 2 * kSphereBytes, ay += 2 * kSphereBytes, az += 2 * kSphereBytes) {
      const bool p0 = slab(ax, ay, az), p1 = slab(ax + kSphereBytes, ay + kSphereBytes, az + kSphereBytes);
      if (p0 | p1) {
        if (p0) sphere(k);
        if (p1) sphere(k + 1);
      }
    }
    if (k < s.ns && slab(ax, ay, az)) sphere(k);
  }
  saddr rec = s.pln;
  for (int j = 0; j < s.np; ++j, rec += kPlaneBytes) {
    if (!slab(rec + offx, rec + offy, rec + offz)) continue;
    float t;
    if (!plane_t(rec, o, d, t) || !(t > 0.0f)) continue;
    if (t < h.t || (t == h.t && h.code >= 0 && lds32i(rec + 108u) < hit_object_index(s, h.code))) {
      h.t = t; h.code = s.ns + j;
    }
  }
  return h;
}

// The cold variant lives behind a call so the hot loop stays small in the instruction cache.
__device__ __noinline__ Hit trace_closest_generic(const Scene& s, f3 o, f3 d, f3 inv, float a) {
  return trace_one<true, false>(s, 0u, o, d, inv, a);
}

// Closest hit for primary / bounce / SSS rays. IEEE 1/dir in both builds: the hit set of these
// rays feeds P and N.
template <bool BIG>
__device__ __forceinline__ Hit trace_closest(const Scene& s, saddr cand_base, f3 o, f3 d) {
  const f3 inv = mk3(rcp_ieee(d.x), rcp_ieee(d.y), rcp_ieee(d.z));
  const float a = dot3(d, d);
  // the whole warp is here, converged, every ray regular: the cluster scan is shared (see trace_one)
  bool coop = false;
  if (BIG && RT_PK_COOP >= 2 && s.nc > 0 && s.wlist != 0u && __activemask() == 0xffffffffu)
    coop = __all_sync(0xffffffffu, dir_regular(d, a));
  if (!dir_regular(d, a)) {
    // rare path: the out-of-line routine takes COPIES, so that the structs of the common path are
    // never address-taken and stay in registers (a by-reference argument of a noinline function
    // pins its object in local memory: measured 6 % of the kernel's instructions as STL/LDL)
    const Scene s2 = s;
    return trace_closest_generic(s2, o, d, inv, a);
  }
  return trace_one<false, BIG>(s, cand_base, o, d, inv, a, coop);
}

// ---------------------------------------------------------------- shadow jobs: cone culling
// All shadow rays of one (shading point, light) — 16 PCSS probes and/or the PCF samples — start
// at the same origin and lie in a cone around lightDir: a probe is normalize(L + 2*rv*ss*(1,1,1))
// (glsl:408-409, |rv| <= 1), a PCF sample is L + fs*(rx*T + ry*B) with rx,ry in [0,1), |T| = 1,
// |B| <= 1 (glsl:366-378); for |L| = 1 and |delta| <= e < 1 the angle to L has sin <= e.
// An object no ray of that cone can touch cannot occlude, whatever its AABB says (a hit needs the
// AABB test AND the primitive test), so the job first scans the scene once with a CONSERVATIVE
// cone-vs-bounding-sphere test and the ray packets then visit only the survivors, with the exact
// per-ray tests. No survivor: every probe / sample is unoccluded, the factor is exactly 1.
// Conservative: sin is over-, cos under-estimated, radii are inflated by 1e-3, and the squared
// comparison carries a 4e-6*|v|^2 slack (40x the rounding of d2 - along^2, and above the
// 1e-7*|v|^2 by which the shader's own discriminant can report a grazing hit).
__device__ __forceinline__ bool cone_misses_plane(const Cone& c, f3 o, saddr rec) {
  const float4 pn = lds128(rec + 48u), pp = lds128(rec + 96u), pb = lds128(rec + 112u);
  const f3 n = xyz(pn);
  const float s0 = dot3(xyz(pp) - o, n);                    // the numerator of glsl:127, same operations
  const float nl = fmaf(n.z, c.L.z, fmaf(n.y, c.L.y, n.x * c.L.x));
  const float slack = pb.y * (c.s + 1.0e-5f);               // |n.delta| <= |n| e (+ rounding)
  // every denominator n.d has one strict sign and the numerator the opposite (or zero): t <= 0
  if ((s0 <= 0.0f && nl > slack) || (s0 >= 0.0f && nl < -slack)) return true;
  // t = s0 / (n.d) with |n.d| <= |n| |d| and |d| <= 1 + e: a plane farther than that from the origin
  // is beyond every ray's limit (far already carries the (1 + e) factor; NaN: not culled)
#if !RT_PK_NO_PLANE_FAR
  if (fabsf(s0) > c.far * pb.y * (1.0f + c.s)) return true;
#endif
  return cone_misses_sphere(c, o, xyz(pp), pb.x);
}

__device__ __forceinline__ saddr record_of(const Scene& s, int code) {
  return code < s.ns ? s.sph + (unsigned)code * kSphereBytes : s.pln + (unsigned)(code - s.ns) * kPlaneBytes;
}
__device__ __forceinline__ int cand_code(const Cands& cd, int i) { return cd.list ? cand_load(cd.base, i) : i; }

// Cell of the light grid a job reads (rt_abi.cu build_light_grids holds the derivation and the layout),
// or -1 when the job is not covered: no grid for this light, a wider cone than the grid was inflated for,
// an origin farther from the light / outside the scene ball than the bound the grid assumes (NaN: -1).
__device__ __forceinline__ int grid_cell(const Scene& s, int light_index, const Cone& cone, f3 o, float lightDistance, int& first_cell, int& shared_cell) {
  const float4 h0 = __ldg(s.grid + 4 * light_index);
  const int mode = __float_as_int(h0.x);
  if (mode == 0 || !(cone.s <= h0.y)) return -1;
  first_cell = __float_as_int(h0.w);
  const float4 h1 = __ldg(s.grid + 4 * light_index + 1);
  if (mode == 1) {                                           // cube map of directions around the light: look up -L
    if (!(lightDistance <= h0.z)) return -1;
    const int G = (int)h1.w;
    shared_cell = 6 * G * G;
    const float x = -cone.L.x, y = -cone.L.y, z = -cone.L.z;
    const float ax = fabsf(x), ay = fabsf(y), az = fabsf(z);
    int face; float a, b, m;
    if (ax >= ay && ax >= az) { face = x >= 0.0f ? 0 : 1; a = y; b = z; m = ax; }
    else if (ay >= az)        { face = y >= 0.0f ? 2 : 3; a = x; b = z; m = ay; }
    else                      { face = z >= 0.0f ? 4 : 5; a = x; b = y; m = az; }
    if (!(m > 0.0f)) return -1;
    const float inv = __fdividef(0.5f * (float)G, m), hg = 0.5f * (float)G;
    int ix = (int)floorf(fmaf(a, inv, hg)), iy = (int)floorf(fmaf(b, inv, hg));
    ix = min(max(ix, 0), G - 1); iy = min(max(iy, 0), G - 1);
    return (face * G + iy) * G + ix;
  }
  const float4 h2 = __ldg(s.grid + 4 * light_index + 2), h3 = __ldg(s.grid + 4 * light_index + 3);
  const f3 v = o - xyz(h3);                                  // orthographic grid: the origin's projection
  if (!(fmaf(v.z, v.z, fmaf(v.y, v.y, v.x * v.x)) <= h0.z)) return -1;
  constexpr int G = 32;
  shared_cell = G * G;
  const float ou = fmaf(o.z, h1.z, fmaf(o.y, h1.y, o.x * h1.x)), ov = fmaf(o.z, h2.z, fmaf(o.y, h2.y, o.x * h2.x));
  int ix = (int)floorf((ou - h1.w) * h3.w), iy = (int)floorf((ov - h2.w) * h3.w);
  ix = min(max(ix, 0), G - 1); iy = min(max(iy, 0), G - 1);
  return iy * G + ix;
}

// One scan of the scene per shadow job. Returns the candidate set for the job's packets.
template <bool BIG>
__device__ __forceinline__ Cands scan_candidates(const Scene& s, saddr cand_base, bool cone_ok, const Cone& cone, f3 o,
                                                 int light_index, float lightDistance, bool warp_converged) {
  Cands cd; cd.base = cand_base; cd.n = s.ns + s.np; cd.first = 0; cd.nsph = s.ns; cd.list = false; cd.prefilter = false;
  int first_cell = 0, shared_cell = 0;
  const int cell = (BIG && s.grid && cone_ok) ? grid_cell(s, light_index, cone, o, lightDistance, first_cell, shared_cell) : -1;
  // warp_converged: the caller vouches that all 32 lanes of the warp get here, each with a job (of any light). When any
  // of them is about to scan the clusters, the warp scans once, together (coop_cluster_scan); nshared < 0: each on its own.
  int nshared = -1;
  if (BIG && RT_PK_COOP && warp_converged) {
    const bool want = cone_ok && s.nc > 0 && s.wlist != 0u && cell < 0;
    if ((threadIdx.x & 31) == 0) PK_STAT(11, 1);
    if (__any_sync(0xffffffffu, want)) {
      if ((threadIdx.x & 31) == 0) PK_STAT(18, 1);
      nshared = coop_cluster_scan(s.clu, s.sph, s.nc, s.ns, s.wlist, want, cone.L, cone.s, cone.far, o);
      if (!want) nshared = -1;
    }
  }
  if (!cone_ok) return cd;
  int n = 0;
  if (BIG && cell >= 0) {                                    // light grid: the cell's spheres + the shared list, exact cone test on each
    const unsigned* offs = reinterpret_cast<const unsigned*>(s.grid + s.grid_off_f4) + first_cell;
    const unsigned short* ents = reinterpret_cast<const unsigned short*>(s.grid + s.grid_ent_f4);
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
      const int cidx = pass ? shared_cell : cell;
      const unsigned e0 = __ldg(offs + cidx), e1 = __ldg(offs + cidx + 1);
#pragma unroll 1
      for (unsigned e = e0; e < e1; ++e) {
        const int k = (int)__ldg(ents + e);
        const float4 sp = lds128(s.sph + (unsigned)k * kSphereBytes + 48u);
        PK_STAT(4, 1);
        if (cone_misses_sphere(cone, o, xyz(sp), sp.w)) continue;
        if (n < kCandCap) cand_store(cand_base, n, k);
        ++n;
      }
    }
  } else if (BIG && nshared >= 0) {                          // the warp's shared list, filtered by this job's own exact cone test
#pragma unroll 2
    for (int i = 0; i < nshared; ++i) {
      const int k = lds_u16(s.wlist + 2u * (unsigned)(kWarpClusterCap + i));
      const float4 sp = lds128(s.sph + (unsigned)k * kSphereBytes + 48u);
      PK_STAT(4, 1);
      if (cone_misses_sphere(cone, o, xyz(sp), sp.w)) continue;
      if (n < kCandCap) cand_store(cand_base, n, k);
      ++n;
    }
  } else if (BIG && s.nc > 0) {                              // cluster bounding spheres first, then members
    saddr crec = s.clu;
#pragma unroll 1
    for (int c = 0; c < s.nc; ++c, crec += kSphereBytes) {
      const float4 cs = lds128(crec + 48u);
      PK_STAT(4, 1);
      if (cone_misses_sphere(cone, o, xyz(cs), cs.w)) continue;
      PK_STAT(4, min(kPkClusterSize, s.ns - c * kPkClusterSize));
      const int first = c * kPkClusterSize, last = min(first + kPkClusterSize, s.ns);
      saddr rec = s.sph + (unsigned)first * kSphereBytes;
#pragma unroll 2
      for (int k = first; k < last; ++k, rec += kSphereBytes) {
        const float4 sp = lds128(rec + 48u);
        if (cone_misses_sphere(cone, o, xyz(sp), sp.w)) continue;
        if (n < kCandCap) cand_store(cand_base, n, k);
        ++n;
      }
    }
  } else {
    saddr rec = s.sph;
#pragma unroll 2
    for (int k = 0; k < s.ns; ++k, rec += kSphereBytes) {
      const float4 sp = lds128(rec + 48u);
      if (cone_misses_sphere<!BIG>(cone, o, xyz(sp), sp.w)) continue;
      if (n < kCandCap) cand_store(cand_base, n, k);
      ++n;
    }
  }
  // (the shared list is rewritten only behind the __syncwarp at the top of the warp's next coop_cluster_scan)
  cd.nsph = n;
  saddr rec = s.pln;
  for (int j = 0; j < s.np; ++j, rec += kPlaneBytes) {
    if (cone_misses_plane(cone, o, rec)) continue;
    if (n < kCandCap) cand_store(cand_base, n, s.ns + j);
    ++n;
  }
  PK_STAT(0, 1);
  if (n == 0) PK_STAT(1, 1);
  if (n <= kCandCap) { cd.list = true; cd.n = n; PK_STAT(3, n); }
  else { cd.prefilter = true; PK_STAT(2, 1); }                                  // too many survivors: full scan, cone test inline
  return cd;
}

// One shadow ray against the job's candidates (packets whose rays disagree in sign, or carry an
// irregular direction, fall back to this ray by ray).
__device__ __noinline__ bool trace_any_one(const Scene& s, const Cands& cd, const Cone& cone, f3 o, f3 d, float lim) {
  const float a = dot3(d, d);
  const bool regular = dir_regular(d, a);
  const f3 inv = regular ? mk3(rcp_shadow(d.x), rcp_shadow(d.y), rcp_shadow(d.z)) : mk3(rcp_ieee(d.x), rcp_ieee(d.y), rcp_ieee(d.z));
  const unsigned ox = (regular && d.x < 0.0f) ? 8u : 0u, oy = (regular && d.y < 0.0f) ? 24u : 16u, oz = (regular && d.z < 0.0f) ? 40u : 32u;
#pragma unroll 1
  for (int i = cd.first; i < cd.n; ++i) {
    const int code = cand_code(cd, i);
    const saddr rec = record_of(s, code);
    if (cd.prefilter && (code < s.ns ? cone_misses_sphere(cone, o, xyz(lds128(rec + 48u)), lds32(rec + 60u)) : cone_misses_plane(cone, o, rec))) continue;
    const float2 X = mul2(add2(lds64(rec + ox), dup2(-o.x)), dup2(inv.x));
    const float2 Y = mul2(add2(lds64(rec + oy), dup2(-o.y)), dup2(inv.y));
    const float2 Z = mul2(add2(lds64(rec + oz), dup2(-o.z)), dup2(inv.z));
    float tmin, tmax;
    if (regular) { tmin = fmax3(X.x, Y.x, Z.x); tmax = fmin3(X.y, Y.y, Z.y); }
    else {
      tmin = fmax3(fminf(X.x, X.y), fminf(Y.x, Y.y), fminf(Z.x, Z.y));
      tmax = fmin3(fmaxf(X.x, X.y), fmaxf(Y.x, Y.y), fmaxf(Z.x, Z.y));
    }
    if (!(tmax >= tmin && tmin < s.maxd && tmax > 0.0f)) continue;
    float t;
    const bool hit = code < s.ns ? sphere_t(lds128(rec + 48u), o, d, a, t) : plane_t(rec, o, d, t);
    if (hit && t > 0.0f && t < lim) return true;
  }
  return false;
}

// 4 shadow rays with a common origin against the job's candidates. Returns the mask of rays that
// hit something with 0 < t < lim. Rays of the packet that agree in all three direction signs are
// traced together (one traversal per sign group: normally one, two when a PCSS packet straddles an
// axis); a ray with a NaN direction cannot hit anything (every comparison is false); a ray with a
// zero / denormal / infinite component goes through the generic single-ray routine.
__device__ __forceinline__ unsigned trace_any_packet(const Scene& s, const Cands& cd, const Cone& cone, f3 o,
                                                     const f3 (&d)[4], unsigned active, float lim, unsigned usign) {
  float a[4], m4a[4], m2al[4];
  float2 ivx[4], ivy[4], ivz[4];
  unsigned signs = 0u, remaining = active, occ = 0u;       // signs: 3 bits per ray
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    a[r] = sdot3(d[r], d[r]);
    m4a[r] = -4.0f * a[r]; m2al[r] = -2.0f * a[r] * lim;
    ivx[r] = dup2(rcp_shadow(d[r].x)); ivy[r] = dup2(rcp_shadow(d[r].y)); ivz[r] = dup2(rcp_shadow(d[r].z));   // FMUL2 takes the scalar broadcast form
  }
  if (usign < 8u) signs = usign * 0x249u;                  // the job proved every ray regular, with these signs (shadow_packets)
  else {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      signs |= ((__float_as_uint(d[r].x) >> 31) | ((__float_as_uint(d[r].y) >> 31) << 1) | ((__float_as_uint(d[r].z) >> 31) << 2)) << (3 * r);
      if (((active >> r) & 1u) && !dir_regular(d[r], a[r])) {
        remaining &= ~(1u << r);
        if (a[r] == a[r]) {                                                             // a is NaN iff a component is
          const Scene s2 = s; const Cands cd2 = cd; const Cone cone2 = cone;            // copies: see trace_closest
          if (trace_any_one(s2, cd2, cone2, o, d[r], lim)) occ |= 1u << r;
        }
      }
    }
  }
  const float nox = -o.x, noy = -o.y, noz = -o.z;
  PK_STAT(5, 1);
#pragma unroll 1
  while (remaining) {
    const unsigned lead = (signs >> (3 * (__ffs(remaining) - 1))) & 7u;
    unsigned group = 0u;
#pragma unroll
    for (int r = 0; r < 4; ++r) if (((signs >> (3 * r)) & 7u) == lead) group |= 1u << r;
    group &= remaining;
    remaining &= ~group;
    const unsigned ox = (lead & 1u) ? 8u : 0u, oy = (lead & 2u) ? 24u : 16u, oz = (lead & 4u) ? 40u : 32u;
#pragma unroll 1
    for (int i = cd.first; i < cd.n; ++i) {
      const int code = cand_code(cd, i);
      PK_STAT(6, 1);
      const saddr rec = record_of(s, code);
      const bool is_sphere = code < s.ns;
      if (cd.prefilter && (is_sphere ? cone_misses_sphere(cone, o, xyz(lds128(rec + 48u)), lds32(rec + 60u)) : cone_misses_plane(cone, o, rec))) continue;
      const float2 X = add2(lds64(rec + ox), dup2(nox)), Y = add2(lds64(rec + oy), dup2(noy)), Z = add2(lds64(rec + oz), dup2(noz));
      bool pass[4], any = false;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float2 tx = mul2(X, ivx[r]), ty = mul2(Y, ivy[r]), tz = mul2(Z, ivz[r]);
        const float tmin = fmax3(tx.x, ty.x, tz.x), tmax = fmin3(tx.y, ty.y, tz.y);
        pass[r] = tmax >= tmin && tmin < s.maxd && tmax > 0.0f && ((group >> r) & 1u);
        any = any || pass[r];
      }
      if (!any) continue;
      if (is_sphere) {
        const float4 sp = lds128(rec + 48u);
        const f3 oc = o - xyz(sp);
        const float cc = smad(-sp.w, sp.w, sdot3(oc, oc));                 // dot(oc,oc) - r*r
#pragma unroll
        for (int r = 0; r < 4; ++r)
          if (pass[r] && !((occ >> r) & 1u) && sphere_any(oc, cc, d[r], a[r], lim, m4a[r], m2al[r])) occ |= 1u << r;
      } else {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          float t;
          if (pass[r] && !((occ >> r) & 1u) && plane_t(rec, o, d[r], t) && t > 0.0f && t < lim) occ |= 1u << r;
        }
      }
      if ((occ & group) == group) break;                    // every ray of the group is occluded
    }
  }
  return occ;
}

// ---------------------------------------------------------------- shading
struct LightRec {
  int type; f3 position, direction, color; float intensity, shadowSoftness; int shadowType, pcfSamples; float lightSize;
};
__device__ __forceinline__ LightRec load_light(const Scene& s, int i) {
  const saddr l = s.lights + (unsigned)i * 96u;          // rt_light offsets (include/rt_abi.h)
  const float4 q1 = lds128(l + 16u), q2 = lds128(l + 32u), q3 = lds128(l + 48u);
  LightRec r;
  r.type = lds32i(l);
  r.position = xyz(q1); r.direction = xyz(q2); r.color = xyz(q3); r.intensity = q3.w;
  r.shadowSoftness = lds32(l + 72u); r.shadowType = lds32i(l + 76u);
  r.pcfSamples = lds32i(l + 80u); r.lightSize = lds32(l + 84u);
  return r;
}
__device__ __forceinline__ Mat load_material(const Scene& s, int index) {
  float4 m0, m1, m2, m3;
  if (s.material_smem) {
    const saddr a = s.material_smem + (unsigned)index * 64u;
    m0 = lds128(a); m1 = lds128(a + 16u); m2 = lds128(a + 32u); m3 = lds128(a + 48u);
  } else {
    m0 = __ldg(s.material + 4 * index + 0); m1 = __ldg(s.material + 4 * index + 1);
    m2 = __ldg(s.material + 4 * index + 2); m3 = __ldg(s.material + 4 * index + 3);
  }
  Mat m;
  m.albedo = xyz(m0); m.metallic = m0.w;
  m.roughness = m1.x; m.diffuseStrength = m1.y; m.ior = m1.z; m.transparency = m1.w;
  m.subsurfaceColor = xyz(m2); m.subsurfaceScatter = m2.w; m.scatterDistance = m3.x;
  return m;
}

#if !RT_PARITY
// Fast build only: the 16 PCSS probes of one job decided analytically. Probe k leaves the job's
// origin along u(x) = L + x*(1,1,1), x = 2*rv_k*searchSize (glsl:408-409; the normalisation does
// not change what it hits). For a sphere (centre c, radius r), with oc = o - c and cc = oc.oc - r^2:
//     B(x) = oc.u = A0 + A1 x,   U(x) = u.u = LL + 2 Lw x + 3 x^2,
//     the ray hits the sphere (near root > 0)  <=>  cc > 0,  B(x) < 0,  p(x) = B^2 - U cc >= 0,
// and p is a quadratic in x whose three coefficients are per (job, sphere) constants. Every probe
// is classified as a sure hit (p > tol and B < -tol), a sure miss (p < -tol or B > tol) or
// undecided, with tol = 1e-4 x the magnitude of the terms that cancel (100 x their rounding). A
// sure hit on any candidate means "a blocker exists" (glsl:425); sure misses on all of them mean
// "fully lit" (glsl:428-430); one undecided probe, a sphere that reaches beyond the light
// (t < lightDistance would matter) or a sphere too small for its distance sends the job down the
// exact per-ray packets; plane candidates always take the exact packets (over the planes alone
// when every sphere is a sure miss). Out of line, scalar arguments: small scenes without a PCSS
// light keep their instruction footprint. The host guarantees box >= sphere (fan_ok), so a
// sure hit also passes the slab test.  Returns 0 none blocked, 1 blocked, 2 undecided.
__device__ __noinline__ int probe_fan(saddr sph, saddr cand_base, int nsph, f3 o, f3 L, float searchSize, float lim) {
  const float LL = sdot3(L, L), Lw = (L.x + L.y) + L.z, xs = 2.0f * searchSize;
  bool undecided = false;
#pragma unroll 1
  for (int i = 0; i < nsph; ++i) {
    const float4 sp = lds128(sph + (unsigned)cand_load(cand_base, i) * kSphereBytes + 48u);
    const f3 oc = o - xyz(sp);
    const float d2 = sdot3(oc, oc);
    const float cc = smad(-sp.w, sp.w, d2);                             // the packets' own expression: same bits
    if (!(cc > 0.0f)) continue;                                         // origin inside / on the sphere: the near root is <= 0
    const float r = fabsf(sp.w), dist = sqrt_shade(d2);
    if (!(dist + r < lim * 0.9999f) || !(r * 1000.0f > dist)) { undecided = true; continue; }
    const float A0 = sdot3(oc, L), A1 = (oc.x + oc.y) + oc.z;
    const float q2 = smad(A1, A1, -3.0f * cc), q1 = smad(A0, A1, -Lw * cc), q0 = smad(A0, A0, -LL * cc);
    const float acc = fabsf(cc);
    const float m2 = smad(A1, A1, 3.0f * acc) * 1.0e-4f, m1 = (fabsf(A0 * A1) + fabsf(Lw) * acc) * 2.0e-4f, m0 = smad(A0, A0, LL * acc) * 1.0e-4f;
    const float aA0 = fabsf(A0) * 1.0e-4f, aA1 = fabsf(A1) * 1.0e-4f;
#pragma unroll 4
    for (int k = 0; k < 16; ++k) {
      const float x = c_probe_rv[k] * xs, ax = fabsf(x);
      const float p = smad(smad(q2, x, 2.0f * q1), x, q0);
      const float tol = smad(smad(m2, ax, m1), ax, m0);
      const float B = smad(A1, x, A0), tolB = smad(aA1, ax, aA0);
      if (p > tol && B < -tolB) return 1;
      if (!(p < -tol || B > tolB)) undecided = true;
    }
  }
  return undecided ? 2 : 0;
}
#endif

// calculateShadow (glsl:442-455) for shadowType 1 (pcfShadow, glsl:342-397) and 2 (pcssShadow,
// glsl:400-440 = "fully lit unless one of 16 fixed probes is blocked, else PCF"). Probes and PCF
// samples run through ONE packet site: jobs [0,nprobe) are probes, [nprobe,nprobe+pcf) samples.
// MODE 0: the whole job. MODE 1: the prescan only — returns with done = true when the factor is already known (nothing in
// the cone: 1.0), else done = false and the candidate set in cdio. MODE 2: the packets only, over the candidate set cdio
// of an earlier MODE 1 call with the SAME arguments (cdio.base may be another thread's scratch list). Everything between
// — the cone, the limits — is a pure function of the arguments, so 1 + 2 computes exactly what 0 does: the regrouping
// megakernel (rt_regroup.cuh) uses the split to run the packets on survivors only.
template <bool BIG, bool FAN, int MODE>
__device__ __forceinline__ float shadow_packets_m(const Scene& s, saddr cand_base, float noise_r, f3 origin, const LightRec& light,
                                                  int light_index, f3 lightDir, float lightDistance, Cands& cdio, bool& done,
                                                  bool warp_converged = false) {
  const int nprobe = light.shadowType == 2 ? 16 : 0;
  const int total = nprobe + (light.pcfSamples > 0 ? light.pcfSamples : 0);
  const float filterSize = light.shadowSoftness * 0.005f;                        // glsl:364 (overwrites :352)
  const float searchSize = light.lightSize * 0.1f;                               // glsl:404
  // t < lightDistance && t < maxRayDistance as one bound; a NaN distance stays NaN (never occluded)
  const float lim_light = lightDistance > s.maxd ? s.maxd : lightDistance;
  const float lim_pcf = (light.type == 0 || light.type == 2) ? lim_light : s.maxd;   // glsl:388-392
  const float lim_probe = light.type != 1 ? lim_light : s.maxd;                      // glsl:419-423
  // the job's cone (see "shadow jobs: cone culling"): e bounds |direction - lightDir|
  const float e_pcf = 2.0f * fabsf(filterSize), e_probe = nprobe ? 3.4642f * fabsf(searchSize) : 0.0f;
  Cone cone;
  cone.L = lightDir;
  cone.s = fmaxf(e_pcf, e_probe) * 1.01f + 1.0e-5f;
  cone.c2 = (1.0f - cone.s * cone.s) * 0.999f;
  cone.far = fmaxf(lim_pcf, lim_probe) * (1.0f + cone.s) * 1.0001f;              // NaN limits: fmaxf keeps the other
  cone.fn = mk3(0.0f, 0.0f, 0.0f);
  cone.fan = false;
  cone.fe = e_pcf * 1.01f + 2.0e-5f;              // + the rounding of a probe direction off its plane
  if (nprobe) {
    const f3 fx = cross3(lightDir, mk3(0.57735026f, 0.57735026f, 0.57735026f));
    const float f2 = dot3(fx, fx);
    if (f2 > 1.0e-4f) { cone.fn = fx * rsqrtf(f2); cone.fan = true; }
  }
  const float ll = dot3(lightDir, lightDir);
  const bool cone_ok = (unsigned)light.type <= 2u && cone.s < 0.7f && ll > 0.99f && ll < 1.01f;
  const Cands cd = MODE == 2 ? cdio : scan_candidates<BIG>(s, cand_base, cone_ok, cone, origin, light_index, lightDistance, warp_converged);
  if (MODE == 1) {
    cdio = cd;
    done = cd.list && cd.n == 0 && (nprobe || light.pcfSamples > 0);
    return 1.0f;
  }
  if (cd.list && cd.n == 0 && (nprobe || light.pcfSamples > 0)) return 1.0f;    // nothing can occlude: n/n, or no blocker
  int unoccluded = 0;
  int j = 0;
  int probe_first = 0;                                   // exact probe packets visit candidates [probe_first, n)
#if !RT_PARITY
  if (FAN && nprobe && cd.list && cd.nsph > 0) {
    const int verdict = probe_fan(s.sph, cd.base, cd.nsph, origin, lightDir, searchSize, lim_probe);
    if (verdict == 1) j = nprobe;                        // a blocker exists: straight to the PCF phase
    else if (verdict == 0) {
      if (cd.nsph == cd.n) return 1.0f;                  // glsl:428-430: no blocker -> fully lit
      probe_first = cd.nsph;                             // no sphere blocks: the planes decide, exactly
    }
  }
#endif
  // (the frame only steers shadow rays: single-MUFU normalisation in the fast build; 0 * inf is NaN in both)
  const f3 tangent = normalize_shadow(cross3(lightDir, mk3(0.0f, 1.0f, 0.0f)));   // NaN when lightDir || Y (glsl:349)
  const f3 bitangent = cross3(lightDir, tangent);
#if !RT_PARITY
  const f3 tangent_fs = tangent * filterSize, bitangent_fs = bitangent * filterSize;
#endif
  // Every ray of the job is lightDir + delta with |delta_i| <= e <= cone.s per component (glsl:371-378, :409;
  // normalising scales by a positive number): when every |lightDir_i| clears that bound, all rays are
  // regular and share lightDir's signs, and the packets skip their per-ray sign / regularity bookkeeping.
  unsigned usign = 8u;                                   // 8: not proven uniform
  if (cone_ok && fmin3(fabsf(lightDir.x), fabsf(lightDir.y), fabsf(lightDir.z)) > cone.s * 1.05f + 1.0e-6f)
    usign = (__float_as_uint(lightDir.x) >> 31) | ((__float_as_uint(lightDir.y) >> 31) << 1) | ((__float_as_uint(lightDir.z) >> 31) << 2);
  while (j < total) {
    const bool probing = j < nprobe;
    const int end = probing ? nprobe : total;
    f3 d[4];
    unsigned active = 0u;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int idx = j + r;
      if (idx < end) {
        active |= 1u << r;
        if (probing) {
          const float rv = c_probe_rv[idx];                                                  // glsl:408
          d[r] = normalize_shadow((lightDir + rv * searchSize) + rv * searchSize);           // glsl:409
        } else {
          const int i = idx - nprobe;
          const float rx = fract_(halton2(i) + noise_r);                                     // glsl:366-367
          const float ry = fract_(halton3(i) + 0.0f);
#if RT_PARITY
          f3 v = (lightDir + (rx * tangent) * filterSize) + (ry * bitangent) * filterSize;   // glsl:371-378
#else
          f3 v = smad3(bitangent_fs, ry, smad3(tangent_fs, rx, lightDir));                   // same point, fused
#endif
          if (light.type != 1) v = normalize_shadow(v);
          d[r] = v;
        }
      } else {
        d[r] = d[0];
      }
    }
    Cands cdj = cd;
    if (FAN) cdj.first = probing ? probe_first : 0;
    const unsigned occ = trace_any_packet(s, cdj, cone, origin, d, active, probing ? lim_probe : lim_pcf, usign) & active;
    if (probing) {
      if (occ) { j = nprobe; continue; }                 // a blocker exists: go to the PCF phase
      j += 4;
      if (j >= nprobe) return 1.0f;                      // glsl:428-430: no blocker -> fully lit
    } else {
      unoccluded += __popc(active & ~occ);
      j += 4;
    }
  }
  return div_ieee((float)unoccluded, (float)light.pcfSamples);     // 0/0 = NaN when pcfSamples == 0
}
template <bool BIG, bool FAN>
__device__ __forceinline__ float shadow_packets(const Scene& s, saddr cand_base, float noise_r, f3 origin, const LightRec& light,
                                                int light_index, f3 lightDir, float lightDistance, bool warp_converged = false) {
  Cands cd; bool done;
  return shadow_packets_m<BIG, FAN, 0>(s, cand_base, noise_r, origin, light, light_index, lightDir, lightDistance, cd, done, warp_converged);
}

// computeSubsurfaceScattering (glsl:316-339); unreachable from the reference's UI and scene files.
__device__ __noinline__ f3 subsurface(const Scene& s, f3 P, f3 N, const Mat& mat) {
  f3 sss = mk3(0.0f, 0.0f, 0.0f);
  for (int i = 0; i < 4; ++i) {
    const f3 d = cosine_hemisphere((float)i / 4.0f, halton(i, 2), N);
    const f3 o = P + N * 0.001f;
    const f3 inv = mk3(rcp_ieee(d.x), rcp_ieee(d.y), rcp_ieee(d.z));
    const Hit h = trace_closest_generic(s, o, d, inv, dot3(d, d));
    if (h.code >= 0) {
      const float attenuation = expf(div_shade(-h.t, mat.scatterDistance));
      sss = sss + xyz(__ldg(s.material + 4 * hit_object_index(s, h.code))) * attenuation;
    }
  }
  const f3 v = (sss * mat.subsurfaceColor) * mat.subsurfaceScatter;
  return mk3(v.x / 4.0f, v.y / 4.0f, v.z / 4.0f);
}

// Phase barrier of the CTA-coherent mode (see the kernel): keeps the warps that share an SM
// sub-partition inside the same stretch of code. `coh` is kernel-uniform.
// RT_PK_SYNC_WARPS (dev A/B): phase barriers over groups of that many warps (named barriers
// 1 + warp / group) instead of the whole CTA.
#ifndef RT_PK_SYNC_WARPS
#define RT_PK_SYNC_WARPS 0
#endif
__device__ __forceinline__ void phase_sync(bool on) {
  if (!on) return;
#if RT_PK_SYNC_WARPS > 0
  asm volatile("bar.sync %0, %1;" ::"r"(1u + (threadIdx.x >> 5) / RT_PK_SYNC_WARPS), "r"(32u * RT_PK_SYNC_WARPS) : "memory");
#else
  __syncthreads();
#endif
}
__device__ __forceinline__ bool phase_any(bool pred) {
#if RT_PK_SYNC_WARPS > 0
  unsigned r;
  asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.u32 q, %3, 0;\n\tbar.red.or.pred p, %1, %2, q;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(r) : "r"(1u + (threadIdx.x >> 5) / RT_PK_SYNC_WARPS), "r"(32u * RT_PK_SYNC_WARPS), "r"((unsigned)pred) : "memory");
  return r != 0u;
#else
  return __syncthreads_or(pred) != 0;
#endif
}

// length(v) and normalize(v) of a light vector (glsl:463-466, :477-479). Parity build: the shader's
// sqrt(dot) and v * (1 / sqrt(dot)), IEEE. Fast build: r = rsqrt(dot), length = dot * r, unit = v * r.
__device__ __forceinline__ void unit_and_length(f3& v, float& len) {
#if RT_PARITY
  len = sqrt_ieee(dot3(v, v));
  v = normalize_ieee(v);
#else
  const float dd = sdot3(v, v);
  const float r = inversesqrt_shade(dd);
  len = dd * r;
  v = v * r;
#endif
}
__device__ __forceinline__ f3 normalize_light(f3 v) {
#if RT_PARITY
  return normalize_ieee(v);
#else
  return normalize_color(v);
#endif
}

// computeLighting (glsl:457-507). The BRDF term is evaluated before the shadow factor so that
// lights which cannot contribute skip their shadow rays (Lo + 0*finite == Lo exactly).
// Every lane of the CTA walks the light loop (lit = this lane has a surface to shade).
template <bool BIG, bool FAN>
__device__ __forceinline__ f3 compute_lighting(const Scene& s, saddr cand_base, int coh, bool lit, float noise_r, f3 P, f3 N, const Mat& mat, f3 V) {
  f3 Lo = mk3(0.0f, 0.0f, 0.0f);
  const f3 shadow_origin = smad3(N, 0.001f, P);          // glsl:381, :412 (P + N * 0.001)
#pragma unroll 1
  for (int i = 0; i < s.nl; ++i) {
    phase_sync(coh >= 3);
    if (!lit) continue;
    const LightRec light = load_light(s, i);
    f3 lightDir = mk3(0.0f, 0.0f, 0.0f);
    float attenuation = 1.0f, lightDistance = 0.0f;
    // lightDir / lightDistance feed the BRDF (colour) and the shadow rays (a boolean each), never P, N
    // or a bounce direction: the fast build takes length and unit vector from one MUFU.RSQ.
    if (light.type == 0) {
      lightDir = light.position - P;
      unit_and_length(lightDir, lightDistance);
      attenuation = div_shade(1.0f, smad(0.01f * lightDistance, lightDistance, smad(0.1f, lightDistance, 1.0f)));
    } else if (light.type == 1) {
      lightDir = normalize_light(-light.direction);
      lightDistance = 1e6f;
    } else if (light.type == 2) {
      lightDir = light.position - P;
      unit_and_length(lightDir, lightDistance);
      attenuation = div_shade(1.0f, lightDistance * lightDistance);
      const f3 lightNormal = normalize_color(light.direction);
      attenuation *= fmaxf(sdot3(lightDir, lightNormal), 0.0f);   // P->light vs facing
    }
#if RT_PARITY
    const f3 L = normalize_shade(lightDir);              // glsl:493: normalises the unit vector once more
#else
    const f3 L = lightDir;                               // (a second normalisation moves it by <= 1 ulp)
#endif
    const f3 H = normalize_color(V + L);
    const f3 radiance = (light.color * attenuation) * light.intensity;
    const f3 brdf = compute_pbr(mat, N, V, L, H, radiance);
    const bool sampled = light.shadowType == 1 || light.shadowType == 2;
    // exactly zero (not NaN) and the shadow factor is finite (pcfSamples == 0 gives 0/0)
    if (brdf.x == 0.0f && brdf.y == 0.0f && brdf.z == 0.0f && (!sampled || light.pcfSamples > 0)) continue;
    float shadowFactor = 0.0f;                           // calculateShadow (glsl:442-455)
    if (light.shadowType == 0) shadowFactor = 1.0f;
    else if (sampled) shadowFactor = shadow_packets<BIG, FAN>(s, cand_base, noise_r, shadow_origin, light, i, lightDir, lightDistance,
                                                              RT_PK_COOP >= 1 && __activemask() == 0xffffffffu);
    Lo = smad3(brdf, shadowFactor, Lo);
  }
  if (lit && mat.subsurfaceScatter > 0.0f) {
    const Scene s2 = s; const Mat mat2 = mat;                                           // copies: see trace_closest
    Lo = Lo + subsurface(s2, P, N, mat2);
  }
  return Lo;
}

// ---------------------------------------------------------------- main (glsl:509-584)
struct PixelResult { f3 color, P, N; int object_id; };

// One invocation of the shader's main() per lane. Whole warps (coh: the whole CTA) call this
// together; `alive` = this lane still has a path to extend (false from the start for lanes
// outside the image). The bounce loop ends when no lane of the warp / CTA is alive.
template <bool BIG, bool FAN>
__device__ __forceinline__ PixelResult shade_pixel(const KernelParams& kp, const Scene& s, const Textures& tx, saddr cand_base, int coh,
                                                   bool alive, unsigned px, unsigned py, int frame_count) {
  const rt_frame_params& fp = kp.fp;
  const float noise_r = fetch_noise(tx, (float)(px + (unsigned)frame_count) * fp.noise_scale[0],
                                    (float)(py + (unsigned)frame_count) * fp.noise_scale[1]);
  // generateCameraRay (glsl:198-217); jitter = texel.xy*2-1 with .y = 0 for an R8 texture.
  const float jx = noise_r * 2.0f - 1.0f, jy = 0.0f * 2.0f - 1.0f;
  float ux = div_ieee((float)(int)px + 0.5f + jx, (float)fp.width);
  float uy = div_ieee((float)(int)py + 0.5f + jy, (float)fp.height);
  ux = ux * 2.0f - 1.0f;
  uy = uy * 2.0f - 1.0f;
  ux *= kp.aspect * kp.tan_fov * fp.focal_length;
  uy *= kp.tan_fov * fp.focal_length;
  const f3 camDir = mk3(fp.camera_dir[0], fp.camera_dir[1], fp.camera_dir[2]);
  const f3 camRight = mk3(fp.camera_right[0], fp.camera_right[1], fp.camera_right[2]);
  const f3 camUp = mk3(fp.camera_up[0], fp.camera_up[1], fp.camera_up[2]);
  f3 ro = mk3(fp.camera_pos[0], fp.camera_pos[1], fp.camera_pos[2]);
  f3 rd = normalize_ieee((camDir + ux * camRight) + uy * camUp);

  PixelResult out;
  out.color = mk3(0.0f, 0.0f, 0.0f);
  out.P = mk3(0.0f, 0.0f, 0.0f);               // P = N = 0 on a primary miss
  out.N = mk3(0.0f, 0.0f, 0.0f);
  out.object_id = -1;
  f3 throughput = mk3(1.0f, 1.0f, 1.0f);

#pragma unroll 1
  for (int depth = 0; depth < fp.max_ray_depth; ++depth) {
    if (!(coh >= 2 ? phase_any(alive) : __any_sync(0xffffffffu, alive) != 0u)) break;
    bool lit = false;
    f3 N = mk3(0.0f, 0.0f, 0.0f), P = N, V = N;
    Mat mat = {};
    if (alive) {
      const Hit h = trace_closest<BIG>(s, cand_base, ro, rd);
      if (h.code < 0) {                         // miss path (glsl:530-534)
        const f3 sky = fp.use_skybox ? fetch_sky(tx, rd) : mk3(0.0f, 0.0f, 0.0f);
        out.color = smad3(throughput, sky, out.color);
        alive = false;
      } else {
        lit = true;
        const int index = hit_object_index(s, h.code);
        if (h.code < s.ns) N = normalize_ieee((ro + rd * h.t) - xyz(lds128(s.sph + (unsigned)h.code * kSphereBytes + 48u)));   // glsl:186-191
        else N = xyz(lds128(s.pln + (unsigned)(h.code - s.ns) * kPlaneBytes + 48u));
        mat = load_material(s, index);
        P = ro + rd * h.t;
        V = normalize_color(-rd);
        out.P = P; out.N = N; out.object_id = index;
      }
    }
    const f3 Lo = compute_lighting<BIG, FAN>(s, cand_base, coh, lit, noise_r, P, N, mat, V);
    if (!lit) continue;
    out.color = smad3(throughput, Lo, out.color);

    if (depth > 2) {                            // Russian roulette (glsl:544-549)
      const float diffuseWeight = sqrt_ieee(dot3(mat.albedo, mat.albedo)) * mat.diffuseStrength;
      const float continueProb =
          fminf(fmaxf(throughput.x, fmaxf(throughput.y, throughput.z)) * 0.95f + diffuseWeight, 0.99f);
      if (random2((float)(px + (unsigned)depth), (float)(py + (unsigned)depth)) > continueProb) { alive = false; continue; }
      throughput = mk3(div_ieee(throughput.x, continueProb), div_ieee(throughput.y, continueProb),
                       div_ieee(throughput.z, continueProb));
    }
    if (depth + 1 >= fp.max_ray_depth) { alive = false; continue; }   // the next ray would never be traced

    const float F = fresnel_schlick(fmaxf(dot3(V, N), 0.0f), mat.ior);
    if (mat.diffuseStrength > 0.0f) {           // glsl:555-567
      const int hi = depth * 64 + frame_count;
      const f3 specularDir = reflect3(rd, N);
      const f3 diffuseDir = cosine_hemisphere(div_ieee((float)hi, 64.0f), halton(hi, 2), N);
      rd = normalize_ieee(mix3(specularDir, diffuseDir, mat.roughness));
      ro = P + N * 0.001f;
      throughput = throughput * (mat.albedo * mat.diffuseStrength);
    } else if (mat.transparency > 0.0f) {       // glsl:568-571
      rd = refraction_dir(rd, N, mat.ior);
      ro = P - N * 0.001f;
      throughput = throughput * ((mat.albedo * (1.0f - F)) * mat.transparency);
    } else {                                    // glsl:572-576
      rd = reflect3(rd, N);
      ro = P + N * 0.001f;
      throughput = throughput * (mat.albedo * F);
    }
  }
  return out;
}

// ---------------------------------------------------------------- kernel
#ifndef RT_PK_MINBLOCKS
#define RT_PK_MINBLOCKS 2
#endif
// BIG = false: small scenes (CTA-coherent phases, no cluster code in the instruction stream);
// BIG = true: large scenes (independent warps, two-level boxes). One source, two compact kernels:
// on C2 every extra KB of hot code costs (measured +9 % when the cluster loops were compiled in).
// BANDS = true: interleaved row bands (rt_frame_params.row_band > 0). A separate instantiation
// because the register allocation of the whole kernel moves with this one expression (measured
// on C2: 0.571 ms without the band arithmetic compiled in, 0.626 ms with it).
// MINB = resident CTAs per SM the register budget is sized for: 2 (128 registers, 16 warps) or, for
// the large-scene kernel, also 3 (80 registers, 24 warps, more spills). Same source, same
// arithmetic (-fmad=false: nothing contracts differently), bit-identical images; which one is
// faster depends on the scene (B200: C3 39.5 -> 35.1 ms with 3, C5 10.9 -> 12.0 ms), so the host
// times both on the first frames of a scene and keeps the faster (rt_abi.cu).
// RT_PK_MINCTAS (dev A/B): resident CTAs per SM the small-scene kernel's registers are sized for.
#ifdef RT_PK_MINCTAS
#define PK_MIN_CTAS(big, minb) ((big) ? (minb) * 8 / RT_PK_WARPS : RT_PK_MINCTAS)
#else
#define PK_MIN_CTAS(big, minb) ((minb) * 8 / RT_PK_WARPS)
#endif
// FAN = true: the fast build's analytic PCSS probe classification (probe_fan) is compiled in; chosen
// by the host when the scene has a PCSS light and every sphere's box contains its sphere. A
// separate instantiation because the mere presence of that code costs the small-scene kernel 3.5 %
// on scenes that never run it (C1 0.1075 -> 0.111 ms, same box).
template <bool BIG, bool BANDS, int MINB, bool FAN>
__global__ void __launch_bounds__(kThreads, PK_MIN_CTAS(BIG, MINB)) rt_packet_kernel(const __grid_constant__ KernelParams kp) {
  extern __shared__ float4 smem[];
  const DeviceScene& sc = kp.sc;
  const rt_frame_params& fp = kp.fp;
#if RT_PK_STATS
  g_stats = kp.counters;
#endif

  // ---- stage the scene once per CTA: one contiguous copy of the host-built blob
  for (int i = threadIdx.x; i < sc.pk_float4s; i += kThreads) smem[i] = __ldg(sc.pk_blob + i);
  __syncthreads();
  Scene s;
  const saddr base = (saddr)__cvta_generic_to_shared(smem);
  s.clu = base;
  s.sph = base + (unsigned)sc.pk_num_clusters * kSphereBytes;
  s.pln = s.sph + (unsigned)sc.pk_num_spheres * kSphereBytes;
  s.sph_index = s.pln + (unsigned)sc.pk_num_planes * kPlaneBytes;
  s.lights = base + (unsigned)sc.pk_lights_offset;
  s.material = sc.material;
  s.material_smem = (!BIG && sc.pk_material_offset >= 0) ? base + (unsigned)sc.pk_material_offset : 0u;
  s.nc = BIG ? sc.pk_num_clusters : 0; s.ns = sc.pk_num_spheres; s.np = sc.pk_num_planes; s.nl = sc.num_lights;
  s.maxd = fp.max_ray_distance;
  s.grid = BIG ? sc.pk_grid : nullptr; s.grid_off_f4 = sc.pk_grid_off_f4; s.grid_ent_f4 = sc.pk_grid_ent_f4;
  // per-thread candidate list behind the scene image: entry i of thread t at +(i*kThreads + t)*2
  const saddr cand_base = base + (unsigned)sc.pk_float4s * 16u + threadIdx.x * 2u;
  s.wlist = base + (unsigned)sc.pk_float4s * 16u + (unsigned)(kCandCap * kThreads * 2) + (threadIdx.x >> 5) * (unsigned)kWarpListBytes;

  Textures tx;
  tx.sky = sc.sky; tx.sky_tex = sc.sky_tex; tx.sky_size = sc.sky_size;
  tx.sky_hw = (fp.flags & RT_FLAG_SKYBOX_HW) != 0u && sc.sky_tex != 0ull;
  tx.noise = sc.noise; tx.noise_w = sc.noise_w; tx.noise_h = sc.noise_h;

  const int lane = threadIdx.x & 31;
  const int lx = lane & (kTileW - 1), ly = lane >> 3;
  const unsigned num_tiles = kp.tiles_x * kp.tiles_y;
  const int nsamp = fp.num_samples > 0 ? fp.num_samples : 1;
  const float inv_ns = rcp_ieee((float)nsamp);

  // Two work distributions over the same 8x4-pixel warp tiles (kp.sc.pk_coherent, host-chosen):
  //  * coherent (small scenes): the CTA takes 8 tiles forming a 32x8-pixel block and walks the
  //    shader in phases separated by CTA barriers, so that the warps sharing an SM sub-partition
  //    fetch the same instructions at about the same time. Measured on B200 (C2): the
  //    independent-warp loop is instruction-fetch bound (ncu: 62 % of warp samples stalled on
  //    no_instruction), tile-level coherence alone gave +52 % throughput.
  //  * independent (large scenes): every warp pulls its own tiles from the global counter; the
  //    object loops are long, cheap sky tiles and expensive reflective tiles balance dynamically.
  const int coh = sc.pk_coherent;   // 0 independent warps; 1 CTA blocks; 2 + barrier per bounce; 3 + barrier per light
  __shared__ unsigned s_tile;
  unsigned tile = 0, next = 0;
  if (!coh) {
    if (lane == 0) tile = atomicAdd(kp.tile_counter, 1u) - kp.tile_base;
    tile = __shfl_sync(0xffffffffu, tile, 0);
  }
  for (;;) {
    if (coh) {
      __syncthreads();
      if (threadIdx.x == 0) s_tile = atomicAdd(kp.tile_counter, 1u) - kp.tile_base;
      __syncthreads();
      const unsigned blk = s_tile;                       // block id: 4x2 tiles = 32x8 pixels
      if (blk >= kp.blocks_x * kp.blocks_y) break;
      const unsigned by = blk / kp.blocks_x, bx = blk - by * kp.blocks_x;
      const unsigned w = threadIdx.x >> 5;
      const unsigned tyi = by * (unsigned)kBlockTilesY + (w >> 2), txi = bx * (unsigned)kBlockTilesX + (w & 3u);
      tile = (tyi < kp.tiles_y && txi < kp.tiles_x) ? tyi * kp.tiles_x + txi : num_tiles;
    } else {
      if (tile >= num_tiles) break;
      if (lane == 0) next = atomicAdd(kp.tile_counter, 1u) - kp.tile_base;   // prefetch the next tile id
    }
    const unsigned ty = tile / kp.tiles_x, txi = tile - ty * kp.tiles_x;
    const int px = (int)txi * kTileW + lx;
    const int py = (BANDS ? window_row(fp, (int)ty * kTileH) : fp.row_begin + (int)ty * kTileH) + ly;
    const bool inside = tile < num_tiles && px < fp.width && py < fp.row_end;
    PixelResult r;                                       // position/normal/id: those of sample 0
#pragma unroll 1
    for (int sidx = 0; sidx < nsamp; ++sidx) {           // colour: sequential FP32 sum over samples
      const PixelResult r2 = shade_pixel<BIG, FAN>(kp, s, tx, cand_base, coh, inside, (unsigned)px, (unsigned)py, fp.frame_count + sidx);
      if (sidx == 0) r = r2;
      else r.color = r.color + r2.color;
    }
    if (inside) {
      const size_t at = (size_t)py * fp.width + px;
      if (kp.color) {
        float4 c = make_float4(r.color.x, r.color.y, r.color.z, 1.0f);
        if (fp.flags & RT_FLAG_ACCUMULATE) {             // partial sums for sample-range sharding
          const float4 old = kp.color[at];
          c = make_float4(old.x + c.x, old.y + c.y, old.z + c.z, 1.0f);
        } else if (nsamp > 1 && !(fp.flags & RT_FLAG_SUM_SAMPLES)) {
          c = make_float4(c.x * inv_ns, c.y * inv_ns, c.z * inv_ns, 1.0f);
        }
        kp.color[at] = c;
      }
      if (kp.position) kp.position[at] = make_float4(r.P.x, r.P.y, r.P.z, 1.0f);
      if (kp.normal) {
        if (fp.flags & RT_FLAG_NORMAL_FP32) reinterpret_cast<float4*>(kp.normal)[at] = make_float4(r.N.x, r.N.y, r.N.z, 1.0f);
        else reinterpret_cast<uint2*>(kp.normal)[at] = pack_half4(r.N.x, r.N.y, r.N.z, 1.0f);
      }
      if (kp.object_id) kp.object_id[at] = r.object_id;
    }
    if (!coh) tile = __shfl_sync(0xffffffffu, next, 0);
  }
}

inline cudaError_t upload_probe_order(const float* halton3, const int* order16) {
  float rv[16];
  for (int i = 0; i < 16; ++i) rv[i] = halton3[order16[i]] * 2.0f - 1.0f;      // host TU: -ffp-contract=off
  return cudaMemcpyToSymbol(c_probe_rv, rv, sizeof(rv));
}

template <bool BIG, bool BANDS, int MINB, bool FAN>
inline cudaError_t launch_packet_t(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched) {
  const size_t smem = (size_t)kp.sc.pk_float4s * sizeof(float4) + (size_t)kScratchBytes;
  int per_sm = 0;
  cudaError_t e = cached_occupancy(rt_packet_kernel<BIG, BANDS, MINB, FAN>, kThreads, smem, &per_sm);
  if (e != cudaSuccess) return e;
  const unsigned num_tiles = kp.tiles_x * kp.tiles_y;
  unsigned ctas = (unsigned)(sm_count * per_sm);
  const unsigned needed = (num_tiles + kWarpsPerCta - 1) / kWarpsPerCta;
  if (ctas > needed) ctas = needed;
  if (ctas < 1) ctas = 1;
  // how far the work counter advances during this launch (reported through *warps_launched
  // as "advance - num_tiles", the form rt_render adds back): independent mode = one fetch per
  // tile + one failing fetch per warp; coherent mode = one per block + one per CTA.
  if (kp.sc.pk_coherent) *warps_launched = kp.blocks_x * kp.blocks_y + ctas - num_tiles;
  else *warps_launched = ctas * kWarpsPerCta;
  rt_packet_kernel<BIG, BANDS, MINB, FAN><<<ctas, kThreads, smem, stream>>>(kp);
  return cudaGetLastError();
}

// kp.sc.pk_big (host-chosen by object count) selects the kernel; kp.sc.pk_fan_ok the FAN instantiation
// (fast build only: the parity build has no analytic path and instantiates FAN = false alone).
template <bool BANDS, bool FAN>
inline cudaError_t launch_packet_f(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched) {
  constexpr int kMinB = RT_PK_MINBLOCKS;
  if (!kp.sc.pk_big) return launch_packet_t<false, BANDS, kMinB, FAN>(kp, sm_count, stream, warps_launched);
  return kp.sc.pk_dense ? launch_packet_t<true, BANDS, 3, FAN>(kp, sm_count, stream, warps_launched)
                        : launch_packet_t<true, BANDS, kMinB, FAN>(kp, sm_count, stream, warps_launched);
}
inline cudaError_t launch_packet(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched) {
#if !RT_PARITY
  if (kp.sc.pk_fan_ok)
    return kp.fp.row_band > 0 ? launch_packet_f<true, true>(kp, sm_count, stream, warps_launched)
                              : launch_packet_f<false, true>(kp, sm_count, stream, warps_launched);
#endif
  return kp.fp.row_band > 0 ? launch_packet_f<true, false>(kp, sm_count, stream, warps_launched)
                            : launch_packet_f<false, false>(kp, sm_count, stream, warps_launched);
}

}  // namespace pk
}  // namespace RT_VARIANT_NS
}  // namespace rt
