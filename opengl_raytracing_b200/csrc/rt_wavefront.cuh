// rt_wavefront.cuh — the wavefront variant of the ray-tracing pass (/root/reference/shader/raytracingCs.glsl):
// the SAME per-pixel arithmetic as the persistent megakernel of rt_packet.cuh (it calls that file's trace,
// shadow-job and BRDF routines), scheduled differently. The shader's bounce loop (glsl:519-579) diverges at
// three places — miss -> break (:526-534), Russian roulette (:544-549), the reflect / refract / diffuse
// branch (:555-576) — and its light loop (:462-504) at a fourth (which lights need shadow rays, and how many
// candidates their cones keep). In the megakernel a lane whose path ended, or whose light needs no rays,
// idles until its warp / CTA votes the phase over. Here the rays that survive a bounce are regrouped:
//
//   bounce(d)  one lane per path alive after depth d-1: accumulate that hit's lighting from the (BRDF term,
//              shadow factor) records, Russian roulette, next ray, closest hit. Paths that end are written to
//              the images there and then; the hits are COMPACTED — __ballot_sync + prefix popcount, one
//              atomicAdd per warp — into the hit queue of depth d, and every (hit, light) that needs shadow
//              rays is compacted the same way into that light's job queue.
//   shadow(d)  one lane per shadow job, light by light (all lanes of a warp run the same light's code):
//              shadow_packets of rt_packet.cuh, the factor stored next to the job's BRDF term.
//
// so every warp of every launch starts full, and each kernel has its own register budget (the shadow jobs do
// not carry a path's state, the bounce kernel does not carry the packets'). Bit-identical to the megakernel in
// both math builds by construction: the per-light terms are added in light order (glsl:502), the samples of a
// multi-sample render in sample order, exactly as there (tests/test_wavefront_gpu.py).
#pragma once
#include "rt_packet.cuh"

namespace rt {
namespace RT_VARIANT_NS {
namespace pk {

#ifndef RT_WF_BOUNCE_MINB
#define RT_WF_BOUNCE_MINB 3     // resident CTAs per SM the bounce kernel's registers are sized for (3 -> 80 registers)
#endif
#ifndef RT_WF_SHADOW_MINB
#define RT_WF_SHADOW_MINB 3     // ... and the shadow kernel's, small scenes
#endif
// large scenes: 2 (128 registers). Measured on B200 (ms, 3 / 2 / 4 CTAs): C3 13.40 / 13.21 / 14.25, C5 7.73 / 7.54 / 7.85;
// small scenes do not care (C2 0.4927 / 0.4907 / 0.5151). The bounce kernel stays at 3 (2: C3 14.13, C5 7.73).
#ifndef RT_WF_SHADOW_COOP
#define RT_WF_SHADOW_COOP 0
#endif
#ifndef RT_WF_SHADOW_MINB_BIG
#define RT_WF_SHADOW_MINB_BIG 2
#endif

// The scene image in shared memory, as rt_packet_kernel stages it.
template <bool BIG>
__device__ __forceinline__ void wf_stage_scene(const KernelParams& kp, float4* smem, Scene& s, saddr& cand_base, Textures& tx) {
  const DeviceScene& sc = kp.sc;
#if RT_PK_STATS
  g_stats = kp.counters;
#endif
  for (int i = threadIdx.x; i < sc.pk_float4s; i += kThreads) smem[i] = __ldg(sc.pk_blob + i);
  __syncthreads();
  const saddr base = (saddr)__cvta_generic_to_shared(smem);
  s.clu = base;
  s.sph = base + (unsigned)sc.pk_num_clusters * kSphereBytes;
  s.pln = s.sph + (unsigned)sc.pk_num_spheres * kSphereBytes;
  s.sph_index = s.pln + (unsigned)sc.pk_num_planes * kPlaneBytes;
  s.lights = base + (unsigned)sc.pk_lights_offset;
  s.material = sc.material;
  s.material_smem = (!BIG && sc.pk_material_offset >= 0) ? base + (unsigned)sc.pk_material_offset : 0u;
  s.nc = BIG ? sc.pk_num_clusters : 0; s.ns = sc.pk_num_spheres; s.np = sc.pk_num_planes; s.nl = sc.num_lights;
  s.maxd = kp.fp.max_ray_distance;
  s.grid = BIG ? sc.pk_grid : nullptr; s.grid_off_f4 = sc.pk_grid_off_f4; s.grid_ent_f4 = sc.pk_grid_ent_f4;
  cand_base = base + (unsigned)sc.pk_float4s * 16u + threadIdx.x * 2u;
  s.wlist = base + (unsigned)sc.pk_float4s * 16u + (unsigned)(kCandCap * kThreads * 2) + (threadIdx.x >> 5) * (unsigned)kWarpListBytes;
  tx.sky = sc.sky; tx.sky_tex = sc.sky_tex; tx.sky_size = sc.sky_size;
  tx.sky_hw = (kp.fp.flags & RT_FLAG_SKYBOX_HW) != 0u && sc.sky_tex != 0ull;
  tx.noise = sc.noise; tx.noise_w = sc.noise_w; tx.noise_h = sc.noise_h;
}

// Light vector of one (shading point, light): computeLighting's head (glsl:463-489), the same routines as
// compute_lighting in rt_packet.cuh so that the bounce kernel (BRDF) and the shadow kernel (rays) see the
// same bits. Returns false for an unknown light type (lightDir stays 0, glsl:460).
__device__ __forceinline__ void wf_light_vector(const LightRec& light, f3 P, f3& lightDir, float& lightDistance, float& attenuation) {
  lightDir = mk3(0.0f, 0.0f, 0.0f);
  attenuation = 1.0f; lightDistance = 0.0f;
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
}

// The pixel of a finished path: main()'s three stores (glsl:581-583) with the sample bookkeeping of
// rt_packet_kernel's epilogue — colour of sample s is added to the running sum in sample order, position /
// normal / id are those of sample 0, the last sample applies ACCUMULATE / the 1/n mean and stores.
__device__ __forceinline__ void wf_finish_pixel(const KernelParams& kp, const WfParams& wp, unsigned pxy, f3 color, f3 P, f3 N, int object_id) {
  const rt_frame_params& fp = kp.fp;
  const int px = (int)(pxy & 0xffffu), py = (int)(pxy >> 16);
  const size_t at = (size_t)py * fp.width + px;
  const int nsamp = fp.num_samples > 0 ? fp.num_samples : 1;
  if (wp.sample == 0) {
    if (kp.position) kp.position[at] = make_float4(P.x, P.y, P.z, 1.0f);
    if (kp.normal) {
      if (fp.flags & RT_FLAG_NORMAL_FP32) reinterpret_cast<float4*>(kp.normal)[at] = make_float4(N.x, N.y, N.z, 1.0f);
      else reinterpret_cast<uint2*>(kp.normal)[at] = pack_half4(N.x, N.y, N.z, 1.0f);
    }
    if (kp.object_id) kp.object_id[at] = object_id;
  }
  if (nsamp > 1) {
    if (wp.sample > 0) { const float4 a = wp.acc[at]; color = mk3(a.x, a.y, a.z) + color; }     // r.color + r2.color
    if (wp.sample < nsamp - 1) { wp.acc[at] = make_float4(color.x, color.y, color.z, 1.0f); return; }
  }
  if (kp.color) {
    float4 c = make_float4(color.x, color.y, color.z, 1.0f);
    if (fp.flags & RT_FLAG_ACCUMULATE) {
      const float4 old = kp.color[at];
      c = make_float4(old.x + c.x, old.y + c.y, old.z + c.z, 1.0f);
    } else if (nsamp > 1 && !(fp.flags & RT_FLAG_SUM_SAMPLES)) {
      const float inv_ns = rcp_ieee((float)nsamp);
      c = make_float4(c.x * inv_ns, c.y * inv_ns, c.z * inv_ns, 1.0f);
    }
    kp.color[at] = c;
  }
}

template <bool BIG, bool BANDS>
__global__ void __launch_bounds__(kThreads, RT_WF_BOUNCE_MINB * 8 / RT_PK_WARPS)
wf_bounce_kernel(const __grid_constant__ KernelParams kp, const __grid_constant__ WfParams wp) {
  extern __shared__ float4 smem[];
  const rt_frame_params& fp = kp.fp;
  Scene s; saddr cand_base; Textures tx;
  wf_stage_scene<BIG>(kp, smem, s, cand_base, tx);

  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  const int depth = wp.depth;
  const bool first = depth == 0;
  const bool shade_only = depth >= fp.max_ray_depth;      // bounce(D): the hits of the last depth are shaded, nothing is traced
  const unsigned cap = wp.cap;
  // queues are ping-ponged by depth parity: this launch reads the records of depth d-1 and writes those of depth d
  const float4* __restrict__ qin = wp.path[(depth + 1) & 1];
  float4* __restrict__ qout = wp.path[depth & 1];
  const float4* __restrict__ hin = wp.hit[(depth + 1) & 1];
  float4* __restrict__ hout = wp.hit[depth & 1];
  const float4* __restrict__ bin = wp.brdf[(depth + 1) & 1];
  float4* __restrict__ bout = wp.brdf[depth & 1];
  const unsigned n_in = first ? wp.tile_rows * kp.tiles_x * 32u : wp.counters[wf_c_nq(depth - 1)];
  const int frame_count = fp.frame_count + wp.sample;
  unsigned* work = wp.counters + wf_c_work(depth);
  unsigned* nq = wp.counters + wf_c_nq(depth);

  for (;;) {
    unsigned base = 0;
    if (lane == 0) base = atomicAdd(work, 32u);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= n_in) break;
    const unsigned i = base + (unsigned)lane;

    bool alive = false;
    unsigned pxy = 0u;
    float noise_r = 0.0f;
    f3 ro = mk3(0.0f, 0.0f, 0.0f), rd = ro, throughput = mk3(1.0f, 1.0f, 1.0f), color = ro;
    if (first) {
      // ---- generateCameraRay (glsl:198-217), as shade_pixel does it; one warp = one 8x4-pixel tile
      const unsigned tile = i >> 5;
      const unsigned ty = tile / kp.tiles_x, txi = tile - ty * kp.tiles_x;
      const int px = (int)txi * kTileW + (lane & (kTileW - 1));
      const int lr = (int)(ty + wp.tile_row0) * kTileH;
      const int py = (BANDS ? window_row(fp, lr) : fp.row_begin + lr) + (lane >> 3);
      if (px < fp.width && py < fp.row_end) {
        pxy = (unsigned)px | ((unsigned)py << 16);
        noise_r = fetch_noise(tx, (float)((unsigned)px + (unsigned)frame_count) * fp.noise_scale[0],
                              (float)((unsigned)py + (unsigned)frame_count) * fp.noise_scale[1]);
        const float jx = noise_r * 2.0f - 1.0f, jy = 0.0f * 2.0f - 1.0f;
        float ux = div_ieee((float)px + 0.5f + jx, (float)fp.width);
        float uy = div_ieee((float)py + 0.5f + jy, (float)fp.height);
        ux = ux * 2.0f - 1.0f;
        uy = uy * 2.0f - 1.0f;
        ux *= kp.aspect * kp.tan_fov * fp.focal_length;
        uy *= kp.tan_fov * fp.focal_length;
        const f3 camDir = mk3(fp.camera_dir[0], fp.camera_dir[1], fp.camera_dir[2]);
        const f3 camRight = mk3(fp.camera_right[0], fp.camera_right[1], fp.camera_right[2]);
        const f3 camUp = mk3(fp.camera_up[0], fp.camera_up[1], fp.camera_up[2]);
        ro = mk3(fp.camera_pos[0], fp.camera_pos[1], fp.camera_pos[2]);
        rd = normalize_ieee((camDir + ux * camRight) + uy * camUp);
        alive = true;
        if (shade_only) {                              // max_ray_depth == 0: the loop body never runs (glsl:519)
          wf_finish_pixel(kp, wp, pxy, color, mk3(0.0f, 0.0f, 0.0f), mk3(0.0f, 0.0f, 0.0f), -1);
          alive = false;
        }
      }
    } else if (i < n_in) {
      // ---- the tail of the loop body for the hit of depth d-1 (glsl:537-578)
      const int pd = depth - 1;
      const float4 q0 = qin[i], q1 = qin[cap + i], q2 = qin[2u * cap + i];
      const float4 h0 = hin[i], h1 = hin[cap + i];
      rd = xyz(q0); pxy = __float_as_uint(q0.w);
      throughput = xyz(q1); noise_r = q1.w;
      color = xyz(q2);
      const f3 P = xyz(h0), N = xyz(h1);
      const int index = __float_as_int(h0.w);
      unsigned lmask = __float_as_uint(h1.w);
      const Mat mat = load_material(s, index);
      f3 Lo = mk3(0.0f, 0.0f, 0.0f);
      while (lmask) {                                  // Lo += BRDF * shadow, in light order (glsl:502)
        const int l = __ffs(lmask) - 1;
        lmask &= lmask - 1u;
        const float4 b = bin[(size_t)l * cap + i];
        Lo = smad3(xyz(b), b.w, Lo);
      }
      if (mat.subsurfaceScatter > 0.0f) {
        const Scene s2 = s; const Mat mat2 = mat;      // copies: see trace_closest
        Lo = Lo + subsurface(s2, P, N, mat2);
      }
      color = smad3(throughput, Lo, color);
      bool cont = true;
      const unsigned px = pxy & 0xffffu, py = pxy >> 16;
      if (pd > 2) {                                    // Russian roulette (glsl:544-549)
        const float diffuseWeight = sqrt_ieee(dot3(mat.albedo, mat.albedo)) * mat.diffuseStrength;
        const float continueProb =
            fminf(fmaxf(throughput.x, fmaxf(throughput.y, throughput.z)) * 0.95f + diffuseWeight, 0.99f);
        if (random2((float)(px + (unsigned)pd), (float)(py + (unsigned)pd)) > continueProb) cont = false;
        else throughput = mk3(div_ieee(throughput.x, continueProb), div_ieee(throughput.y, continueProb),
                              div_ieee(throughput.z, continueProb));
      }
      if (pd + 1 >= fp.max_ray_depth) cont = false;   // the next ray would never be traced
      if (cont) {
        const f3 V = normalize_color(-rd);
        const float F = fresnel_schlick(fmaxf(dot3(V, N), 0.0f), mat.ior);
        if (mat.diffuseStrength > 0.0f) {             // glsl:555-567
          const int hi = pd * 64 + frame_count;
          const f3 specularDir = reflect3(rd, N);
          const f3 diffuseDir = cosine_hemisphere(div_ieee((float)hi, 64.0f), halton(hi, 2), N);
          rd = normalize_ieee(mix3(specularDir, diffuseDir, mat.roughness));
          ro = P + N * 0.001f;
          throughput = throughput * (mat.albedo * mat.diffuseStrength);
        } else if (mat.transparency > 0.0f) {         // glsl:568-571
          rd = refraction_dir(rd, N, mat.ior);
          ro = P - N * 0.001f;
          throughput = throughput * ((mat.albedo * (1.0f - F)) * mat.transparency);
        } else {                                      // glsl:572-576
          rd = reflect3(rd, N);
          ro = P + N * 0.001f;
          throughput = throughput * (mat.albedo * F);
        }
        alive = true;
      } else {
        wf_finish_pixel(kp, wp, pxy, color, P, N, index);
      }
    }
    if (shade_only) continue;

    // ---- closest hit of the new ray (glsl:522-534)
    bool hit = false;
    Hit h; h.code = -1; h.t = 0.0f;
    if (alive) {
      h = trace_closest<BIG>(s, cand_base, ro, rd);
      if (h.code < 0) {
        const f3 sky = fp.use_skybox ? fetch_sky(tx, rd) : mk3(0.0f, 0.0f, 0.0f);
        color = smad3(throughput, sky, color);
        f3 P = mk3(0.0f, 0.0f, 0.0f), N = P;          // P = N = 0 on a primary miss; else those of the last hit
        int index = -1;
        if (!first) {
          const float4 h0 = hin[i], h1 = hin[cap + i];
          P = xyz(h0); N = xyz(h1); index = __float_as_int(h0.w);
        }
        wf_finish_pixel(kp, wp, pxy, color, P, N, index);
      } else {
        hit = true;
      }
    }

    // ---- regroup: the surviving paths of this warp take consecutive slots of the hit queue
    const unsigned hm = __ballot_sync(0xffffffffu, hit);
    if (!hm) continue;
    unsigned slot = 0;
    if (lane == 0) slot = atomicAdd(nq, (unsigned)__popc(hm));
    slot = __shfl_sync(0xffffffffu, slot, 0) + (unsigned)__popc(hm & lt);

    f3 P = mk3(0.0f, 0.0f, 0.0f), N = P, V = P;
    Mat mat = {};
    int index = -1;
    if (hit) {
      index = hit_object_index(s, h.code);
      if (h.code < s.ns) N = normalize_ieee((ro + rd * h.t) - xyz(lds128(s.sph + (unsigned)h.code * kSphereBytes + 48u)));   // glsl:186-191
      else N = xyz(lds128(s.pln + (unsigned)(h.code - s.ns) * kPlaneBytes + 48u));
      mat = load_material(s, index);
      P = ro + rd * h.t;
      V = normalize_color(-rd);
      qout[slot] = make_float4(rd.x, rd.y, rd.z, __uint_as_float(pxy));
      qout[cap + slot] = make_float4(throughput.x, throughput.y, throughput.z, noise_r);
      qout[2u * cap + slot] = make_float4(color.x, color.y, color.z, 0.0f);
    }
    // ---- computeLighting's per-light BRDF terms (glsl:462-504); lights that need rays become shadow jobs
    unsigned lmask = 0u;
#pragma unroll 1
    for (int l = 0; l < s.nl; ++l) {
      bool job = false;
      if (hit) {
        const LightRec light = load_light(s, l);
        f3 lightDir; float lightDistance, attenuation;
        wf_light_vector(light, P, lightDir, lightDistance, attenuation);
#if RT_PARITY
        const f3 L = normalize_shade(lightDir);              // glsl:493: normalises the unit vector once more
#else
        const f3 L = lightDir;
#endif
        const f3 H = normalize_color(V + L);
        const f3 radiance = (light.color * attenuation) * light.intensity;
        const f3 brdf = compute_pbr(mat, N, V, L, H, radiance);
        const bool sampled = light.shadowType == 1 || light.shadowType == 2;
        // exactly zero (not NaN) with a finite shadow factor: the term cannot change Lo (compute_lighting)
        if (!(brdf.x == 0.0f && brdf.y == 0.0f && brdf.z == 0.0f && (!sampled || light.pcfSamples > 0))) {
          lmask |= 1u << l;
          job = sampled;
          bout[(size_t)l * cap + slot] = make_float4(brdf.x, brdf.y, brdf.z, light.shadowType == 0 ? 1.0f : 0.0f);   // calculateShadow (glsl:442-455)
        }
      }
      const unsigned jm = __ballot_sync(0xffffffffu, job);
      if (jm) {
        unsigned jb = 0;
        if (lane == 0) jb = atomicAdd(wp.counters + wf_c_jobn(depth, l), (unsigned)__popc(jm));
        jb = __shfl_sync(0xffffffffu, jb, 0);
        if (job) wp.jobq[(size_t)l * cap + jb + (unsigned)__popc(jm & lt)] = slot;
      }
    }
    if (hit) {
      hout[slot] = make_float4(P.x, P.y, P.z, __int_as_float(index));
      hout[cap + slot] = make_float4(N.x, N.y, N.z, __uint_as_float(lmask));
    }
  }
}

template <bool BIG, bool FAN>
__global__ void __launch_bounds__(kThreads, (BIG ? RT_WF_SHADOW_MINB_BIG : RT_WF_SHADOW_MINB) * 8 / RT_PK_WARPS)
wf_shadow_kernel(const __grid_constant__ KernelParams kp, const __grid_constant__ WfParams wp) {
  extern __shared__ float4 smem[];
  Scene s; saddr cand_base; Textures tx;
  wf_stage_scene<BIG>(kp, smem, s, cand_base, tx);
  const int lane = threadIdx.x & 31;
  const int depth = wp.depth;
  const unsigned cap = wp.cap;
  const float4* __restrict__ q = wp.path[depth & 1];
  const float4* __restrict__ hq = wp.hit[depth & 1];
  float* sf_out = reinterpret_cast<float*>(wp.brdf[depth & 1]);
#pragma unroll 1
  for (int l = 0; l < s.nl; ++l) {                    // light by light: the warps of an SM run the same light's code
    const unsigned n = wp.counters[wf_c_jobn(depth, l)];
    if (n == 0u) continue;
    const LightRec light = load_light(s, l);
    unsigned* cur = wp.counters + wf_c_jobcur(depth, l);
    for (;;) {
      unsigned base = 0;
      if (lane == 0) base = atomicAdd(cur, 32u);
      base = __shfl_sync(0xffffffffu, base, 0);
      if (base >= n) break;
      const unsigned j = base + (unsigned)lane;
      // warp-uniform: all 32 lanes hold a job of light l. The shared prescan is OFF here by default: the job queue is filled
      // by whole warps in arrival order, a warp's 32 jobs come from unrelated pixel blocks more often than in the regrouping
      // megakernel (14 % of the scans found no common cone, 7 % overflowed) and C3 measured 13.3 -> 14.3 ms with it on.
      const bool full_warp = RT_WF_SHADOW_COOP && base + 32u <= n;
      if (j >= n) continue;
      const unsigned slot = wp.jobq[(size_t)l * cap + j];
      const float4 h0 = hq[slot], h1 = hq[cap + slot];
      const float noise_r = q[cap + slot].w;
      const f3 P = xyz(h0), N = xyz(h1);
      f3 lightDir; float lightDistance, attenuation;
      wf_light_vector(light, P, lightDir, lightDistance, attenuation);
      const f3 shadow_origin = smad3(N, 0.001f, P);          // glsl:381, :412 (P + N * 0.001)
      const float sf = shadow_packets<BIG, FAN>(s, cand_base, noise_r, shadow_origin, light, l, lightDir, lightDistance, full_warp);
      sf_out[((size_t)l * cap + slot) * 4u + 3u] = sf;
    }
  }
}

template <typename K>
inline cudaError_t wf_launch(K kernel, const KernelParams& kp, const WfParams& wp, int sm_count, cudaStream_t stream) {
  const size_t smem = (size_t)kp.sc.pk_float4s * sizeof(float4) + (size_t)kScratchBytes;
  int per_sm = 0;
  cudaError_t e = cached_occupancy(kernel, kThreads, smem, &per_sm);
  if (e != cudaSuccess) return e;
  kernel<<<(unsigned)(sm_count * per_sm), kThreads, smem, stream>>>(kp, wp);
  return cudaGetLastError();
}

inline cudaError_t launch_wf_bounce(const KernelParams& kp, const WfParams& wp, int sm_count, cudaStream_t stream) {
  if (kp.sc.pk_big)
    return kp.fp.row_band > 0 ? wf_launch(wf_bounce_kernel<true, true>, kp, wp, sm_count, stream)
                              : wf_launch(wf_bounce_kernel<true, false>, kp, wp, sm_count, stream);
  return kp.fp.row_band > 0 ? wf_launch(wf_bounce_kernel<false, true>, kp, wp, sm_count, stream)
                            : wf_launch(wf_bounce_kernel<false, false>, kp, wp, sm_count, stream);
}
inline cudaError_t launch_wf_shadow(const KernelParams& kp, const WfParams& wp, int sm_count, cudaStream_t stream) {
#if !RT_PARITY
  if (kp.sc.pk_fan_ok)
    return kp.sc.pk_big ? wf_launch(wf_shadow_kernel<true, true>, kp, wp, sm_count, stream)
                        : wf_launch(wf_shadow_kernel<false, true>, kp, wp, sm_count, stream);
#endif
  return kp.sc.pk_big ? wf_launch(wf_shadow_kernel<true, false>, kp, wp, sm_count, stream)
                      : wf_launch(wf_shadow_kernel<false, false>, kp, wp, sm_count, stream);
}

}  // namespace pk
}  // namespace RT_VARIANT_NS
}  // namespace rt
