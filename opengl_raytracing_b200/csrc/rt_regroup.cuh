// rt_regroup.cuh — the persistent megakernel WITH ray regrouping: the phases of the wavefront pipeline
// (rt_wavefront.cuh) inside one CTA, their queues in shared memory. north_star's "warp-level ballot /
// compaction to regroup rays that diverge at the reflect / refract / terminate branch inside a persistent
// megakernel" (/root/reference/shader/raytracingCs.glsl:526-534 miss, :544-549 Russian roulette, :555-576 the
// next-ray branch; :462-504 the light loop).
//
// A CTA takes a 32x8-pixel block (as the coherent mode of rt_packet_kernel does) and walks it in rounds:
//   A  lane i < n_live: [shade the hit of the previous depth from its (BRDF term, shadow factor) records,
//      Russian roulette, next ray] -> closest hit. Ended paths go to the block's colour sums; the surviving
//      hits are compacted with __ballot_sync + popc into the first n_hit slots of the shared-memory path / hit
//      records, the (hit, light) pairs that need shadow rays into per-light job lists.
//   B  lane j < n_jobs: one shadow job each (light-major order: a warp runs one light's code), factor -> record.
// Whole warps beyond n_live / n_jobs skip the round. The path state is PARKED in shared memory between
// rounds, so neither round carries the other's registers: the kernel is sized for 3 CTAs (24 warps) per SM at
// 80 registers with a few hundred bytes of spills, where rt_packet_kernel needs 128 registers (16 warps).
// Same per-pixel arithmetic as rt_packet_kernel / the wavefront kernels: bit-identical images.
#pragma once
#include "rt_wavefront.cuh"

namespace rt {
namespace RT_VARIANT_NS {
namespace pk {

// Resident CTAs per SM the registers are sized for, measured on B200 (ms, 3 CTAs / 80 registers vs 2 CTAs / 128 registers):
// small scenes C1 0.0911 / 0.0901, C2 0.3644 / 0.3686 (a wash: 3 kept, 24 warps); large scenes C3 13.48 / 12.35,
// C5 13.01 / 12.59 (the long object loops spill at 80 registers: 2).
#ifndef RT_RG_MINB
#define RT_RG_MINB 3
#endif
#ifndef RT_RG_MINB_BIG
#define RT_RG_MINB_BIG 2
#endif
// RT_RG_SPLIT=1: round B in two phases (B1 prescan, B2 the packets of the surviving jobs on compacted lanes). Measured on
// B200 and REJECTED: C1 0.088 -> 0.094, C2 0.360 -> 0.395, C3 12.37 -> 13.45 ms — three more CTA barriers per batch and
// the recomputed job data cost more than the idle lanes of the one-pass form. Bit-identical either way.
#ifndef RT_RG_SPLIT
#define RT_RG_SPLIT 0
#endif
static_assert(!RT_RG_SPLIT || kThreads <= 256, "RT_RG_SPLIT packs the owner lane into 8 bits: build it with RT_RG_WARPS_SMALL=8 RT_RG_WARPS_BIG=8");

// bytes of dynamic shared memory behind the scene blob and the candidate scratch, for nl lights
__host__ __device__ inline size_t regroup_queue_bytes(int nl) {
  return (size_t)kThreads * (3 * 16 + 2 * 16 + 16 + 8) + (size_t)(nl > 0 ? nl : 1) * kThreads * (16 + 2);
}

template <bool BIG, bool BANDS, bool FAN>
__global__ void __launch_bounds__(kThreads, ((BIG ? RT_RG_MINB_BIG : RT_RG_MINB) * 8 / RT_PK_WARPS > 0 ? (BIG ? RT_RG_MINB_BIG : RT_RG_MINB) * 8 / RT_PK_WARPS : 1)) rt_regroup_kernel(const __grid_constant__ KernelParams kp) {
  extern __shared__ float4 smem[];
  const rt_frame_params& fp = kp.fp;
  Scene s; saddr cand_base; Textures tx;
  wf_stage_scene<BIG>(kp, smem, s, cand_base, tx);
  const int nl = s.nl;
  // queue records behind the scene blob and the per-thread candidate lists (16-byte aligned: both are multiples of 16)
  float4* q = smem + kp.sc.pk_float4s + kScratchBytes / 16;
  float4* q_rd = q;                      // (rd, pixel bits)
  float4* q_tp = q + kThreads;           // (throughput, noise_r)
  float4* q_col = q + 2 * kThreads;      // (colour so far, block-local pixel bits)
  float4* h_p = q + 3 * kThreads;        // (P, object index bits)
  float4* h_n = q + 4 * kThreads;        // (N, contributing-light mask bits)
  float4* sum = q + 5 * kThreads;        // per pixel of the block: running colour sum over the samples
  uint2* surv_rec = reinterpret_cast<uint2*>(q + 6 * kThreads);   // round B: the surviving jobs of a batch (kThreads x 8 B = half a plane)
  float4* brdf = q + 6 * kThreads + kThreads / 2;                 // nl planes: (BRDF term, shadow factor)
  unsigned short* jobq = reinterpret_cast<unsigned short*>(brdf + (size_t)(nl > 0 ? nl : 1) * kThreads);   // nl planes of hit slots
  __shared__ unsigned s_tile, s_nhit, s_njob[kRgMaxLights], s_wsurv[kWarpsPerCta];
  __shared__ unsigned s_jobcur;                       // round B: the next unclaimed batch of 32 jobs
  __shared__ unsigned char s_lorder[kRgMaxLights];    // round B visits the lights in this order: the expensive ones (PCSS) first
  if (threadIdx.x == 0) {
    int k = 0;
    for (int pass = 0; pass < 2; ++pass)
      for (int l = 0; l < nl; ++l)
        if ((load_light(s, l).shadowType == 2) == (pass == 0)) s_lorder[k++] = (unsigned char)l;
  }

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned lt = (1u << lane) - 1u;
  const int nsamp = fp.num_samples > 0 ? fp.num_samples : 1;
  const float inv_ns = rcp_ieee((float)nsamp);
  const int D = fp.max_ray_depth;

  for (;;) {
    __syncthreads();
    if (tid == 0) s_tile = atomicAdd(kp.tile_counter, 1u) - kp.tile_base;
    __syncthreads();
    const unsigned blk = s_tile;                         // block id: 4 x 2 tiles = 32 x 8 pixels
    if (blk >= kp.blocks_x * kp.blocks_y) break;
    const unsigned by = blk / kp.blocks_x, bx = blk - by * kp.blocks_x;
    const unsigned tyi = by * (unsigned)kBlockTilesY + ((unsigned)warp >> 2), txi = bx * (unsigned)kBlockTilesX + ((unsigned)warp & 3u);
    const bool tile_ok = tyi < kp.tiles_y && txi < kp.tiles_x;
    const int px0 = (int)txi * kTileW + (lane & (kTileW - 1));
    const int py0 = (BANDS ? window_row(fp, (int)tyi * kTileH) : fp.row_begin + (int)tyi * kTileH) + (lane >> 3);
    const bool inside = tile_ok && px0 < fp.width && py0 < fp.row_end;

#pragma unroll 1
    for (int sidx = 0; sidx < nsamp; ++sidx) {
      const int frame_count = fp.frame_count + sidx;
      unsigned n_live = kThreads;                        // depth 0: every lane of the block owns its pixel
#pragma unroll 1
      for (int depth = 0; depth <= D; ++depth) {
        const bool first = depth == 0, shade_only = depth >= D;
        // ================= round A: read the parked state (all lanes), then barrier, then work and re-park
        bool alive = false;
        unsigned pxy = 0u, local = (unsigned)tid;
        float noise_r = 0.0f;
        f3 ro = mk3(0.0f, 0.0f, 0.0f), rd = ro, throughput = mk3(1.0f, 1.0f, 1.0f), color = ro;
        f3 pP = ro, pN = ro;                             // the previous hit (for the pixel of a path that now misses)
        int pindex = -1;
        bool finish = false;                             // this lane's path ends in this round
        if (first) {
          if (inside) {
            pxy = (unsigned)px0 | ((unsigned)py0 << 16);
            noise_r = fetch_noise(tx, (float)((unsigned)px0 + (unsigned)frame_count) * fp.noise_scale[0],
                                  (float)((unsigned)py0 + (unsigned)frame_count) * fp.noise_scale[1]);
            const float jx = noise_r * 2.0f - 1.0f, jy = 0.0f * 2.0f - 1.0f;
            float ux = div_ieee((float)px0 + 0.5f + jx, (float)fp.width);
            float uy = div_ieee((float)py0 + 0.5f + jy, (float)fp.height);
            ux = ux * 2.0f - 1.0f;
            uy = uy * 2.0f - 1.0f;
            ux *= kp.aspect * kp.tan_fov * fp.focal_length;
            uy *= kp.tan_fov * fp.focal_length;
            const f3 camDir = mk3(fp.camera_dir[0], fp.camera_dir[1], fp.camera_dir[2]);
            const f3 camRight = mk3(fp.camera_right[0], fp.camera_right[1], fp.camera_right[2]);
            const f3 camUp = mk3(fp.camera_up[0], fp.camera_up[1], fp.camera_up[2]);
            ro = mk3(fp.camera_pos[0], fp.camera_pos[1], fp.camera_pos[2]);
            rd = normalize_ieee((camDir + ux * camRight) + uy * camUp);
            alive = !shade_only;
            finish = shade_only;                         // max_ray_depth == 0: the loop body never runs (glsl:519)
          }
        } else if ((unsigned)tid < n_live) {
          const int pd = depth - 1;
          const float4 q0 = q_rd[tid], q1 = q_tp[tid], q2 = q_col[tid];
          const float4 h0 = h_p[tid], h1 = h_n[tid];
          rd = xyz(q0); pxy = __float_as_uint(q0.w);
          throughput = xyz(q1); noise_r = q1.w;
          color = xyz(q2); local = __float_as_uint(q2.w);
          const f3 P = xyz(h0), N = xyz(h1);
          const int index = __float_as_int(h0.w);
          unsigned lmask = __float_as_uint(h1.w);
          pP = P; pN = N; pindex = index;
          const Mat mat = load_material(s, index);
          f3 Lo = mk3(0.0f, 0.0f, 0.0f);
          while (lmask) {                                // Lo += BRDF * shadow, in light order (glsl:502)
            const int l = __ffs(lmask) - 1;
            lmask &= lmask - 1u;
            const float4 b = brdf[l * kThreads + tid];
            Lo = smad3(xyz(b), b.w, Lo);
          }
          if (mat.subsurfaceScatter > 0.0f) {
            const Scene s2 = s; const Mat mat2 = mat;    // copies: see trace_closest
            Lo = Lo + subsurface(s2, P, N, mat2);
          }
          color = smad3(throughput, Lo, color);
          bool cont = true;
          const unsigned px = pxy & 0xffffu, py = pxy >> 16;
          if (pd > 2) {                                  // Russian roulette (glsl:544-549)
            const float diffuseWeight = sqrt_ieee(dot3(mat.albedo, mat.albedo)) * mat.diffuseStrength;
            const float continueProb =
                fminf(fmaxf(throughput.x, fmaxf(throughput.y, throughput.z)) * 0.95f + diffuseWeight, 0.99f);
            if (random2((float)(px + (unsigned)pd), (float)(py + (unsigned)pd)) > continueProb) cont = false;
            else throughput = mk3(div_ieee(throughput.x, continueProb), div_ieee(throughput.y, continueProb),
                                  div_ieee(throughput.z, continueProb));
          }
          if (pd + 1 >= D) cont = false;                 // the next ray would never be traced
          if (cont) {
            const f3 V = normalize_color(-rd);
            const float F = fresnel_schlick(fmaxf(dot3(V, N), 0.0f), mat.ior);
            if (mat.diffuseStrength > 0.0f) {            // glsl:555-567
              const int hi = pd * 64 + frame_count;
              const f3 specularDir = reflect3(rd, N);
              const f3 diffuseDir = cosine_hemisphere(div_ieee((float)hi, 64.0f), halton(hi, 2), N);
              rd = normalize_ieee(mix3(specularDir, diffuseDir, mat.roughness));
              ro = P + N * 0.001f;
              throughput = throughput * (mat.albedo * mat.diffuseStrength);
            } else if (mat.transparency > 0.0f) {        // glsl:568-571
              rd = refraction_dir(rd, N, mat.ior);
              ro = P - N * 0.001f;
              throughput = throughput * ((mat.albedo * (1.0f - F)) * mat.transparency);
            } else {                                     // glsl:572-576
              rd = reflect3(rd, N);
              ro = P + N * 0.001f;
              throughput = throughput * (mat.albedo * F);
            }
            alive = true;
          } else {
            finish = true;
          }
        }
        if (tid == 0) { s_nhit = 0u; s_jobcur = 0u; }
        if (tid < kRgMaxLights) s_njob[tid] = 0u;
        __syncthreads();                                 // every parked record of the previous depth has been read

        // ---- closest hit of the new ray (glsl:522-534)
        bool hit = false;
        Hit h; h.code = -1; h.t = 0.0f;
        if (alive) {
          h = trace_closest<BIG>(s, cand_base, ro, rd);
          if (h.code < 0) {
            const f3 sky = fp.use_skybox ? fetch_sky(tx, rd) : mk3(0.0f, 0.0f, 0.0f);
            color = smad3(throughput, sky, color);
            finish = true;
          } else {
            hit = true;
          }
        }
        if (finish) {
          // the pixel of a finished path: position / normal / id of its LAST hit (sample 0), colour into the block's sums
          if (sidx == 0) {
            const int px = (int)(pxy & 0xffffu), py = (int)(pxy >> 16);
            const size_t at = (size_t)py * fp.width + px;
            if (kp.position) kp.position[at] = make_float4(pP.x, pP.y, pP.z, 1.0f);
            if (kp.normal) {
              if (fp.flags & RT_FLAG_NORMAL_FP32) reinterpret_cast<float4*>(kp.normal)[at] = make_float4(pN.x, pN.y, pN.z, 1.0f);
              else reinterpret_cast<uint2*>(kp.normal)[at] = pack_half4(pN.x, pN.y, pN.z, 1.0f);
            }
            if (kp.object_id) kp.object_id[at] = pindex;
          }
          if (sidx > 0) { const float4 a = sum[local]; color = mk3(a.x, a.y, a.z) + color; }     // r.color + r2.color
          sum[local] = make_float4(color.x, color.y, color.z, __uint_as_float(pxy));
        }
        if (shade_only) break;                           // uniform: depth == D

        // ---- regroup: the surviving paths take the first n_hit slots
        const unsigned hm = __ballot_sync(0xffffffffu, hit);
        unsigned slot = 0;
        if (hm) {
          if (lane == 0) slot = atomicAdd(&s_nhit, (unsigned)__popc(hm));
          slot = __shfl_sync(0xffffffffu, slot, 0) + (unsigned)__popc(hm & lt);
        }
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
          q_rd[slot] = make_float4(rd.x, rd.y, rd.z, __uint_as_float(pxy));
          q_tp[slot] = make_float4(throughput.x, throughput.y, throughput.z, noise_r);
          q_col[slot] = make_float4(color.x, color.y, color.z, __uint_as_float(local));
        }
        // ---- computeLighting's per-light BRDF terms (glsl:462-504); lights that need rays become shadow jobs
        unsigned lmask = 0u;
        if (hm) {                                        // warp-uniform
#pragma unroll 1
          for (int l = 0; l < nl; ++l) {
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
              const f3 b = compute_pbr(mat, N, V, L, H, radiance);
              const bool sampled = light.shadowType == 1 || light.shadowType == 2;
              if (!(b.x == 0.0f && b.y == 0.0f && b.z == 0.0f && (!sampled || light.pcfSamples > 0))) {
                lmask |= 1u << l;
                job = sampled;
                brdf[l * kThreads + (int)slot] = make_float4(b.x, b.y, b.z, light.shadowType == 0 ? 1.0f : 0.0f);   // calculateShadow (glsl:442-455)
              }
            }
            const unsigned jm = __ballot_sync(0xffffffffu, job);
            if (jm) {
              unsigned jb = 0;
              if (lane == 0) jb = atomicAdd(&s_njob[l], (unsigned)__popc(jm));
              jb = __shfl_sync(0xffffffffu, jb, 0);
              if (job) jobq[l * kThreads + (int)(jb + (unsigned)__popc(jm & lt))] = (unsigned short)slot;
            }
          }
        }
        if (hit) {
          h_p[slot] = make_float4(P.x, P.y, P.z, __int_as_float(index));
          h_n[slot] = make_float4(N.x, N.y, N.z, __uint_as_float(lmask));
        }
        __syncthreads();
        n_live = s_nhit;
        if (n_live == 0u) break;                         // uniform: every path of the block has ended

        // ================= round B: the shadow jobs, light-major, in batches of one job per lane
        //   B1  every job runs its PRESCAN (light grid / cluster scan with the cone test -> the lane's candidate list). Jobs
        //       with nothing in their cone are done: factor 1 (C2: half of all jobs). The survivors are compacted CTA-wide
        //       in job order (still light-major: a warp of B2 runs one light's code).
        //   B2  lane k < n_surv runs the PACKETS of survivor k, reading the candidate list where its owner lane left it.
        // so the lanes that trace rays are consecutive and whole warps without survivors skip the packets (RT_RG_SPLIT=0:
        // one pass, prescan and packets in the same lane — the A/B baseline).
        {
          unsigned total = 0;
          for (int l = 0; l < nl; ++l) total += s_njob[l];
          // the job of index j: (light, hit slot) and everything shadow_packets needs — recomputed in B2 from the same records
          auto job_of = [&](unsigned j, int& l, int& slot2, LightRec& light, float& nr, f3& origin, f3& lightDir, float& lightDistance) {
            int li = 0; unsigned jj = j;
            l = s_lorder[0];
            while (jj >= s_njob[l]) { jj -= s_njob[l]; l = s_lorder[++li]; }
            slot2 = (int)jobq[l * kThreads + (int)jj];
            light = load_light(s, l);
            const float4 h0 = h_p[slot2], h1 = h_n[slot2];
            nr = q_tp[slot2].w;
            float attenuation;
            wf_light_vector(light, xyz(h0), lightDir, lightDistance, attenuation);
            origin = smad3(xyz(h1), 0.001f, xyz(h0));                            // glsl:381, :412 (P + N * 0.001)
          };
#if RT_RG_SPLIT
#pragma unroll 1
          for (unsigned j0 = 0; j0 < total; j0 += kThreads) {
            const unsigned j = j0 + (unsigned)tid;
            bool surv = false;
            Cands cd; cd.base = cand_base; cd.n = 0; cd.first = 0; cd.nsph = 0; cd.list = true; cd.prefilter = false;
            if (j < total) {                             // ---- B1
              int l, slot2; LightRec light; float nr, lightDistance; f3 origin, lightDir;
              job_of(j, l, slot2, light, nr, origin, lightDir, lightDistance);
              bool done = false;
              const float sf = shadow_packets_m<BIG, FAN, 1>(s, cand_base, nr, origin, light, l, lightDir, lightDistance, cd, done);
              if (done) reinterpret_cast<float*>(brdf)[(l * kThreads + slot2) * 4 + 3] = sf;
              else surv = true;
            }
            const unsigned sm = __ballot_sync(0xffffffffu, surv);
            if (lane == 0) s_wsurv[warp] = (unsigned)__popc(sm);
            __syncthreads();
            unsigned before = 0, n_surv = 0;
#pragma unroll
            for (int w2 = 0; w2 < kWarpsPerCta; ++w2) { const unsigned c = s_wsurv[w2]; n_surv += c; if (w2 < warp) before += c; }
            if (surv) {                                  // record: job | owner lane | list kind, candidate counts
              const unsigned k = before + (unsigned)__popc(sm & lt);
              surv_rec[k] = make_uint2(j | ((unsigned)tid << 16) | (cd.list ? 1u << 24 : 0u) | (cd.prefilter ? 1u << 25 : 0u),
                                       (unsigned)cd.n | ((unsigned)cd.nsph << 16));
            }
            __syncthreads();
            if ((unsigned)tid < n_surv) {                // ---- B2
              const uint2 rec = surv_rec[tid];
              int l, slot2; LightRec light; float nr, lightDistance; f3 origin, lightDir;
              job_of(rec.x & 0xffffu, l, slot2, light, nr, origin, lightDir, lightDistance);
              Cands cd2;
              cd2.base = cand_base - (unsigned)tid * 2u + ((rec.x >> 16) & 0xffu) * 2u;     // the owner lane's scratch list
              cd2.n = (int)(rec.y & 0xffffu); cd2.nsph = (int)(rec.y >> 16); cd2.first = 0;
              cd2.list = (rec.x >> 24) & 1u; cd2.prefilter = (rec.x >> 25) & 1u;
              bool done = false;
              const float sf = shadow_packets_m<BIG, FAN, 2>(s, cand_base, nr, origin, light, l, lightDir, lightDistance, cd2, done);
              reinterpret_cast<float*>(brdf)[(l * kThreads + slot2) * 4 + 3] = sf;
            }
            __syncthreads();                             // the candidate lists are free for the next batch
          }
#else
          // Batches of 32 consecutive jobs, claimed by whichever warp is free (one shared-memory atomic per batch): a warp
          // that drew a batch of wide-cone jobs takes fewer batches, and the expensive lights go first, so the barrier
          // behind the round waits for a short tail (ncu, C3 before: 17 % of all warp samples sat at that barrier).
#pragma unroll 1
          for (;;) {
            unsigned j0 = 0;
            if (lane == 0) j0 = atomicAdd(&s_jobcur, 32u);
            j0 = __shfl_sync(0xffffffffu, j0, 0);
            if (j0 >= total) break;
            const unsigned j = j0 + (unsigned)lane;
            const bool full_warp = j0 + 32u <= total;                           // warp-uniform: all 32 lanes hold a job
            if (j < total) {
              int l, slot2; LightRec light; float nr, lightDistance; f3 origin, lightDir;
              job_of(j, l, slot2, light, nr, origin, lightDir, lightDistance);
              const float sf = shadow_packets<BIG, FAN>(s, cand_base, nr, origin, light, l, lightDir, lightDistance, full_warp);
              reinterpret_cast<float*>(brdf)[(l * kThreads + slot2) * 4 + 3] = sf;
            }
          }
#endif
        }
        __syncthreads();
      }
      __syncthreads();                                   // the sums of this sample are complete
    }
    // ---- the block's colour image: one pixel per lane, 8 px = one 128-B line per row of a tile
    if (inside && kp.color) {
      const float4 a = sum[tid];
      float4 c = make_float4(a.x, a.y, a.z, 1.0f);
      const size_t at = (size_t)py0 * fp.width + px0;
      if (fp.flags & RT_FLAG_ACCUMULATE) {               // partial sums for sample-range sharding
        const float4 old = kp.color[at];
        c = make_float4(old.x + c.x, old.y + c.y, old.z + c.z, 1.0f);
      } else if (nsamp > 1 && !(fp.flags & RT_FLAG_SUM_SAMPLES)) {
        c = make_float4(c.x * inv_ns, c.y * inv_ns, c.z * inv_ns, 1.0f);
      }
      kp.color[at] = c;
    }
  }
}

template <bool BIG, bool BANDS, bool FAN>
inline cudaError_t launch_regroup_t(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* advance) {
  const size_t smem = (size_t)kp.sc.pk_float4s * sizeof(float4) + (size_t)kScratchBytes + regroup_queue_bytes(kp.sc.num_lights);
  int per_sm = 0;
  cudaError_t e = cached_occupancy(rt_regroup_kernel<BIG, BANDS, FAN>, kThreads, smem, &per_sm);
  if (e != cudaSuccess) return e;
  unsigned ctas = (unsigned)(sm_count * per_sm);
  const unsigned blocks = kp.blocks_x * kp.blocks_y;
  if (ctas > blocks) ctas = blocks;
  if (ctas < 1) ctas = 1;
  *advance = blocks + ctas - kp.tiles_x * kp.tiles_y;    // one fetch per block + one failing fetch per CTA (the form rt_render adds back)
  rt_regroup_kernel<BIG, BANDS, FAN><<<ctas, kThreads, smem, stream>>>(kp);
  return cudaGetLastError();
}

template <bool BANDS, bool FAN>
inline cudaError_t launch_regroup_f(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* advance) {
#if defined(RT_RG_ONLY_BIG) && RT_RG_ONLY_BIG == 1        // this translation unit holds the large-scene kernel only
  return launch_regroup_t<true, BANDS, FAN>(kp, sm_count, stream, advance);
#elif defined(RT_RG_ONLY_BIG)
  return launch_regroup_t<false, BANDS, FAN>(kp, sm_count, stream, advance);
#else
  return kp.sc.pk_big ? launch_regroup_t<true, BANDS, FAN>(kp, sm_count, stream, advance)
                      : launch_regroup_t<false, BANDS, FAN>(kp, sm_count, stream, advance);
#endif
}
inline cudaError_t launch_regroup(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* advance) {
#if !RT_PARITY
  if (kp.sc.pk_fan_ok)
    return kp.fp.row_band > 0 ? launch_regroup_f<true, true>(kp, sm_count, stream, advance) : launch_regroup_f<false, true>(kp, sm_count, stream, advance);
#endif
  return kp.fp.row_band > 0 ? launch_regroup_f<true, false>(kp, sm_count, stream, advance) : launch_regroup_f<false, false>(kp, sm_count, stream, advance);
}

}  // namespace pk
}  // namespace RT_VARIANT_NS
}  // namespace rt
