// rt_kernel.cuh — the sm_100a ray-tracing kernel: a from-scratch CUDA implementation of
// the reference's per-pixel compute pass (/root/reference/shader/raytracingCs.glsl).
//
// Included by exactly two translation units:
//   rt_kernel_fast.cu    RT_PARITY 0: approximate SFU forms (rsqrt, fast divide, pow5 by
//                        multiplication) and explicit fmaf only in colour-only shading code;
//                        the geometric chain ray -> t -> P -> N -> next ray is IEEE and unfused.
//   rt_kernel_parity.cu  RT_PARITY 1: IEEE div/sqrt everywhere, same operation order as the
//                        shader -> agrees with the CPU oracle to the last bit except for
//                        sin/cos/pow/exp library ulps.
// Both are compiled with -fmad=false (no implicit FMA contraction): measured on B200, letting
// nvcc contract the sphere/reflect arithmetic moved the last-hit position by up to 5e-2
// relative after three mirror bounces (profiles/README.md, round 1).
//
// Design: a persistent kernel, one CTA per SM slot. Each CTA stages the
// scene once into shared memory as an SoA "intersection set" (48 B/object) that every
// lane reads at the same index -> pure LDS broadcast. Warps are independent workers that
// pull 8x4-pixel tiles from a global atomic counter (cheap sky tiles and expensive
// reflective tiles balance dynamically). Shadow rays use an exact any-hit early-out.
// Stores are one float4 per pixel per image: 8 consecutive pixels = one 128-B line.
#pragma once
#include "rt_math.cuh"

namespace rt {
namespace RT_VARIANT_NS {

constexpr int kWarpsPerCta = 8;
constexpr int kThreads = kWarpsPerCta * 32;
constexpr int kTileW = 8, kTileH = 4;            // one warp = 8x4 pixels

// ---------------------------------------------------------------- counters (flop model v0)
template <bool COUNT> struct Cnt;
template <> struct Cnt<false> {
  __device__ __forceinline__ void add(int, unsigned = 1u) {}
  __device__ __forceinline__ void zero() {}
  __device__ __forceinline__ unsigned get(int) const { return 0u; }
};
template <> struct Cnt<true> {
  unsigned c[RT_CNT_COUNT];
  __device__ __forceinline__ void add(int i, unsigned n = 1u) { c[i] += n; }
  __device__ __forceinline__ void zero() { for (int i = 0; i < RT_CNT_COUNT; ++i) c[i] = 0u; }
  __device__ __forceinline__ unsigned get(int i) const { return c[i]; }
};

// ---------------------------------------------------------------- shared-memory scene view
struct SceneView {
  const float4* isect;      // 3 float4 per object
  const float4* plane;      // 3 float4 per plane
  const float4* lights;     // 6 float4 per light (raw rt_light)
  const float4* material;   // 4 float4 per object, global memory (gathered on closest hit only)
  int num_objects, num_lights;
  float max_dist;
};

struct Hit { int index; float t; };

// ---------------------------------------------------------------- intersection (glsl:91-196)

// intersectAABB (glsl:91-103) with 1/dir hoisted per ray. fminf/fmaxf are IEEE minNum/maxNum
// (single FMNMX each), the semantics pinned for GLSL min/max. The subtraction and the
// multiplication are kept separate (no FMA can form): +-inf slabs must stay +-inf.
__device__ __forceinline__ bool slab_test(float4 lo, float4 hi, f3 o, f3 inv, float max_dist) {
  float t0x = (lo.x - o.x) * inv.x, t1x = (hi.x - o.x) * inv.x;
  float t0y = (lo.y - o.y) * inv.y, t1y = (hi.y - o.y) * inv.y;
  float t0z = (lo.z - o.z) * inv.z, t1z = (hi.z - o.z) * inv.z;
  float tMin = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
  float tMax = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
  return tMax >= tMin && tMin < max_dist && tMax > 0.0f;
}

// intersectSphere (glsl:105-118): near root only. a = dot(d,d) is hoisted per ray.
__device__ __forceinline__ bool sphere_test(float4 sh, f3 o, f3 d, float a, float& t) {
  f3 oc = o - xyz(sh);
  float b = 2.0f * dot3(oc, d);
  float c = dot3(oc, oc) - sh.w;                 // sh.w = radius * radius
  float disc = b * b - 4.0f * a * c;
  if (disc < 0.0f) return false;
  t = div_ieee(-b - sqrt_ieee(disc), 2.0f * a);
  return t > 0.0f;
}

// intersectPlane (glsl:120-153) with the per-object frame hoisted to staging time.
// *full tells the counters whether the frame section (glsl:126-150) was reached.
__device__ __forceinline__ bool plane_test(const float4* pl, float4 sh, f3 o, f3 d, float& t, bool& full) {
  full = false;
  float4 pn = pl[0];
  f3 n = xyz(pn);
  float denom = dot3(n, d);
  if (!(fabsf(denom) > 1e-6f)) return false;
  t = div_ieee(dot3(xyz(sh) - o, n), denom);
  if (t < 0.0f) return false;
  full = true;
  float4 pr = pl[1], pf = pl[2];
  f3 local = (o + d * t) - xyz(sh);
  float x = dot3(local, xyz(pr));
  float z = dot3(local, xyz(pf));
  if (fabsf(x) > pn.w || fabsf(z) > pr.w) return false;   // size.x/2, size.y/2
  return true;
}

// intersectObjects (glsl:155-196): closest hit over ALL objects in array order; strict `<`
// keeps the first object on ties. The reference copies the material and computes the normal
// at every improvement; doing both once for the final index yields identical values.
template <bool COUNT>
__device__ __forceinline__ Hit trace_closest(const SceneView& s, f3 o, f3 d, Cnt<COUNT>& k) {
  const f3 inv = mk3(rcp_ieee(d.x), rcp_ieee(d.y), rcp_ieee(d.z));
  const float a = dot3(d, d);
  Hit h; h.index = -1; h.t = s.max_dist;
  k.add(RT_CNT_RAYS);
  k.add(RT_CNT_AABB, (unsigned)s.num_objects);
#pragma unroll 2
  for (int i = 0; i < s.num_objects; ++i) {
    const float4 lo = s.isect[3 * i + 0], hi = s.isect[3 * i + 1];
    if (!slab_test(lo, hi, o, inv, s.max_dist)) continue;
    const float4 sh = s.isect[3 * i + 2];
    const int type = __float_as_int(lo.w);
    float t = 0.0f; bool is_hit = false;
    if (type == 0) {
      k.add(RT_CNT_SPHERE);
      is_hit = sphere_test(sh, o, d, a, t);
    } else if (type == 1) {
      bool full;
      is_hit = plane_test(s.plane + 3 * __float_as_int(hi.w), sh, o, d, t, full);
      k.add(full ? RT_CNT_PLANE_FULL : RT_CNT_PLANE_EARLY);
    }
    if (is_hit && t > 0.0f && t < h.t) {
      h.t = t; h.index = i;
      k.add(type == 0 ? RT_CNT_HIT_UPDATE_SPHERE : RT_CNT_HIT_UPDATE_PLANE);
    }
  }
  return h;
}

// Shadow / probe rays only ask "is there a hit with t < maxRayDistance (and t < limit)?"
// (glsl:388-392, :419-423): `hit && closestT < L` <=> `exists hit with t < L`, so stopping
// at the first such object is exactly equivalent. use_limit is false for directional lights.
// With COUNT the literal full loop runs instead so the op counters match the reference.
template <bool COUNT>
__device__ __forceinline__ bool trace_any(const SceneView& s, f3 o, f3 d, bool use_limit, float limit, Cnt<COUNT>& k) {
  if (COUNT) {
    Hit h = trace_closest<COUNT>(s, o, d, k);
    return h.index >= 0 && (!use_limit || h.t < limit);
  }
  const f3 inv = mk3(rcp_ieee(d.x), rcp_ieee(d.y), rcp_ieee(d.z));
  const float a = dot3(d, d);
#pragma unroll 2
  for (int i = 0; i < s.num_objects; ++i) {
    const float4 lo = s.isect[3 * i + 0], hi = s.isect[3 * i + 1];
    if (!slab_test(lo, hi, o, inv, s.max_dist)) continue;
    const float4 sh = s.isect[3 * i + 2];
    const int type = __float_as_int(lo.w);
    float t = 0.0f; bool is_hit = false;
    if (type == 0) {
      is_hit = sphere_test(sh, o, d, a, t);
    } else if (type == 1) {
      bool full;
      is_hit = plane_test(s.plane + 3 * __float_as_int(hi.w), sh, o, d, t, full);
    }
    if (is_hit && t > 0.0f && t < s.max_dist && (!use_limit || t < limit)) return true;
  }
  return false;
}

// Surface normal of the closest hit (glsl:186-191).
__device__ __forceinline__ f3 hit_normal(const SceneView& s, Hit h, f3 o, f3 d) {
  const float4 lo = s.isect[3 * h.index + 0];
  if (__float_as_int(lo.w) == 0) {
    const float4 sh = s.isect[3 * h.index + 2];
    return normalize_ieee((o + d * h.t) - xyz(sh));
  }
  const float4 hi = s.isect[3 * h.index + 1];
  return xyz(s.plane[3 * __float_as_int(hi.w)]);
}

__device__ __forceinline__ Mat load_material(const SceneView& s, int index) {
  const float4 m0 = __ldg(s.material + 4 * index + 0), m1 = __ldg(s.material + 4 * index + 1);
  const float4 m2 = __ldg(s.material + 4 * index + 2), m3 = __ldg(s.material + 4 * index + 3);
  Mat m;
  m.albedo = xyz(m0); m.metallic = m0.w;
  m.roughness = m1.x; m.diffuseStrength = m1.y; m.ior = m1.z; m.transparency = m1.w;
  m.subsurfaceColor = xyz(m2); m.subsurfaceScatter = m2.w; m.scatterDistance = m3.x;
  return m;
}




// ---------------------------------------------------------------- shadows (glsl:341-455)
struct PixelCtx {
  unsigned px, py;
  int frame_count;
  float noise_r;             // texture(blueNoiseTex, (pix+frameCount)*noiseScale).r — same uv at
                             // both call sites (glsl:360, :513), fetched once per sample
};

struct LightRec {
  int type; f3 position, direction, color; float intensity, shadowSoftness; int shadowType, pcfSamples; float lightSize;
};
__device__ __forceinline__ LightRec load_light(const SceneView& s, int i) {
  const rt_light* l = reinterpret_cast<const rt_light*>(s.lights) + i;
  LightRec r;
  r.type = l->type;
  r.position = mk3(l->position[0], l->position[1], l->position[2]);
  r.direction = mk3(l->direction[0], l->direction[1], l->direction[2]);
  r.color = mk3(l->color[0], l->color[1], l->color[2]);
  r.intensity = l->intensity; r.shadowSoftness = l->shadowSoftness;
  r.shadowType = l->shadowType; r.pcfSamples = l->pcfSamples; r.lightSize = l->lightSize;
  return r;
}

// pcfShadow (glsl:342-397)
template <bool COUNT>
__device__ __forceinline__ float pcf_shadow(const SceneView& s, const PixelCtx& pc, f3 origin, const LightRec& light,
                                            f3 lightDir, float lightDistance, Cnt<COUNT>& k) {
  k.add(RT_CNT_PCF_SETUP);
  const f3 tangent = normalize_ieee(cross3(lightDir, mk3(0.0f, 1.0f, 0.0f)));   // NaN when lightDir || Y
  const f3 bitangent = cross3(lightDir, tangent);
  const float filterSize = light.shadowSoftness * 0.005f;                        // glsl:364 (overwrites :352)
  float shadow = 0.0f;
  for (int i = 0; i < light.pcfSamples; ++i) {
    k.add(RT_CNT_PCF_SAMPLE);
    const float rx = fract_(halton2(i) + pc.noise_r);
    const float ry = fract_(halton3(i) + 0.0f);
    f3 d = (lightDir + (rx * tangent) * filterSize) + (ry * bitangent) * filterSize;
    if (light.type != 1) d = normalize_ieee(d);
    const bool occluded = trace_any<COUNT>(s, origin, d, light.type == 0 || light.type == 2, lightDistance, k);
    shadow += occluded ? 0.0f : 1.0f;
  }
  return div_ieee(shadow, (float)light.pcfSamples);      // 0/0 = NaN when pcfSamples == 0
}

// pcssShadow (glsl:400-440): 16 fixed probes answer "any blocker?"; then plain PCF.
template <bool COUNT>
__device__ __forceinline__ float pcss_shadow(const SceneView& s, const PixelCtx& pc, f3 origin, const LightRec& light,
                                             f3 lightDir, float lightDistance, Cnt<COUNT>& k) {
  const float searchSize = light.lightSize * 0.1f;
  bool any_blocker = false;
  for (int i = 0; i < 16; ++i) {
    if (!COUNT && any_blocker) break;                    // remaining probes cannot change the result
    k.add(RT_CNT_PCSS_PROBE);
    const float rv = halton3(i) * 2.0f - 1.0f;
    const f3 d = normalize_ieee((lightDir + rv * searchSize) + rv * searchSize);
    if (trace_any<COUNT>(s, origin, d, light.type != 1, lightDistance, k)) any_blocker = true;
  }
  if (!any_blocker) return 1.0f;
  return pcf_shadow<COUNT>(s, pc, origin, light, lightDir, lightDistance, k);
}

// computeSubsurfaceScattering (glsl:316-339)
template <bool COUNT>
__device__ __forceinline__ f3 subsurface(const SceneView& s, f3 P, f3 N, const Mat& mat, Cnt<COUNT>& k) {
  f3 sss = mk3(0.0f, 0.0f, 0.0f);
  for (int i = 0; i < 4; ++i) {
    k.add(RT_CNT_SSS_SAMPLE);
    const f3 d = cosine_hemisphere((float)i / 4.0f, halton(i, 2), N);
    const f3 o = P + N * 0.001f;
    const Hit h = trace_closest<COUNT>(s, o, d, k);
    if (h.index >= 0) {
      const float attenuation = expf(div_shade(-h.t, mat.scatterDistance));
      sss = sss + xyz(__ldg(s.material + 4 * h.index)) * attenuation;
    }
  }
  const f3 v = (sss * mat.subsurfaceColor) * mat.subsurfaceScatter;
  return mk3(v.x / 4.0f, v.y / 4.0f, v.z / 4.0f);
}

// computeLighting (glsl:457-507). Shadow rays are cast even when the light contributes 0.
template <bool COUNT>
__device__ __forceinline__ f3 compute_lighting(const SceneView& s, const PixelCtx& pc, f3 P, f3 N, const Mat& mat, f3 V,
                                               Cnt<COUNT>& k) {
  f3 Lo = mk3(0.0f, 0.0f, 0.0f);
  const f3 shadow_origin = P + N * 0.001f;               // glsl:381, :412
  for (int i = 0; i < s.num_lights; ++i) {
    const LightRec light = load_light(s, i);
    f3 lightDir = mk3(0.0f, 0.0f, 0.0f);
    float attenuation = 1.0f, lightDistance = 0.0f;
    if (light.type == 0) {
      k.add(RT_CNT_LIGHT_POINT);
      lightDir = light.position - P;
      lightDistance = sqrt_ieee(dot3(lightDir, lightDir));
      attenuation = div_shade(1.0f, 1.0f + 0.1f * lightDistance + 0.01f * lightDistance * lightDistance);
      lightDir = normalize_ieee(lightDir);
    } else if (light.type == 1) {
      k.add(RT_CNT_LIGHT_DIR);
      lightDir = normalize_ieee(-light.direction);
      lightDistance = 1e6f;
    } else if (light.type == 2) {
      k.add(RT_CNT_LIGHT_AREA);
      lightDir = light.position - P;
      lightDistance = sqrt_ieee(dot3(lightDir, lightDir));
      lightDir = normalize_ieee(lightDir);
      attenuation = div_shade(1.0f, lightDistance * lightDistance);
      const f3 lightNormal = normalize_shade(light.direction);
      attenuation *= fmaxf(dot3(lightDir, lightNormal), 0.0f);   // P->light vs facing
    }
    float shadowFactor = 0.0f;                           // calculateShadow (glsl:442-455)
    if (light.shadowType == 0) shadowFactor = 1.0f;
    else if (light.shadowType == 1) shadowFactor = pcf_shadow<COUNT>(s, pc, shadow_origin, light, lightDir, lightDistance, k);
    else if (light.shadowType == 2) shadowFactor = pcss_shadow<COUNT>(s, pc, shadow_origin, light, lightDir, lightDistance, k);

    const f3 L = normalize_shade(lightDir);
    const f3 H = normalize_shade(V + L);
    const f3 radiance = (light.color * attenuation) * light.intensity;
    k.add(RT_CNT_PBR);
    Lo = Lo + compute_pbr(mat, N, V, L, H, radiance) * shadowFactor;
  }
  if (mat.subsurfaceScatter > 0.0f) Lo = Lo + subsurface<COUNT>(s, P, N, mat, k);
  return Lo;
}

// ---------------------------------------------------------------- main (glsl:509-584)
struct PixelResult { f3 color, P, N; int object_id; };

// One invocation of the shader's main() for pixel (px,py) and one frameCount.
template <bool COUNT>
__device__ __forceinline__ PixelResult shade_pixel(const KernelParams& kp, const SceneView& s, const Textures& tx,
                                                   unsigned px, unsigned py, int frame_count, Cnt<COUNT>& k) {
  const rt_frame_params& fp = kp.fp;
  PixelCtx pc;
  pc.px = px; pc.py = py; pc.frame_count = frame_count;
  pc.noise_r = fetch_noise(tx, (float)(px + (unsigned)frame_count) * fp.noise_scale[0],
                           (float)(py + (unsigned)frame_count) * fp.noise_scale[1]);
  // generateCameraRay (glsl:198-217); jitter = texel.xy*2-1 with .y = 0 for an R8 texture.
  k.add(RT_CNT_CAMERA);
  const float jx = pc.noise_r * 2.0f - 1.0f, jy = 0.0f * 2.0f - 1.0f;
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

  for (int depth = 0; depth < fp.max_ray_depth; ++depth) {
    const Hit h = trace_closest<COUNT>(s, ro, rd, k);
    if (h.index < 0) {                          // miss path (glsl:530-534)
      k.add(RT_CNT_MISS);
      const f3 sky = fp.use_skybox ? fetch_sky(tx, rd) : mk3(0.0f, 0.0f, 0.0f);
      out.color = out.color + throughput * sky;
      break;
    }
    const f3 N = hit_normal(s, h, ro, rd);
    const Mat mat = load_material(s, h.index);
    const f3 P = ro + rd * h.t;
    const f3 V = normalize_shade(-rd);
    out.P = P; out.N = N; out.object_id = h.index;

    const f3 Lo = compute_lighting<COUNT>(s, pc, P, N, mat, V, k);
    out.color = out.color + throughput * Lo;

    if (depth > 2) {                            // Russian roulette (glsl:544-549)
      k.add(RT_CNT_RR);
      const float diffuseWeight = sqrt_ieee(dot3(mat.albedo, mat.albedo)) * mat.diffuseStrength;
      const float continueProb =
          fminf(fmaxf(throughput.x, fmaxf(throughput.y, throughput.z)) * 0.95f + diffuseWeight, 0.99f);
      if (random2((float)(px + (unsigned)depth), (float)(py + (unsigned)depth)) > continueProb) break;
      throughput = mk3(div_ieee(throughput.x, continueProb), div_ieee(throughput.y, continueProb),
                       div_ieee(throughput.z, continueProb));
    }
    if (depth + 1 >= fp.max_ray_depth) {        // the next ray would never be traced
      k.add(mat.diffuseStrength > 0.0f ? RT_CNT_BOUNCE_DIFFUSE
                                       : (mat.transparency > 0.0f ? RT_CNT_BOUNCE_REFRACT : RT_CNT_BOUNCE_MIRROR));
      break;
    }

    const float F = fresnel_schlick(fmaxf(dot3(V, N), 0.0f), mat.ior);
    if (mat.diffuseStrength > 0.0f) {           // glsl:555-567
      k.add(RT_CNT_BOUNCE_DIFFUSE);
      const int hi = depth * 64 + frame_count;
      const f3 specularDir = reflect3(rd, N);
      const f3 diffuseDir = cosine_hemisphere(div_ieee((float)hi, 64.0f), halton(hi, 2), N);
      rd = normalize_ieee(mix3(specularDir, diffuseDir, mat.roughness));
      ro = P + N * 0.001f;
      throughput = throughput * (mat.albedo * mat.diffuseStrength);
    } else if (mat.transparency > 0.0f) {       // glsl:568-571
      k.add(RT_CNT_BOUNCE_REFRACT);
      rd = refraction_dir(rd, N, mat.ior);
      ro = P - N * 0.001f;
      throughput = throughput * ((mat.albedo * (1.0f - F)) * mat.transparency);
    } else {                                    // glsl:572-576
      k.add(RT_CNT_BOUNCE_MIRROR);
      rd = reflect3(rd, N);
      ro = P + N * 0.001f;
      throughput = throughput * (mat.albedo * F);
    }
  }
  return out;
}

// ---------------------------------------------------------------- kernel

template <bool COUNT>
__global__ void __launch_bounds__(kThreads) rt_render_kernel(const __grid_constant__ KernelParams kp) {
  extern __shared__ float4 smem[];
  const DeviceScene& sc = kp.sc;
  const rt_frame_params& fp = kp.fp;

  // ---- stage the scene once per CTA (SoA intersection set, plane frames, lights)
  SceneView s;
  s.material = sc.material;
  s.num_objects = sc.num_objects; s.num_lights = sc.num_lights;
  s.max_dist = fp.max_ray_distance;
  if (sc.scene_in_smem) {
    float4* d_isect = smem;
    float4* d_plane = d_isect + 3 * sc.num_objects;
    float4* d_light = d_plane + 3 * sc.num_planes;
    for (int i = threadIdx.x; i < 3 * sc.num_objects; i += kThreads) d_isect[i] = __ldg(sc.isect + i);
    for (int i = threadIdx.x; i < 3 * sc.num_planes; i += kThreads) d_plane[i] = __ldg(sc.plane + i);
    for (int i = threadIdx.x; i < 6 * sc.num_lights; i += kThreads) d_light[i] = __ldg(sc.lights + i);
    __syncthreads();
    s.isect = d_isect; s.plane = d_plane; s.lights = d_light;
  } else {
    s.isect = sc.isect; s.plane = sc.plane; s.lights = sc.lights;
  }
  Textures tx;
  tx.sky = sc.sky; tx.sky_tex = sc.sky_tex; tx.sky_size = sc.sky_size;
  tx.sky_hw = (fp.flags & RT_FLAG_SKYBOX_HW) != 0u && sc.sky_tex != 0ull;
  tx.noise = sc.noise; tx.noise_w = sc.noise_w; tx.noise_h = sc.noise_h;

  const int lane = threadIdx.x & 31;
  const int lx = lane & (kTileW - 1), ly = lane >> 3;
  const unsigned num_tiles = kp.tiles_x * kp.tiles_y;
  const int ns = fp.num_samples > 0 ? fp.num_samples : 1;
  const float inv_ns = rcp_ieee((float)ns);

  Cnt<COUNT> k;
  k.zero();

  // ---- persistent warp loop: each warp pulls 8x4 tiles from the global counter
  unsigned tile = 0;
  if (lane == 0) tile = atomicAdd(kp.tile_counter, 1u) - kp.tile_base;
  tile = __shfl_sync(0xffffffffu, tile, 0);
  while (tile < num_tiles) {
    unsigned next = 0;
    if (lane == 0) next = atomicAdd(kp.tile_counter, 1u) - kp.tile_base;   // prefetch the next tile id
    const unsigned ty = tile / kp.tiles_x, txi = tile - ty * kp.tiles_x;
    const int px = (int)txi * kTileW + lx;
    const int py = window_row(fp, (int)ty * kTileH) + ly;
    if (px < fp.width && py < fp.row_end) {
      PixelResult r = shade_pixel<COUNT>(kp, s, tx, (unsigned)px, (unsigned)py, fp.frame_count, k);
      for (int sidx = 1; sidx < ns; ++sidx) {            // colour: sequential FP32 sum over samples
        const PixelResult r2 = shade_pixel<COUNT>(kp, s, tx, (unsigned)px, (unsigned)py, fp.frame_count + sidx, k);
        r.color = r.color + r2.color;
      }
      const size_t at = (size_t)py * fp.width + px;
      if (kp.color) {
        float4 c = make_float4(r.color.x, r.color.y, r.color.z, 1.0f);
        if (fp.flags & RT_FLAG_ACCUMULATE) {             // partial sums for sample-range sharding
          const float4 old = kp.color[at];
          c = make_float4(old.x + c.x, old.y + c.y, old.z + c.z, 1.0f);
        } else if (ns > 1 && !(fp.flags & RT_FLAG_SUM_SAMPLES)) {
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
    tile = __shfl_sync(0xffffffffu, next, 0);
  }

  if (COUNT) {                                           // warp-reduce, one atomic per counter per warp
    for (int i = 0; i < RT_CNT_COUNT; ++i) {
      unsigned long long v = k.get(i);
      for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
      if (lane == 0 && v) atomicAdd(kp.counters + i, v);
    }
  }
}


inline cudaError_t launch_render(const KernelParams& kp, int sm_count, cudaStream_t stream, unsigned* warps_launched) {
  const bool count = (kp.fp.flags & RT_FLAG_COUNTERS) != 0u;
  const size_t smem = kp.sc.scene_in_smem ? scene_smem_bytes(kp.sc.num_objects, kp.sc.num_planes, kp.sc.num_lights) : 0;
  auto kern = count ? rt_render_kernel<true> : rt_render_kernel<false>;
  int per_sm = 0;
  cudaError_t e = count ? cached_occupancy(rt_render_kernel<true>, kThreads, smem, &per_sm) : cached_occupancy(rt_render_kernel<false>, kThreads, smem, &per_sm);
  if (e != cudaSuccess) return e;
  const unsigned num_tiles = kp.tiles_x * kp.tiles_y;
  unsigned ctas = (unsigned)(sm_count * per_sm);
  const unsigned needed = (num_tiles + kWarpsPerCta - 1) / kWarpsPerCta;
  if (ctas > needed) ctas = needed;
  if (ctas < 1) ctas = 1;
  *warps_launched = ctas * kWarpsPerCta;
  kern<<<ctas, kThreads, smem, stream>>>(kp);
  return cudaGetLastError();
}

}  // namespace RT_VARIANT_NS
}  // namespace rt
