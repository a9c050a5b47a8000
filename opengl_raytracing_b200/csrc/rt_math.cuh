// rt_math.cuh — device helpers shared by the two kernels of librt_b200.so: vector algebra,
// the math policy (IEEE "decision" forms vs SFU "shading" forms), texture fetches (noise,
// cubemap), the shader's samplers and BRDF. Each function cites the shader lines it restates
// (/root/reference/shader/raytracingCs.glsl). Included inside a per-variant namespace:
// RT_PARITY 1 = IEEE everywhere, RT_PARITY 0 = SFU approximations in colour-only code.
#pragma once
#include <cuda_fp16.h>
#include "rt_device.h"

#ifndef RT_PARITY
#error "define RT_PARITY to 0 or 1 before including rt_math.cuh"
#endif

namespace rt {
namespace RT_VARIANT_NS {

constexpr float kPi = 3.14159265359f;            // glsl:6

// Row of the frame for local row `lr` of this launch's tile grid (lr is a multiple of 4):
// contiguous window, or interleaved bands (rt_frame_params.row_band / row_period / row_phase).
__device__ __forceinline__ int window_row(const rt_frame_params& fp, int lr) {
  if (fp.row_band <= 0) return fp.row_begin + lr;
  const int k = lr / fp.row_band;
  return fp.row_begin + (k * fp.row_period + fp.row_phase) * fp.row_band + (lr - k * fp.row_band);
}
// ---------------------------------------------------------------- small vector algebra
struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ f3 operator*(float s, f3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ f3 operator+(f3 a, float s) { return mk3(a.x + s, a.y + s, a.z + s); }
__device__ __forceinline__ f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ f3 cross3(f3 a, f3 b) {
  return mk3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
__device__ __forceinline__ f3 xyz(float4 v) { return mk3(v.x, v.y, v.z); }

// Math policy. "Decision" sites (slab test, sphere/plane roots, TIR test, Russian roulette)
// always use IEEE forms; shading-only sites may use the SFU approximations in the fast TU.
__device__ __forceinline__ float rcp_ieee(float x) { return __frcp_rn(x); }            // == 1.0f / x
__device__ __forceinline__ float div_ieee(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float sqrt_ieee(float x) { return __fsqrt_rn(x); }
#if RT_PARITY
__device__ __forceinline__ float div_shade(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float sqrt_shade(float x) { return __fsqrt_rn(x); }
__device__ __forceinline__ float inversesqrt_shade(float x) { return __frcp_rn(__fsqrt_rn(x)); }
__device__ __forceinline__ float pow5(float x) { return powf(x, 5.0f); }
#else
// Colour-only sites: single MUFU forms (.ftz: no denormal-range fix-up sequence; a denormal
// operand or result is a colour term below 1.2e-38). Measured on C2: -4 % kernel time for the
// shadow-ray pair of these alone (profiles/README.md).
#if RT_PK_NO_FTZ
__device__ __forceinline__ float div_shade(float a, float b) { return __fdividef(a, b); }
__device__ __forceinline__ float sqrt_shade(float x) { return sqrtf(x); }
__device__ __forceinline__ float inversesqrt_shade(float x) { return rsqrtf(x); }
#else
__device__ __forceinline__ float div_shade(float a, float b) { float r; asm("div.approx.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float sqrt_shade(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float inversesqrt_shade(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
#endif
__device__ __forceinline__ float pow5(float x) { float x2 = x * x; return x2 * x2 * x; }
#endif
__device__ __forceinline__ float pow2(float x) { return x * x; }                       // pow(x, 2.0)
// normalize(v) = v * inversesqrt(dot(v, v))  (pinned, glsl_shim.h)
__device__ __forceinline__ f3 normalize_ieee(f3 a) { return a * __frcp_rn(__fsqrt_rn(dot3(a, a))); }
__device__ __forceinline__ f3 normalize_shade(f3 a) { return a * inversesqrt_shade(dot3(a, a)); }
__device__ __forceinline__ float fract_(float a) { return a - floorf(a); }
__device__ __forceinline__ f3 mix3(f3 a, f3 b, float t) { return a * (1.0f - t) + b * t; }
__device__ __forceinline__ f3 reflect3(f3 I, f3 N) { return I - (2.0f * dot3(N, I)) * N; }

// Multiply-add of colour-only and boolean-only code (BRDF terms, light evaluation, texture
// filtering, shadow-ray set-up: nothing that feeds P, N or a bounce direction). Parity build: the
// shader's two unfused operations, same operands (a*b + c; IEEE addition commutes, so the call
// sites keep their bits). Fast build: one FFMA. The -fmad=false of the translation unit still
// protects everything written with plain * and +.
#if RT_PARITY
__device__ __forceinline__ float smad(float a, float b, float c) { return a * b + c; }
#else
__device__ __forceinline__ float smad(float a, float b, float c) { return fmaf(a, b, c); }
#endif
__device__ __forceinline__ float sdot3(f3 a, f3 b) { return smad(a.z, b.z, smad(a.y, b.y, a.x * b.x)); }   // == dot3 in the parity build
__device__ __forceinline__ f3 smad3(f3 a, float s, f3 c) { return mk3(smad(a.x, s, c.x), smad(a.y, s, c.y), smad(a.z, s, c.z)); }
__device__ __forceinline__ f3 smad3(f3 a, f3 b, f3 c) { return mk3(smad(a.x, b.x, c.x), smad(a.y, b.y, c.y), smad(a.z, b.z, c.z)); }
__device__ __forceinline__ f3 smix3(f3 a, f3 b, float t) { return smad3(b, t, a * (1.0f - t)); }                // == mix3
__device__ __forceinline__ f3 normalize_color(f3 a) { return a * inversesqrt_shade(sdot3(a, a)); }              // == normalize_shade


struct Mat {
  f3 albedo; float metallic, roughness, diffuseStrength, ior, transparency;
  f3 subsurfaceColor; float subsurfaceScatter, scatterDistance;
};

// ---------------------------------------------------------------- texture fetches
struct Textures {
  const uint2* sky; unsigned long long sky_tex; int sky_size; bool sky_hw;
  const unsigned char* noise; int noise_w, noise_h;
};

__device__ __forceinline__ int wrap_repeat(int i, int n) {
  if ((n & (n - 1)) == 0) return i & (n - 1);            // power of two (the usual 1024): same value as the modulo
  int m = i % n; return m < 0 ? m + n : m;
}

// texture(blueNoiseTex, uv).r — GL_R8, REPEAT, LINEAR with FP32 weights
// (src/ForwardShadingPipeline.cpp:39-49); unbound -> 0. .g is always 0.
__device__ __forceinline__ float fetch_noise(const Textures& tx, float uvx, float uvy) {
  if (!tx.noise) return 0.0f;
  float u = uvx * (float)tx.noise_w - 0.5f, v = uvy * (float)tx.noise_h - 0.5f;
  if (!(u == u) || fabsf(u) > 1.0e9f) u = 0.0f;
  if (!(v == v) || fabsf(v) > 1.0e9f) v = 0.0f;
  const float fu = floorf(u), fv = floorf(v);
  const float a = u - fu, w = v - fv;
  const int i0 = wrap_repeat((int)fu, tx.noise_w), i1 = wrap_repeat((int)fu + 1, tx.noise_w);
  const int j0 = wrap_repeat((int)fv, tx.noise_h), j1 = wrap_repeat((int)fv + 1, tx.noise_h);
  const float k = 1.0f / 255.0f;
  const float t00 = (float)__ldg(tx.noise + j0 * tx.noise_w + i0) * k, t10 = (float)__ldg(tx.noise + j0 * tx.noise_w + i1) * k;
  const float t01 = (float)__ldg(tx.noise + j1 * tx.noise_w + i0) * k, t11 = (float)__ldg(tx.noise + j1 * tx.noise_w + i1) * k;
  const float top = t00 * (1.0f - a) + t10 * a;
  const float bot = t01 * (1.0f - a) + t11 * a;
  return top * (1.0f - w) + bot * w;
}

__device__ __forceinline__ f3 unpack_rgb16f(uint2 p) {
  const float2 rg = __half22float2(*reinterpret_cast<const __half2*>(&p.x));
  const float2 ba = __half22float2(*reinterpret_cast<const __half2*>(&p.y));
  return mk3(rg.x, rg.y, ba.x);
}

// texture(skybox, dir).rgb — GL cube map face selection (OpenGL 4.3 table 8.19, ties
// x > y > z), per-face CLAMP_TO_EDGE, LINEAR, not seamless (src/TextureLoader.cpp:143-147).
// Default: 4 RGBA16F texel loads + FP32 lerp (exact). RT_FLAG_SKYBOX_HW: the CUDA cubemap
// texture object (9-bit hardware filter weights, seamless edges; deviation in DESIGN.md).
__device__ __forceinline__ f3 fetch_sky(const Textures& tx, f3 r) {
  if (tx.sky_hw) {
    const float4 v = texCubemap<float4>((cudaTextureObject_t)tx.sky_tex, r.x, r.y, r.z);
    return mk3(v.x, v.y, v.z);
  }
  if (!tx.sky) return mk3(0.0f, 0.0f, 0.0f);
  const float ax = fabsf(r.x), ay = fabsf(r.y), az = fabsf(r.z);
  int face; float sc, tc, ma;
  if (ax >= ay && ax >= az) { ma = ax; if (r.x >= 0.0f) { face = 0; sc = -r.z; tc = -r.y; } else { face = 1; sc = r.z; tc = -r.y; } }
  else if (ay >= az)        { ma = ay; if (r.y >= 0.0f) { face = 2; sc = r.x; tc = r.z; } else { face = 3; sc = r.x; tc = -r.z; } }
  else                      { ma = az; if (r.z >= 0.0f) { face = 4; sc = r.x; tc = -r.y; } else { face = 5; sc = -r.x; tc = -r.y; } }
  const int S = tx.sky_size;
  const float rma = div_shade(1.0f, ma);               // colour-only: one MUFU.RCP in the fast build, 1/ma * sc ...
#if RT_PARITY
  float u = smad(0.5f * (div_ieee(sc, ma) + 1.0f), (float)S, -0.5f);
  float v = smad(0.5f * (div_ieee(tc, ma) + 1.0f), (float)S, -0.5f);
#else
  float u = smad(smad(sc, rma, 1.0f) * 0.5f, (float)S, -0.5f);
  float v = smad(smad(tc, rma, 1.0f) * 0.5f, (float)S, -0.5f);
#endif
  if (!(u == u)) u = 0.0f;
  if (!(v == v)) v = 0.0f;
  const float hi = (float)(S - 1);
  float fu = floorf(u), fv = floorf(v);
  const float a = u - fu, w = v - fv;
  float fu1 = fu + 1.0f, fv1 = fv + 1.0f;
  fu = fu < 0.0f ? 0.0f : (fu > hi ? hi : fu);  fu1 = fu1 < 0.0f ? 0.0f : (fu1 > hi ? hi : fu1);
  fv = fv < 0.0f ? 0.0f : (fv > hi ? hi : fv);  fv1 = fv1 < 0.0f ? 0.0f : (fv1 > hi ? hi : fv1);
  const int i0 = (int)fu, i1 = (int)fu1, j0 = (int)fv, j1 = (int)fv1;
  const uint2* f = tx.sky + (size_t)face * S * S;
  const f3 t00 = unpack_rgb16f(__ldg(f + j0 * S + i0)), t10 = unpack_rgb16f(__ldg(f + j0 * S + i1));
  const f3 t01 = unpack_rgb16f(__ldg(f + j1 * S + i0)), t11 = unpack_rgb16f(__ldg(f + j1 * S + i1));
  return smix3(smix3(t00, t10, a), smix3(t01, t11, a), w);
}

// ---------------------------------------------------------------- samplers (glsl:273-313)

// random (glsl:273-275). The sine is evaluated in FP64 and rounded: it decides Russian
// roulette, and 1 ulp of sinf moves fract(sin*43758.5) by 2.6e-3.
#ifndef RT_INLINE_RANDOM
#define RT_INLINE_RANDOM 1
#endif
#if RT_INLINE_RANDOM
__device__ __forceinline__
#else
__device__ __noinline__
#endif
float random2(float sx, float sy) {
  const float arg = sx * 12.9898f + sy * 78.233f;
  return fract_((float)sin((double)arg) * 43758.5453123f);
}

// haltonSequence (glsl:278-288): radical inverse by repeated FP32 division (IEEE).
__device__ __forceinline__ float halton(int index, int base) {
  float result = 0.0f;
  const float fb = (float)base;
  float f = div_ieee(1.0f, fb);
  int i = index;
  while (i > 0) {
    result = result + f * (float)(i % base);
    i = i / base;
    f = div_ieee(f, fb);
  }
  return result;
}

// cosineWeightedHemisphere (glsl:291-308). Its result is the diffuse-bounce (and SSS) ray DIRECTION, i.e. part of
// the geometric chain ray -> t -> P -> N: IEEE square roots and normalisation in both builds, and sin / cos
// evaluated in FP64 and rounded (like random2 below): sincosf is 2 ulp off, and every curved mirror after a
// diffuse bounce multiplies that by ~2 * distance / radius. Out of line: it runs once per diffuse bounce, and
// scenes without diffuse materials (the shipped ones) keep it out of their instruction stream and registers.
__device__ __noinline__ f3 cosine_hemisphere(float rx, float ry, f3 normal) {
  const float phi = 2.0f * kPi * rx;
  const float cosTheta = sqrt_ieee(ry);
  const float sinTheta = sqrt_ieee(1.0f - ry);
  double sd, cd;
  sincos((double)phi, &sd, &cd);
  const float sp = (float)sd, cp = (float)cd;
  const f3 h = mk3(sinTheta * cp, cosTheta, sinTheta * sp);
  const f3 tangent = normalize_ieee(cross3(normal, mk3(0.0f, 1.0f, 1.0f)));
  const f3 bitangent = cross3(normal, tangent);
  return normalize_ieee(tangent * h.x + bitangent * h.z + normal * h.y);
}

// Halton tables for the fixed-index call sites (PCF i<pcfSamples bases 2,3; PCSS i<16 base 3),
// filled by the host with the same IEEE FP32 operations as halton() above.
constexpr int kHaltonTable = 64;
__constant__ float c_halton2[kHaltonTable];
__constant__ float c_halton3[kHaltonTable];
__device__ __forceinline__ float halton2(int i) { return i < kHaltonTable ? c_halton2[i] : halton(i, 2); }
__device__ __forceinline__ float halton3(int i) { return i < kHaltonTable ? c_halton3[i] : halton(i, 3); }

// ---------------------------------------------------------------- BRDF (glsl:219-270)

// fresnelSchlick (glsl:220-223)
__device__ __forceinline__ float fresnel_schlick(float cosTheta, float ior) {
  const float r0 = pow2(div_shade(1.0f - ior, 1.0f + ior));
  return smad(1.0f - r0, pow5(1.0f - cosTheta), r0);
}

// computePBR (glsl:226-253): GGX NDF, Schlick-GGX geometry, Schlick Fresnel, kD/F split.
__device__ __forceinline__ f3 compute_pbr(const Mat& mat, f3 N, f3 V, f3 L, f3 H, f3 radiance) {
  const float alpha = pow2(mat.roughness);
  const float a2 = alpha * alpha;
  const float NdotH = fmaxf(sdot3(N, H), 0.0f);
  const float NDF = div_shade(a2, kPi * pow2(smad(pow2(NdotH), a2 - 1.0f, 1.0f)));
  const float k = pow2(mat.roughness + 1.0f) / 8.0f;
  const float NdotV = fmaxf(sdot3(N, V), 0.0f), NdotL = fmaxf(sdot3(N, L), 0.0f);
  float G = div_shade(NdotV, smad(NdotV, 1.0f - k, k));
  G *= div_shade(NdotL, smad(NdotL, 1.0f - k, k));
  const f3 F0 = smix3(mk3(0.04f, 0.04f, 0.04f), mat.albedo, mat.metallic);
  const float p5 = pow5(1.0f - fmaxf(sdot3(H, V), 0.0f));
  const f3 F = smad3(mk3(1.0f - F0.x, 1.0f - F0.y, 1.0f - F0.z), p5, F0);
  const f3 numerator = (NDF * G) * F;
  const float denominator = 4.0f * NdotV * NdotL;
  const float dd = fmaxf(denominator, 0.001f);
  const f3 specular = mk3(div_shade(numerator.x, dd), div_shade(numerator.y, dd), div_shade(numerator.z, dd));
  const f3 kD = mk3(1.0f - F.x, 1.0f - F.y, 1.0f - F.z) * (1.0f - mat.metallic);
  const f3 kda = kD * mat.albedo;
  const f3 diffuse = mk3(div_shade(kda.x, kPi), div_shade(kda.y, kPi), div_shade(kda.z, kPi));
  return ((diffuse + specular) * radiance) * NdotL;
}

// calculateRefraction (glsl:256-270), refract() per GLSL 4.30 §8.5. The TIR decision and the
// new direction feed the next intersection -> IEEE forms in both variants.
#ifndef RT_INLINE_REFRACT
#define RT_INLINE_REFRACT 1
#endif
#if RT_INLINE_REFRACT
__device__ __forceinline__
#else
__device__ __noinline__
#endif
f3 refraction_dir(f3 d, f3 N, float ior) {
  const bool entering = dot3(d, N) < 0.0f;
  const float eta = entering ? rcp_ieee(ior) : ior;
  const f3 n = entering ? N : -N;
  const f3 I = normalize_ieee(d);
  const float c = dot3(n, I);
  const float kk = 1.0f - eta * eta * (1.0f - c * c);
  f3 r = mk3(0.0f, 0.0f, 0.0f);
  if (!(kk < 0.0f)) r = eta * I - (eta * c + sqrt_ieee(kk)) * n;
  if (dot3(r, r) < 0.001f) r = reflect3(d, n);          // total internal reflection
  return r;
}

__device__ __forceinline__ uint2 pack_half4(float x, float y, float z, float w) {
  const __half2 a = __floats2half2_rn(x, y), b = __floats2half2_rn(z, w);
  uint2 r;
  r.x = *reinterpret_cast<const unsigned*>(&a);
  r.y = *reinterpret_cast<const unsigned*>(&b);
  return r;
}

inline cudaError_t upload_tables(const float* h2, const float* h3) {
  cudaError_t e = cudaMemcpyToSymbol(c_halton2, h2, sizeof(float) * kHaltonTable);
  if (e != cudaSuccess) return e;
  return cudaMemcpyToSymbol(c_halton3, h3, sizeof(float) * kHaltonTable);
}

}  // namespace RT_VARIANT_NS
}  // namespace rt
