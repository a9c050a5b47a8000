// rt_oracle.cpp — TEST INFRASTRUCTURE: the CPU oracle ("port").
//
// A plain scalar C++ restatement of the reference's per-pixel ray-tracing compute pass,
// /root/reference/shader/raytracingCs.glsl, function by function, each citing the lines it
// follows. It is the checker the CUDA path is compared with and the source of the
// reference-literal op counters (flop model v0); it is NOT product code:
// only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load it, and
// the product library (librt_b200.so) fails loudly rather than fall back to it.
//
// Parity status: PINNED against the reference's own shader source compiled as C++
// (oracle/_ref, built by make_ref.py): tests/test_oracle.py requires this file and _ref to
// agree bit-for-bit on every config, and tests/golden/ holds _ref-generated tiles. The
// reference ships no tests or golden vectors of its own, and its GLSL cannot be executed
// through a GL driver on these machines (unavailable on this box).
//
// Arithmetic: FP32 only, evaluated in the shader's written order; build with
// -ffp-contract=off. Implementation-defined GLSL behaviour is pinned as in glsl_shim.h
// (min/max = IEEE minNum/maxNum, normalize = v * (1/sqrt(v.v)), C powf, zero-initialised
// P/N, FP32 bilinear texture filtering).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <omp.h>
#include "oracle_api.h"

namespace {

struct V3 { float x, y, z; };
struct V2 { float x, y; };

inline V3 mk(float x, float y, float z) { V3 v = {x, y, z}; return v; }
inline V3 ld3(const float* p) { return mk(p[0], p[1], p[2]); }
inline V3 add(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 sub(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 mul(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 scale(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline V3 scale_l(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
inline V3 divs(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
inline V3 adds(V3 a, float s) { return mk(a.x + s, a.y + s, a.z + s); }
inline V3 neg(V3 a) { return mk(-a.x, -a.y, -a.z); }
inline float dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross3(V3 a, V3 b) {
  return mk(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
// GLSL built-ins with the pinned semantics (glsl_shim.h header comment).
inline float fmin_(float a, float b) { return (a != a) ? b : ((b != b) ? a : (b < a ? b : a)); }
inline float fmax_(float a, float b) { return (a != a) ? b : ((b != b) ? a : (a < b ? b : a)); }
inline V3 normalize3(V3 a) { float inv = 1.0f / std::sqrt(dot3(a, a)); return scale(a, inv); }
inline float fract_(float a) { return a - std::floor(a); }
inline V3 mix3(V3 a, V3 b, float t) { return add(scale(a, 1.0f - t), scale(b, t)); }
inline V3 reflect3(V3 I, V3 N) { return sub(I, scale_l(2.0f * dot3(N, I), N)); }
inline V3 refract3(V3 I, V3 N, float eta) {
  float d = dot3(N, I);
  float k = 1.0f - eta * eta * (1.0f - d * d);
  if (k < 0.0f) return mk(0.f, 0.f, 0.f);
  return sub(scale_l(eta, I), scale_l(eta * d + std::sqrt(k), N));
}

const float PI = 3.14159265359f;   // glsl:6

struct Ray { V3 origin, direction; };

// Everything one oracle_render call binds: the SSBOs, uniforms and samplers of glsl:60-89.
struct Bindings {
  const rt_object* objects; int numObjects;
  const rt_light* lights; int numLights;
  V3 cameraPos, cameraDir, cameraUp, cameraRight;
  float fov, focalLength, maxRayDistance;
  bool useSkybox;
  const float* sky; int skySize;
  const unsigned char* noise; int noiseW, noiseH;
  V2 noiseScale;
  int frameCount, maxRayDepth, width, height;
};

struct Counters { uint64_t c[RT_CNT_COUNT]; };

// ------------------------------------------------------------------ texture fetches
inline int wrap_repeat(int i, int n) { int m = i % n; return m < 0 ? m + n : m; }

// texture(blueNoiseTex, uv): GL_R8 / REPEAT / LINEAR (src/ForwardShadingPipeline.cpp:39-49);
// unbound => (0,0,0,1). Returns .r; .g is always 0.
float fetch_noise(const Bindings& b, V2 uv) {
  if (!b.noise) return 0.f;
  float u = uv.x * (float)b.noiseW - 0.5f, v = uv.y * (float)b.noiseH - 0.5f;
  if (!(u == u) || std::fabs(u) > 1.0e9f) u = 0.f;
  if (!(v == v) || std::fabs(v) > 1.0e9f) v = 0.f;
  float fu = std::floor(u), fv = std::floor(v);
  float a = u - fu, w = v - fv;
  int i0 = wrap_repeat((int)fu, b.noiseW), i1 = wrap_repeat((int)fu + 1, b.noiseW);
  int j0 = wrap_repeat((int)fv, b.noiseH), j1 = wrap_repeat((int)fv + 1, b.noiseH);
  const float k = 1.0f / 255.0f;
  float t00 = (float)b.noise[j0 * b.noiseW + i0] * k, t10 = (float)b.noise[j0 * b.noiseW + i1] * k;
  float t01 = (float)b.noise[j1 * b.noiseW + i0] * k, t11 = (float)b.noise[j1 * b.noiseW + i1] * k;
  float top = t00 * (1.0f - a) + t10 * a;
  float bot = t01 * (1.0f - a) + t11 * a;
  return top * (1.0f - w) + bot * w;
}

// texture(skybox, dir).rgb: cube map built by src/TextureLoader.cpp:118-194 (LINEAR,
// CLAMP_TO_EDGE, not seamless); face selection per OpenGL 4.3 table 8.19, ties x > y > z.
V3 fetch_sky(const Bindings& b, V3 r) {
  if (!b.sky) return mk(0.f, 0.f, 0.f);
  float ax = std::fabs(r.x), ay = std::fabs(r.y), az = std::fabs(r.z);
  int face; float sc, tc, ma;
  if (ax >= ay && ax >= az) { ma = ax; if (r.x >= 0.f) { face = 0; sc = -r.z; tc = -r.y; } else { face = 1; sc = r.z; tc = -r.y; } }
  else if (ay >= az)        { ma = ay; if (r.y >= 0.f) { face = 2; sc = r.x; tc = r.z; } else { face = 3; sc = r.x; tc = -r.z; } }
  else                      { ma = az; if (r.z >= 0.f) { face = 4; sc = r.x; tc = -r.y; } else { face = 5; sc = -r.x; tc = -r.y; } }
  const int S = b.skySize;
  float u = 0.5f * (sc / ma + 1.0f) * (float)S - 0.5f;
  float v = 0.5f * (tc / ma + 1.0f) * (float)S - 0.5f;
  if (!(u == u)) u = 0.f;
  if (!(v == v)) v = 0.f;
  float hi = (float)(S - 1);
  float fu = std::floor(u), fv = std::floor(v);
  float a = u - fu, w = v - fv;
  float fu1 = fu + 1.0f, fv1 = fv + 1.0f;
  fu = fu < 0.f ? 0.f : (fu > hi ? hi : fu);  fu1 = fu1 < 0.f ? 0.f : (fu1 > hi ? hi : fu1);
  fv = fv < 0.f ? 0.f : (fv > hi ? hi : fv);  fv1 = fv1 < 0.f ? 0.f : (fv1 > hi ? hi : fv1);
  int i0 = (int)fu, i1 = (int)fu1, j0 = (int)fv, j1 = (int)fv1;
  const float* f = b.sky + (size_t)face * S * S * 4;
  const float* t00 = f + ((size_t)j0 * S + i0) * 4; const float* t10 = f + ((size_t)j0 * S + i1) * 4;
  const float* t01 = f + ((size_t)j1 * S + i0) * 4; const float* t11 = f + ((size_t)j1 * S + i1) * 4;
  float o[3];
  for (int c = 0; c < 3; ++c) {
    float top = t00[c] * (1.0f - a) + t10[c] * a;
    float bot = t01[c] * (1.0f - a) + t11[c] * a;
    o[c] = top * (1.0f - w) + bot * w;
  }
  return mk(o[0], o[1], o[2]);
}

// ------------------------------------------------------------------ intersection (glsl:91-196)

// intersectAABB (glsl:91-103): slab test on 1/dir with IEEE inf/NaN semantics; tolerates
// zero-thickness and min > max boxes (planes, src/SceneIO.h:94-103).
bool intersectAABB(const Bindings& b, const Ray& ray, const rt_aabb& box) {
  V3 inv = mk(1.0f / ray.direction.x, 1.0f / ray.direction.y, 1.0f / ray.direction.z);
  V3 t0 = mul(sub(ld3(box.min), ray.origin), inv);
  V3 t1 = mul(sub(ld3(box.max), ray.origin), inv);
  V3 lo = mk(fmin_(t0.x, t1.x), fmin_(t0.y, t1.y), fmin_(t0.z, t1.z));
  V3 hi = mk(fmax_(t0.x, t1.x), fmax_(t0.y, t1.y), fmax_(t0.z, t1.z));
  float tMin = fmax_(fmax_(lo.x, lo.y), lo.z);
  float tMax = fmin_(fmin_(hi.x, hi.y), hi.z);
  return tMax >= tMin && tMin < b.maxRayDistance && tMax > 0.0f;
}

// intersectSphere (glsl:105-118): near root only, unnormalised direction.
bool intersectSphere(const Ray& ray, const rt_object& obj, float& t) {
  V3 oc = sub(ray.origin, ld3(obj.position));
  float a = dot3(ray.direction, ray.direction);
  float bq = 2.0f * dot3(oc, ray.direction);
  float c = dot3(oc, oc) - obj.radius * obj.radius;
  float disc = bq * bq - 4.0f * a * c;
  if (disc < 0.0f) return false;
  t = (-bq - std::sqrt(disc)) / (2.0f * a);
  return t > 0.0f;
}

// intersectPlane (glsl:120-153): finite two-sided rectangle; the in-plane frame is rebuilt
// on every call in the reference. *full reports whether the frame build was reached.
bool intersectPlane(const Ray& ray, const rt_object& obj, float& t, bool* full) {
  *full = false;
  V3 n = ld3(obj.normal);
  float denom = dot3(n, ray.direction);
  if (std::fabs(denom) > 1e-6f) {
    t = dot3(sub(ld3(obj.position), ray.origin), n) / denom;
    if (t < 0.0f) return false;
    *full = true;
    V3 hitPoint = add(ray.origin, scale(ray.direction, t));
    V3 right, forward;
    if (std::fabs(n.y) > 0.9f) {
      right = normalize3(cross3(n, mk(0.f, 0.f, 1.f)));
      forward = normalize3(cross3(right, n));
    } else {
      right = normalize3(cross3(n, mk(0.f, 1.f, 0.f)));
      forward = normalize3(cross3(right, n));
    }
    V3 local = sub(hitPoint, ld3(obj.position));
    float x = dot3(local, right);
    float z = dot3(local, forward);
    if (std::fabs(x) > obj.size[0] / 2.0f || std::fabs(z) > obj.size[1] / 2.0f) return false;
    return true;
  }
  return false;
}

// ---- grazing margin (test instrumentation, does not touch the rendering arithmetic). While a pixel's
// margin is tracked (oracle_margins), every closest-hit ray of its path (primary and bounce rays — the calls
// that pass hitIndex) records the smallest RELATIVE margin by which one of its geometric decisions was taken:
//   sphere      |disc| / max(b^2, |4ac|)                      the hit/miss decision of glsl:112
//   rectangle   | |x| - size/2 | / (size/2), same for z        the in-rectangle decision of glsl:147 (t >= 0 only)
//               (both only for objects whose hit, or near miss, would lie in front of the closest hit)
//   ordering    (t2 - t1) / t2 of the two nearest hits         the closest-hit decision of glsl:178
// A pixel whose object id / hit mask differs between the kernel and the oracle must have a margin below the
// stated epsilon (tests/parity.py GRAZING_EPS): that is the "stated grazing-angle pixels" set of north_star.
thread_local float* tl_margin = nullptr;
inline void margin_min(float v) { if (tl_margin && v == v && v < *tl_margin) *tl_margin = v; }
// limit = the closest hit found (or maxRayDistance): an object whose near-miss lies beyond it decides nothing.
void margin_of_object(const Ray& ray, const rt_object& obj, float limit) {
  if (obj.type == 0) {
    V3 oc = sub(ray.origin, ld3(obj.position));
    float a = dot3(ray.direction, ray.direction);
    float bq = 2.0f * dot3(oc, ray.direction);
    float c = dot3(oc, oc) - obj.radius * obj.radius;
    float disc = bq * bq - 4.0f * a * c;
    float den = std::fmax(bq * bq, std::fabs(4.0f * a * c));
    // spheres behind the ray, or whose closest approach lies beyond the closest hit, decide nothing
    if (den > 0.0f && bq < 0.0f && -bq / (2.0f * a) <= limit * 1.001f) margin_min(std::fabs(disc) / den);
  } else if (obj.type == 1) {
    V3 n = ld3(obj.normal);
    float denom = dot3(n, ray.direction);
    if (!(std::fabs(denom) > 1e-6f)) return;
    float t = dot3(sub(ld3(obj.position), ray.origin), n) / denom;
    if (!(t >= 0.0f) || !(t <= limit * 1.001f)) return;
    V3 right = std::fabs(n.y) > 0.9f ? normalize3(cross3(n, mk(0.f, 0.f, 1.f))) : normalize3(cross3(n, mk(0.f, 1.f, 0.f)));
    V3 forward = normalize3(cross3(right, n));
    V3 local = sub(add(ray.origin, scale(ray.direction, t)), ld3(obj.position));
    float hx = obj.size[0] / 2.0f, hz = obj.size[1] / 2.0f;
    float mx = std::fabs(std::fabs(dot3(local, right)) - hx) / std::fmax(hx, 1e-30f);
    float mz = std::fabs(std::fabs(dot3(local, forward)) - hz) / std::fmax(hz, 1e-30f);
    // only the coordinate that decides: inside in the other one
    if (std::fabs(dot3(local, forward)) <= hz * 1.001f) margin_min(mx);
    if (std::fabs(dot3(local, right)) <= hx * 1.001f) margin_min(mz);
  }
}

// intersectObjects (glsl:155-196): closest hit over ALL objects in array order, strict
// `<` (first object wins ties). On a miss *mat and *N are left untouched.
bool intersectObjects(const Bindings& b, Counters& k, const Ray& ray, const rt_material** mat, V3* N, float* tOut,
                      int* hitIndex) {
  float minT = b.maxRayDistance;
  bool hit = false;
  k.c[RT_CNT_RAYS]++;
  const bool track = tl_margin && hitIndex;
  float t1 = INFINITY, t2 = INFINITY;                       // the two nearest hits (instrumentation only)
  for (int i = 0; i < b.numObjects; ++i) {
    const rt_object& obj = b.objects[i];
    k.c[RT_CNT_AABB]++;
    if (!intersectAABB(b, ray, obj.bounds)) continue;
    float currentT = 0.f;
    bool isHit = false;
    if (obj.type == 0) {
      k.c[RT_CNT_SPHERE]++;
      isHit = intersectSphere(ray, obj, currentT);
    } else if (obj.type == 1) {
      bool full;
      isHit = intersectPlane(ray, obj, currentT, &full);
      k.c[full ? RT_CNT_PLANE_FULL : RT_CNT_PLANE_EARLY]++;
    }
    if (track && isHit && currentT > 0.0f) {
      if (currentT < t1) { t2 = t1; t1 = currentT; } else if (currentT < t2) t2 = currentT;
    }
    if (isHit && currentT > 0.0f && currentT < minT) {
      minT = currentT;
      hit = true;
      *mat = &obj.material;
      if (hitIndex) *hitIndex = i;
      if (obj.type == 0) {
        k.c[RT_CNT_HIT_UPDATE_SPHERE]++;
        *N = normalize3(sub(add(ray.origin, scale(ray.direction, currentT)), ld3(obj.position)));
      } else {
        k.c[RT_CNT_HIT_UPDATE_PLANE]++;
        *N = ld3(obj.normal);
      }
    }
  }
  if (track) {
    if (t2 < INFINITY) margin_min((t2 - t1) / t2);
    for (int i = 0; i < b.numObjects; ++i) margin_of_object(ray, b.objects[i], minT);
  }
  *tOut = minT;
  return hit;
}

// generateCameraRay (glsl:198-217). Pixel (0,0) is the bottom-left: uv.y grows along cameraUp.
Ray generateCameraRay(const Bindings& b, unsigned px, unsigned py, V2 jitter) {
  float ux = ((float)(int)px + 0.5f + jitter.x) / (float)b.width;
  float uy = ((float)(int)py + 0.5f + jitter.y) / (float)b.height;
  ux = ux * 2.0f - 1.0f;
  uy = uy * 2.0f - 1.0f;
  float aspect = (float)b.width / (float)b.height;
  float tanFov = std::tan(b.fov * 0.017453292519943295f * 0.5f);
  ux *= aspect * tanFov * b.focalLength;
  uy *= tanFov * b.focalLength;
  Ray r;
  r.origin = b.cameraPos;
  r.direction = normalize3(add(add(b.cameraDir, scale_l(ux, b.cameraRight)), scale_l(uy, b.cameraUp)));
  return r;
}

// ------------------------------------------------------------------ shading (glsl:219-507)

// fresnelSchlick (glsl:220-223). pow(x, 2.0) follows C powf (negative bases are fine).
float fresnelSchlick(float cosTheta, float ior) {
  float r0 = std::pow((1.0f - ior) / (1.0f + ior), 2.0f);
  return r0 + (1.0f - r0) * std::pow(1.0f - cosTheta, 5.0f);
}

// computePBR (glsl:226-253): GGX NDF, Schlick-GGX G, Schlick F, kD/F split.
V3 computePBR(const rt_material& mat, V3 N, V3 V, V3 L, V3 H, V3 radiance) {
  float alpha = std::pow(mat.roughness, 2.0f);
  float NDF = alpha * alpha /
              (PI * std::pow(std::pow(fmax_(dot3(N, H), 0.0f), 2.0f) * (alpha * alpha - 1.0f) + 1.0f, 2.0f));
  float k = std::pow(mat.roughness + 1.0f, 2.0f) / 8.0f;
  float G = fmax_(dot3(N, V), 0.0f) / (fmax_(dot3(N, V), 0.0f) * (1.0f - k) + k);
  G *= fmax_(dot3(N, L), 0.0f) / (fmax_(dot3(N, L), 0.0f) * (1.0f - k) + k);
  V3 albedo = ld3(mat.albedo);
  V3 F0 = mix3(mk(0.04f, 0.04f, 0.04f), albedo, mat.metallic);
  float p5 = std::pow(1.0f - fmax_(dot3(H, V), 0.0f), 5.0f);
  V3 F = add(F0, scale(mk(1.0f - F0.x, 1.0f - F0.y, 1.0f - F0.z), p5));
  V3 numerator = scale_l(NDF * G, F);
  float denominator = 4.0f * fmax_(dot3(N, V), 0.0f) * fmax_(dot3(N, L), 0.0f);
  V3 specular = divs(numerator, fmax_(denominator, 0.001f));
  V3 kD = scale(mk(1.0f - F.x, 1.0f - F.y, 1.0f - F.z), 1.0f - mat.metallic);
  V3 diffuse = divs(mul(kD, albedo), PI);
  return scale(mul(add(diffuse, specular), radiance), fmax_(dot3(N, L), 0.0f));
}

// calculateRefraction (glsl:256-270); the energy update (:268) is dead and omitted.
V3 calculateRefraction(const Ray& ray, V3 N, const rt_material& mat) {
  bool entering = dot3(ray.direction, N) < 0.0f;
  float eta = entering ? (1.0f / mat.ior) : mat.ior;
  V3 normal = entering ? N : neg(N);
  V3 refractDir = refract3(normalize3(ray.direction), normal, eta);
  if (dot3(refractDir, refractDir) < 0.001f) refractDir = reflect3(ray.direction, normal);   // TIR
  return refractDir;
}

// random (glsl:273-275)
float random2(float sx, float sy) {
  return fract_(std::sin(sx * 12.9898f + sy * 78.233f) * 43758.5453123f);
}

// haltonSequence (glsl:278-288): radical inverse by repeated FP32 division.
float haltonSequence(int index, int base) {
  float result = 0.0f;
  float f = 1.0f / (float)base;
  int i = index;
  while (i > 0) {
    result += f * (float)(i % base);
    i = i / base;
    f = f / (float)base;
  }
  return result;
}

// cosineWeightedHemisphere (glsl:291-308); degenerate when normal is parallel to (0,1,1).
V3 cosineWeightedHemisphere(V2 rnd, V3 normal) {
  float phi = 2.0f * PI * rnd.x;
  float cosTheta = std::sqrt(rnd.y);
  float sinTheta = std::sqrt(1.0f - rnd.y);
  V3 h = mk(sinTheta * std::cos(phi), cosTheta, sinTheta * std::sin(phi));
  V3 tangent = normalize3(cross3(normal, mk(0.f, 1.f, 1.f)));
  V3 bitangent = cross3(normal, tangent);
  return normalize3(add(add(scale(tangent, h.x), scale(bitangent, h.z)), scale(normal, h.y)));
}

// hammersley (glsl:311-313)
V2 hammersley(int i, int N) { V2 r = {(float)i / (float)N, haltonSequence(i, 2)}; return r; }

// computeSubsurfaceScattering (glsl:316-339)
V3 computeSubsurfaceScattering(const Bindings& b, Counters& k, V3 P, V3 N, const rt_material& mat) {
  V3 sss = mk(0.f, 0.f, 0.f);
  for (int i = 0; i < 4; ++i) {
    k.c[RT_CNT_SSS_SAMPLE]++;
    Ray r;
    r.direction = cosineWeightedHemisphere(hammersley(i, 4), N);
    r.origin = add(P, scale(N, 0.001f));
    const rt_material* tm = nullptr; V3 tn; float t;
    if (intersectObjects(b, k, r, &tm, &tn, &t, nullptr)) {
      float attenuation = std::exp(-t / mat.scatterDistance);
      sss = add(sss, scale(ld3(tm->albedo), attenuation));
    }
  }
  return divs(scale(mul(sss, ld3(mat.subsurfaceColor)), mat.subsurfaceScatter), 4.0f);
}

// pcfShadow (glsl:342-397). filterSize is always shadowSoftness*0.005 (:364 overwrites :352);
// directional lights use the unnormalised jittered direction and no distance test.
float pcfShadow(const Bindings& b, Counters& k, unsigned px, unsigned py, V3 point, V3 normal,
                const rt_light& light, V3 lightDir, float lightDistance) {
  k.c[RT_CNT_PCF_SETUP]++;
  float shadow = 0.0f;
  V3 tangent = normalize3(cross3(lightDir, mk(0.f, 1.f, 0.f)));   // NaN when lightDir is parallel to Y
  V3 bitangent = cross3(lightDir, tangent);
  V2 uv = {(float)(px + (unsigned)b.frameCount) * b.noiseScale.x, (float)(py + (unsigned)b.frameCount) * b.noiseScale.y};
  V2 jitter = {fetch_noise(b, uv), 0.0f};                          // GL_R8: .g = 0
  for (int i = 0; i < light.pcfSamples; ++i) {
    k.c[RT_CNT_PCF_SAMPLE]++;
    float filterSize = light.shadowSoftness * 0.005f;
    float rx = fract_(haltonSequence(i, 2) + jitter.x);
    float ry = fract_(haltonSequence(i, 3) + jitter.y);
    V3 d = add(add(lightDir, scale(scale_l(rx, tangent), filterSize)), scale(scale_l(ry, bitangent), filterSize));
    if (light.type != 1) d = normalize3(d);
    Ray r;
    r.origin = add(point, scale(normal, 0.001f));
    r.direction = d;
    const rt_material* tm = nullptr; V3 tn; float t;
    bool occluded = intersectObjects(b, k, r, &tm, &tn, &t, nullptr);
    if (light.type == 0 || light.type == 2) occluded = occluded && (t < lightDistance);
    shadow += occluded ? 0.0f : 1.0f;
  }
  return shadow / (float)light.pcfSamples;                         // 0/0 = NaN when pcfSamples == 0
}

// pcssShadow (glsl:400-440): 16 fixed probe rays decide "any blocker?"; avgBlockerDepth and
// penumbraSize (:432-436) never reach the result. searchSize is added as a SCALAR to all
// three direction components (:409).
float pcssShadow(const Bindings& b, Counters& k, unsigned px, unsigned py, V3 point, V3 normal,
                 const rt_light& light, V3 lightDir, float lightDistance) {
  int blockerCount = 0;
  float searchSize = light.lightSize * 0.1f;
  for (int i = 0; i < 16; ++i) {
    k.c[RT_CNT_PCSS_PROBE]++;
    float rv = haltonSequence(i, 3) * 2.0f - 1.0f;
    V3 sampleDir = adds(adds(lightDir, rv * searchSize), rv * searchSize);
    Ray r;
    r.origin = add(point, scale(normal, 0.001f));
    r.direction = normalize3(sampleDir);
    const rt_material* tm = nullptr; V3 tn; float t;
    bool occluded = intersectObjects(b, k, r, &tm, &tn, &t, nullptr);
    if (light.type != 1) occluded = occluded && (t < lightDistance);
    if (occluded) blockerCount++;
  }
  if (blockerCount == 0) return 1.0f;
  return pcfShadow(b, k, px, py, point, normal, light, lightDir, lightDistance);
}

// calculateShadow (glsl:442-455)
float calculateShadow(const Bindings& b, Counters& k, unsigned px, unsigned py, V3 point, V3 normal, V3 lightDir,
                      float lightDistance, const rt_light& light) {
  if (light.shadowType == 0) return 1.0f;
  if (light.shadowType == 1) return pcfShadow(b, k, px, py, point, normal, light, lightDir, lightDistance);
  if (light.shadowType == 2) return pcssShadow(b, k, px, py, point, normal, light, lightDir, lightDistance);
  return 0.0f;
}

// computeLighting (glsl:457-507). Shadow rays are cast even when the light contributes 0.
V3 computeLighting(const Bindings& b, Counters& k, unsigned px, unsigned py, V3 P, V3 N, const rt_material& mat, V3 V) {
  V3 Lo = mk(0.f, 0.f, 0.f);
  for (int i = 0; i < b.numLights; ++i) {
    const rt_light& light = b.lights[i];
    V3 lightDir = mk(0.f, 0.f, 0.f);
    float attenuation = 1.0f, lightDistance = 0.0f;
    if (light.type == 0) {                                       // point
      k.c[RT_CNT_LIGHT_POINT]++;
      lightDir = sub(ld3(light.position), P);
      lightDistance = std::sqrt(dot3(lightDir, lightDir));
      attenuation = 1.0f / (1.0f + 0.1f * lightDistance + 0.01f * lightDistance * lightDistance);
      lightDir = normalize3(lightDir);
    } else if (light.type == 1) {                                // directional
      k.c[RT_CNT_LIGHT_DIR]++;
      lightDir = normalize3(neg(ld3(light.direction)));
      lightDistance = 1e6f;
    } else if (light.type == 2) {                                // area: a point light with a cosine
      k.c[RT_CNT_LIGHT_AREA]++;
      lightDir = sub(ld3(light.position), P);
      lightDistance = std::sqrt(dot3(lightDir, lightDir));
      lightDir = normalize3(lightDir);
      attenuation = 1.0f / (lightDistance * lightDistance);
      V3 lightNormal = normalize3(ld3(light.direction));
      float lightCos = fmax_(dot3(lightDir, lightNormal), 0.0f);   // P->light vs facing
      attenuation *= lightCos;
    }
    float shadowFactor = calculateShadow(b, k, px, py, P, N, lightDir, lightDistance, light);
    V3 L = normalize3(lightDir);
    V3 H = normalize3(add(V, L));
    V3 radiance = scale(scale(ld3(light.color), attenuation), light.intensity);
    k.c[RT_CNT_PBR]++;
    Lo = add(Lo, scale(computePBR(mat, N, V, L, H, radiance), shadowFactor));
  }
  if (mat.subsurfaceScatter > 0.0f) Lo = add(Lo, computeSubsurfaceScattering(b, k, P, N, mat));
  return Lo;
}

// ------------------------------------------------------------------ main (glsl:509-584)

struct PixelOut { float color[3]; float P[3]; float N[3]; int objectId; };

void shade_pixel(const Bindings& b, Counters& k, unsigned px, unsigned py, PixelOut* out) {
  V2 nuv = {(float)(px + (unsigned)b.frameCount) * b.noiseScale.x, (float)(py + (unsigned)b.frameCount) * b.noiseScale.y};
  float nr = fetch_noise(b, nuv);
  V2 jitter = {nr * 2.0f - 1.0f, 0.0f * 2.0f - 1.0f};     // .xy of an R8 texel: (r, 0)
  k.c[RT_CNT_CAMERA]++;
  Ray ray = generateCameraRay(b, px, py, jitter);

  V3 finalColor = mk(0.f, 0.f, 0.f), throughput = mk(1.f, 1.f, 1.f);
  V3 P = mk(0.f, 0.f, 0.f), N = mk(0.f, 0.f, 0.f);
  int objectId = -1;

  for (int depth = 0; depth < b.maxRayDepth; ++depth) {
    const rt_material* mp = nullptr;
    float t;
    if (!intersectObjects(b, k, ray, &mp, &N, &t, &objectId)) {
      k.c[RT_CNT_MISS]++;
      if (b.useSkybox) finalColor = add(finalColor, mul(throughput, fetch_sky(b, ray.direction)));
      else finalColor = add(finalColor, mul(throughput, mk(0.f, 0.f, 0.f)));
      break;
    }
    const rt_material& mat = *mp;
    P = add(ray.origin, scale(ray.direction, t));
    V3 V = normalize3(neg(ray.direction));

    V3 Lo = computeLighting(b, k, px, py, P, N, mat, V);
    finalColor = add(finalColor, mul(throughput, Lo));

    if (depth > 2) {                                        // Russian roulette (glsl:544-549)
      k.c[RT_CNT_RR]++;
      V3 albedo = ld3(mat.albedo);
      float diffuseWeight = std::sqrt(dot3(albedo, albedo)) * mat.diffuseStrength;
      float continueProb = fmin_(fmax_(throughput.x, fmax_(throughput.y, throughput.z)) * 0.95f + diffuseWeight, 0.99f);
      if (random2((float)(px + (unsigned)depth), (float)(py + (unsigned)depth)) > continueProb) break;
      throughput = divs(throughput, continueProb);
    }

    float F = fresnelSchlick(fmax_(dot3(V, N), 0.0f), mat.ior);
    V3 albedo = ld3(mat.albedo);
    if (mat.diffuseStrength > 0.0f) {                       // glsl:555-567
      k.c[RT_CNT_BOUNCE_DIFFUSE]++;
      V2 rnd = hammersley(depth * 64 + b.frameCount, 64);
      V3 specularDir = reflect3(ray.direction, N);
      V3 diffuseDir = cosineWeightedHemisphere(rnd, N);
      V3 mixedDir = mix3(specularDir, diffuseDir, mat.roughness);
      ray.direction = normalize3(mixedDir);
      ray.origin = add(P, scale(N, 0.001f));
      throughput = mul(throughput, scale(albedo, mat.diffuseStrength));
    } else if (mat.transparency > 0.0f) {                   // glsl:568-571
      k.c[RT_CNT_BOUNCE_REFRACT]++;
      ray.direction = calculateRefraction(ray, N, mat);
      ray.origin = sub(P, scale(N, 0.001f));
      throughput = mul(throughput, scale(scale(albedo, 1.0f - F), mat.transparency));
    } else {                                                // glsl:572-576
      k.c[RT_CNT_BOUNCE_MIRROR]++;
      ray.direction = reflect3(ray.direction, N);
      ray.origin = add(P, scale(N, 0.001f));
      throughput = mul(throughput, scale(albedo, F));
    }
  }
  out->color[0] = finalColor.x; out->color[1] = finalColor.y; out->color[2] = finalColor.z;
  out->P[0] = P.x; out->P[1] = P.y; out->P[2] = P.z;
  out->N[0] = N.x; out->N[1] = N.y; out->N[2] = N.z;
  out->objectId = objectId;
}

}  // namespace

extern "C" const char* oracle_kind(void) { return "port"; }

static int render_impl(const rt_frame_params* p, const void* objects, const void* lights,
                       const float* sky_rgba32f, int face_size,
                       const unsigned char* noise_r8, int noise_w, int noise_h,
                       int x0, int y0, int sx, int sy, int nx, int ny,
                       float* color, float* position, float* normal, int32_t* object_id,
                       uint64_t* counters, int nthreads, float* margin) {
  if (!p || !color || !position || !normal) return 1;
  if (margin) for (size_t o = 0; o < (size_t)nx * ny; ++o) margin[o] = INFINITY;
  Bindings b;
  b.objects = (const rt_object*)objects; b.numObjects = p->num_objects;
  b.lights = (const rt_light*)lights; b.numLights = p->num_lights;
  b.cameraPos = ld3(p->camera_pos); b.cameraDir = ld3(p->camera_dir);
  b.cameraUp = ld3(p->camera_up); b.cameraRight = ld3(p->camera_right);
  b.fov = p->fov_deg; b.focalLength = p->focal_length; b.maxRayDistance = p->max_ray_distance;
  b.useSkybox = p->use_skybox != 0;
  b.sky = sky_rgba32f; b.skySize = face_size;
  b.noise = noise_r8; b.noiseW = noise_w; b.noiseH = noise_h;
  b.noiseScale.x = p->noise_scale[0]; b.noiseScale.y = p->noise_scale[1];
  b.maxRayDepth = p->max_ray_depth; b.width = p->width; b.height = p->height;
  const int ns = p->num_samples > 0 ? p->num_samples : 1;
  if (nthreads <= 0) nthreads = omp_get_max_threads();
  Counters total; std::memset(&total, 0, sizeof(total));
  for (int s = 0; s < ns; ++s) {
    b.frameCount = p->frame_count + s;
#pragma omp parallel num_threads(nthreads)
    {
      Counters local; std::memset(&local, 0, sizeof(local));
#pragma omp for schedule(dynamic, 1)
      for (int j = 0; j < ny; ++j) {
        for (int i = 0; i < nx; ++i) {
          PixelOut o;
          tl_margin = margin ? margin + (size_t)j * nx + i : nullptr;      // min over the samples of the pixel
          shade_pixel(b, local, (unsigned)(x0 + i * sx), (unsigned)(y0 + j * sy), &o);
          tl_margin = nullptr;
          size_t at = ((size_t)j * nx + i) * 4;
          if (s == 0) {
            for (int c = 0; c < 3; ++c) { color[at + c] = o.color[c]; position[at + c] = o.P[c]; normal[at + c] = o.N[c]; }
            color[at + 3] = 1.0f; position[at + 3] = 1.0f; normal[at + 3] = 1.0f;
            if (object_id) object_id[(size_t)j * nx + i] = o.objectId;
          } else {
            for (int c = 0; c < 3; ++c) color[at + c] = color[at + c] + o.color[c];
          }
        }
      }
#pragma omp critical
      for (int c = 0; c < RT_CNT_COUNT; ++c) total.c[c] += local.c[c];
    }
  }
  if (ns > 1) {
    const float inv = 1.0f / (float)ns;
    for (size_t o = 0; o < (size_t)nx * ny; ++o)
      for (int c = 0; c < 3; ++c) color[o * 4 + c] = color[o * 4 + c] * inv;
  }
  if (counters) for (int c = 0; c < RT_CNT_COUNT; ++c) counters[c] = total.c[c];
  return 0;
}

extern "C" int oracle_render(const rt_frame_params* p, const void* objects, const void* lights,
                             const float* sky_rgba32f, int face_size,
                             const unsigned char* noise_r8, int noise_w, int noise_h,
                             int x0, int y0, int sx, int sy, int nx, int ny,
                             float* color, float* position, float* normal, int32_t* object_id,
                             uint64_t* counters, int nthreads) {
  return render_impl(p, objects, lights, sky_rgba32f, face_size, noise_r8, noise_w, noise_h, x0, y0, sx, sy, nx, ny,
                     color, position, normal, object_id, counters, nthreads, nullptr);
}

// Port only: oracle_render plus, per pixel, the grazing margin of its path (see margin_of_object); INFINITY
// when no decision was close to anything (e.g. an empty scene).
extern "C" int oracle_render_margins(const rt_frame_params* p, const void* objects, const void* lights,
                                     const float* sky_rgba32f, int face_size,
                                     const unsigned char* noise_r8, int noise_w, int noise_h,
                                     int x0, int y0, int sx, int sy, int nx, int ny,
                                     float* color, float* position, float* normal, int32_t* object_id,
                                     uint64_t* counters, int nthreads, float* margin) {
  return render_impl(p, objects, lights, sky_rgba32f, face_size, noise_r8, noise_w, noise_h, x0, y0, sx, sy, nx, ny,
                     color, position, normal, object_id, counters, nthreads, margin);
}
