// glsl_shim.h — TEST INFRASTRUCTURE (oracle). Not part of the product; nothing under
// opengl_raytracing_b200/ may include, link or call it.
//
// A minimal GLSL-4.30 vocabulary written fresh so that the reference's own shader
// source (shader/raytracingCs.glsl), after the mechanical rewrite in make_ref.py,
// compiles as C++. Everything here is the "GL driver" side of the
// reference: built-in functions, texture filtering, image stores. Where GLSL 4.30
// leaves behaviour implementation-defined, the choice made here IS the spec the CUDA
// kernel is held to, and DESIGN.md lists each choice:
//   * min/max     : IEEE-754 minNum/maxNum (a NaN operand is ignored) — what an NVIDIA
//                   GPU's FMNMX does and what std::fmin/fmax specify.
//   * normalize   : v * (1 / sqrt(dot(v,v))), IEEE division and square root.
//   * pow         : C powf semantics (pow(-0.2, 2.0) = 0.04, not NaN).
//   * default-constructed vectors are zero (pins q2: P = N = 0 on a primary miss).
//   * texture()   : FP32 bilinear; 2D noise = R8, REPEAT, .g=.b=0, .a=1, or the constant
//                   (0,0,0,1) when no noise texture is bound (q7); cubemap = GL face
//                   selection, per-face CLAMP_TO_EDGE, non-seamless (a20).
// All arithmetic is FP32, evaluated left to right as written in the shader; compile
// with -ffp-contract=off so no FMA is formed.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace glsl {

// ---------------------------------------------------------------- vector types
struct uvec2;
struct ivec2 {
  int x, y;
  ivec2() : x(0), y(0) {}
  ivec2(int a, int b) : x(a), y(b) {}
  explicit ivec2(const uvec2& u);
};
struct uvec2 {
  unsigned x, y;
  uvec2() : x(0), y(0) {}
  uvec2(unsigned a, unsigned b) : x(a), y(b) {}
};
inline ivec2::ivec2(const uvec2& u) : x((int)u.x), y((int)u.y) {}
struct uvec3 {
  unsigned x, y, z;
  uvec3() : x(0), y(0), z(0) {}
  uvec3(unsigned a, unsigned b, unsigned c) : x(a), y(b), z(c) {}
  uvec2 xy() const { return uvec2(x, y); }
};

struct vec2 {
  float x, y;
  vec2() : x(0.f), y(0.f) {}
  explicit vec2(float s) : x(s), y(s) {}
  vec2(float a, float b) : x(a), y(b) {}
  vec2(const uvec2& u) : x((float)u.x), y((float)u.y) {}   // GLSL implicit uvec2 -> vec2
  explicit vec2(const ivec2& i) : x((float)i.x), y((float)i.y) {}
  vec2 xy() const { return *this; }
  vec2& operator*=(float s) { x = x * s; y = y * s; return *this; }
};
struct vec3 {
  float x, y, z;
  vec3() : x(0.f), y(0.f), z(0.f) {}
  explicit vec3(float s) : x(s), y(s), z(s) {}
  vec3(float a, float b, float c) : x(a), y(b), z(c) {}
  vec3& operator+=(const vec3& o) { x = x + o.x; y = y + o.y; z = z + o.z; return *this; }
  vec3& operator*=(const vec3& o) { x = x * o.x; y = y * o.y; z = z * o.z; return *this; }
  vec3& operator*=(float s) { x = x * s; y = y * s; z = z * s; return *this; }
  vec3& operator/=(float s) { x = x / s; y = y / s; z = z / s; return *this; }
};
struct vec4 {
  float x, y, z, w;
  vec4() : x(0.f), y(0.f), z(0.f), w(0.f) {}
  vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
  vec4(const vec3& v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
  vec2 xy() const { return vec2(x, y); }
  vec2 rg() const { return vec2(x, y); }
  vec3 rgb() const { return vec3(x, y, z); }
};

// ---------------------------------------------------------------- operators
inline uvec2 operator+(const uvec2& a, int s) { return uvec2(a.x + (unsigned)s, a.y + (unsigned)s); }

inline vec2 operator+(const vec2& a, const vec2& b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator+(const vec2& a, float s) { return vec2(a.x + s, a.y + s); }
inline vec2 operator-(const vec2& a, float s) { return vec2(a.x - s, a.y - s); }
inline vec2 operator*(const vec2& a, const vec2& b) { return vec2(a.x * b.x, a.y * b.y); }
inline vec2 operator*(const vec2& a, float s) { return vec2(a.x * s, a.y * s); }
inline vec2 operator/(const vec2& a, const vec2& b) { return vec2(a.x / b.x, a.y / b.y); }

inline vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(const vec3& a, const vec3& b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator/(const vec3& a, const vec3& b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator+(const vec3& a, float s) { return vec3(a.x + s, a.y + s, a.z + s); }
inline vec3 operator-(const vec3& a, float s) { return vec3(a.x - s, a.y - s, a.z - s); }
inline vec3 operator*(const vec3& a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator/(const vec3& a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator*(float s, const vec3& a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(float s, const vec3& a) { return vec3(s / a.x, s / a.y, s / a.z); }
inline vec3 operator-(float s, const vec3& a) { return vec3(s - a.x, s - a.y, s - a.z); }
inline vec3 operator+(float s, const vec3& a) { return vec3(s + a.x, s + a.y, s + a.z); }
inline vec3 operator-(const vec3& a) { return vec3(-a.x, -a.y, -a.z); }

// ---------------------------------------------------------------- built-ins (GLSL 4.30 §8)
inline float min(float a, float b) { return (a != a) ? b : ((b != b) ? a : (b < a ? b : a)); }
inline float max(float a, float b) { return (a != a) ? b : ((b != b) ? a : (a < b ? b : a)); }
inline vec3 min(const vec3& a, const vec3& b) { return vec3(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
inline vec3 max(const vec3& a, const vec3& b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
inline float abs(float a) { return std::fabs(a); }
inline float sqrt(float a) { return std::sqrt(a); }
inline float sin(float a) { return std::sin(a); }
inline float cos(float a) { return std::cos(a); }
inline float tan(float a) { return std::tan(a); }
inline float exp(float a) { return std::exp(a); }
inline float pow(float a, float b) { return std::pow(a, b); }
inline float floor(float a) { return std::floor(a); }
inline float fract(float a) { return a - std::floor(a); }
inline vec2 fract(const vec2& a) { return vec2(fract(a.x), fract(a.y)); }
inline float radians(float d) { return d * 0.017453292519943295f; }
inline float dot(const vec2& a, const vec2& b) { return a.x * b.x + a.y * b.y; }
inline float dot(const vec3& a, const vec3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 cross(const vec3& a, const vec3& b) {
  return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline float length(const vec3& a) { return std::sqrt(dot(a, a)); }
inline float inversesqrt(float a) { return 1.0f / std::sqrt(a); }
inline vec3 normalize(const vec3& a) { return a * inversesqrt(dot(a, a)); }
inline float mix(float a, float b, float t) { return a * (1.0f - t) + b * t; }
inline vec3 mix(const vec3& a, const vec3& b, float t) { return a * (1.0f - t) + b * t; }
inline vec3 reflect(const vec3& I, const vec3& N) { return I - 2.0f * dot(N, I) * N; }
inline vec3 refract(const vec3& I, const vec3& N, float eta) {
  float d = dot(N, I);
  float k = 1.0f - eta * eta * (1.0f - d * d);
  if (k < 0.0f) return vec3(0.0f);
  return eta * I - (eta * d + std::sqrt(k)) * N;
}

// ---------------------------------------------------------------- images
struct image2D {
  int slot = 0;            // 0 = outputImage, 1 = gPosition, 2 = gNormal (binding numbers, glsl:61-63)
  int width = 0, height = 0;
};
inline ivec2 imageSize_(const image2D& im) { return ivec2(im.width, im.height); }
// The oracle harness renders arbitrary pixel subsets, so a store is redirected per pixel
// through thread-local RGBA32F slots the harness points at the current output texel.
extern thread_local float* g_store[3];
inline void imageStore(const image2D& im, const ivec2& /*coords*/, const vec4& v) {
  float* dst = g_store[im.slot];
  if (dst) { dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w; }
}

// ---------------------------------------------------------------- samplers
struct sampler2D {
  const unsigned char* r8 = nullptr;  // GL_R8, or null => constant (0,0,0,1)  (q7)
  int width = 0, height = 0;
};
struct samplerCube {
  const float* faces = nullptr;       // 6 x size x size x 4 floats (already FP16-rounded), +X,-X,+Y,-Y,+Z,-Z
  int size = 0;
};

inline int wrap_repeat(int i, int n) { int m = i % n; return m < 0 ? m + n : m; }

// GL_TEXTURE_2D, GL_R8, GL_REPEAT, GL_LINEAR, FP32 weights (OpenGL 4.3 §8.14.2).
inline vec4 texture(const sampler2D& s, const vec2& uv) {
  if (!s.r8) return vec4(0.f, 0.f, 0.f, 1.f);
  float u = uv.x * (float)s.width - 0.5f;
  float v = uv.y * (float)s.height - 0.5f;
  if (!(u == u) || std::fabs(u) > 1.0e9f) u = 0.f;   // NaN/huge coordinates: pinned to 0
  if (!(v == v) || std::fabs(v) > 1.0e9f) v = 0.f;
  float fu = std::floor(u), fv = std::floor(v);
  float a = u - fu, b = v - fv;
  int i0 = wrap_repeat((int)fu, s.width), i1 = wrap_repeat((int)fu + 1, s.width);
  int j0 = wrap_repeat((int)fv, s.height), j1 = wrap_repeat((int)fv + 1, s.height);
  const float k = 1.0f / 255.0f;
  float t00 = (float)s.r8[j0 * s.width + i0] * k, t10 = (float)s.r8[j0 * s.width + i1] * k;
  float t01 = (float)s.r8[j1 * s.width + i0] * k, t11 = (float)s.r8[j1 * s.width + i1] * k;
  float top = t00 * (1.0f - a) + t10 * a;
  float bot = t01 * (1.0f - a) + t11 * a;
  float r = top * (1.0f - b) + bot * b;
  return vec4(r, 0.f, 0.f, 1.f);
}

// GL_TEXTURE_CUBE_MAP, GL_LINEAR, CLAMP_TO_EDGE, non-seamless (OpenGL 4.3 §8.13, table 8.19).
// Major-axis ties resolve x over y over z.
inline vec4 texture(const samplerCube& s, const vec3& r) {
  if (!s.faces) return vec4(0.f, 0.f, 0.f, 1.f);
  float ax = std::fabs(r.x), ay = std::fabs(r.y), az = std::fabs(r.z);
  int face; float sc, tc, ma;
  if (ax >= ay && ax >= az) { ma = ax; if (r.x >= 0.f) { face = 0; sc = -r.z; tc = -r.y; } else { face = 1; sc = r.z; tc = -r.y; } }
  else if (ay >= az)        { ma = ay; if (r.y >= 0.f) { face = 2; sc = r.x; tc = r.z; } else { face = 3; sc = r.x; tc = -r.z; } }
  else                      { ma = az; if (r.z >= 0.f) { face = 4; sc = r.x; tc = -r.y; } else { face = 5; sc = -r.x; tc = -r.y; } }
  float su = 0.5f * (sc / ma + 1.0f);
  float sv = 0.5f * (tc / ma + 1.0f);
  float u = su * (float)s.size - 0.5f;
  float v = sv * (float)s.size - 0.5f;
  if (!(u == u)) u = 0.f;                      // NaN direction: pinned to texel 0
  if (!(v == v)) v = 0.f;
  float hi = (float)(s.size - 1);
  float fu = std::floor(u), fv = std::floor(v);
  float a = u - fu, b = v - fv;
  float fu1 = fu + 1.0f, fv1 = fv + 1.0f;
  fu = fu < 0.f ? 0.f : (fu > hi ? hi : fu);  fu1 = fu1 < 0.f ? 0.f : (fu1 > hi ? hi : fu1);
  fv = fv < 0.f ? 0.f : (fv > hi ? hi : fv);  fv1 = fv1 < 0.f ? 0.f : (fv1 > hi ? hi : fv1);
  int i0 = (int)fu, i1 = (int)fu1, j0 = (int)fv, j1 = (int)fv1;
  const float* f = s.faces + (size_t)face * s.size * s.size * 4;
  const float* t00 = f + ((size_t)j0 * s.size + i0) * 4; const float* t10 = f + ((size_t)j0 * s.size + i1) * 4;
  const float* t01 = f + ((size_t)j1 * s.size + i0) * 4; const float* t11 = f + ((size_t)j1 * s.size + i1) * 4;
  float o[3];
  for (int c = 0; c < 3; ++c) {
    float top = t00[c] * (1.0f - a) + t10[c] * a;
    float bot = t01[c] * (1.0f - a) + t11[c] * a;
    o[c] = top * (1.0f - b) + bot * b;
  }
  return vec4(o[0], o[1], o[2], 1.f);
}

extern thread_local uvec3 gl_GlobalInvocationID;

}  // namespace glsl
