// post_shim.h — TEST INFRASTRUCTURE (oracle). Not part of the product.
//
// The extra GLSL vocabulary the reference's FRAGMENT shaders need on top of glsl_shim.h
// (bloom, TAA, equirect -> cube: shader/*Fs.glsl), so that their own source compiles as C++
// after make_ref.py's mechanical rewrite. This is the "GL driver" side: texture objects with the
// filter / wrap state the reference's host code sets, and the built-ins those shaders call.
// Choices where GL leaves the result to the implementation (DESIGN.md lists them):
//   * GL_LINEAR   : the fractional texel position is snapped to 8 fractional bits before the
//                   blend (GL hardware filters in fixed point); a weight of exactly 0 or 1
//                   returns that texel untouched. A fetch at a texel centre is the texel.
//   * texelFetch  : outside the image -> (0,0,0,0) (robust buffer access).
//   * RGBA16F / RGB16F render targets and uploads round to nearest even.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include "glsl_shim.h"

namespace glslpost {
using namespace glsl;

// ---------------------------------------------------------------- half rounding (RTNE), value-level
inline float round_to_half(float f) {
  uint32_t x;
  std::memcpy(&x, &f, 4);
  const uint32_t sign = x & 0x80000000u;
  uint32_t a = x & 0x7fffffffu;
  if (a >= 0x7f800000u) return f;                       // inf / NaN stay
  if (a >= 0x477ff000u) {                               // >= 65520 rounds to inf
    const uint32_t inf = sign | 0x7f800000u;
    float r; std::memcpy(&r, &inf, 4); return r;
  }
  if (a < 0x38800000u) {                                // below the smallest normal half (2^-14): subnormal grid 2^-24
    const float scaled = std::fabs(f) * 16777216.0f;    // exact
    const float r = std::nearbyint(scaled) * (1.0f / 16777216.0f);   // default rounding mode = nearest even
    return sign ? -r : r;
  }
  const uint32_t lsb = (a >> 13) & 1u;                  // keep 10 mantissa bits
  a += 0xfffu + lsb;
  a &= ~0x1fffu;
  a |= sign;
  float r; std::memcpy(&r, &a, 4); return r;
}

// ---------------------------------------------------------------- texture objects (float texels)
enum Filter { NEAREST = 0, LINEAR = 1 };
enum Wrap { REPEAT = 0, CLAMP_TO_EDGE = 1 };
struct sampler2D {
  const float* texels = nullptr;     // width * height * channels floats, row 0 = t 0
  int width = 0, height = 0, channels = 4;
  Filter filter = LINEAR;
  Wrap wrap = REPEAT;                // the GL default when the host sets none
  bool half_storage = false;         // GL_RGB16F / GL_RGBA16F upload of float data: values round on fetch
};

inline int wrap_index(int i, int n, Wrap w) {
  if (w == CLAMP_TO_EDGE) return i < 0 ? 0 : (i > n - 1 ? n - 1 : i);
  int m = i % n; return m < 0 ? m + n : m;
}
inline vec4 fetch_texel(const sampler2D& s, int i, int j) {
  const float* t = s.texels + ((size_t)j * s.width + i) * s.channels;
  float c[4] = {0.f, 0.f, 0.f, 1.f};
  for (int k = 0; k < s.channels && k < 4; ++k) c[k] = s.half_storage ? round_to_half(t[k]) : t[k];
  return vec4(c[0], c[1], c[2], c[3]);
}
inline float snap_weight(float a) { return std::nearbyint(a * 256.0f) * (1.0f / 256.0f); }
inline float lerp_tap(float t0, float t1, float a) {
  if (a == 0.0f) return t0;
  if (a == 1.0f) return t1;
  return t0 * (1.0f - a) + t1 * a;
}
inline vec4 lerp_tap(const vec4& t0, const vec4& t1, float a) {
  return vec4(lerp_tap(t0.x, t1.x, a), lerp_tap(t0.y, t1.y, a), lerp_tap(t0.z, t1.z, a), lerp_tap(t0.w, t1.w, a));
}
inline vec4 texture(const sampler2D& s, const vec2& uv) {
  if (s.filter == NEAREST) {
    const int i = wrap_index((int)std::floor(uv.x * (float)s.width), s.width, s.wrap);
    const int j = wrap_index((int)std::floor(uv.y * (float)s.height), s.height, s.wrap);
    return fetch_texel(s, i, j);
  }
  const float u = uv.x * (float)s.width - 0.5f, v = uv.y * (float)s.height - 0.5f;
  const float fu = std::floor(u), fv = std::floor(v);
  const float a = snap_weight(u - fu), b = snap_weight(v - fv);
  const int i0 = wrap_index((int)fu, s.width, s.wrap), i1 = wrap_index((int)fu + 1, s.width, s.wrap);
  const int j0 = wrap_index((int)fv, s.height, s.wrap), j1 = wrap_index((int)fv + 1, s.height, s.wrap);
  const vec4 top = lerp_tap(fetch_texel(s, i0, j0), fetch_texel(s, i1, j0), a);
  const vec4 bot = lerp_tap(fetch_texel(s, i0, j1), fetch_texel(s, i1, j1), a);
  return lerp_tap(top, bot, b);
}
inline vec4 texelFetch(const sampler2D& s, const ivec2& p, int /*lod*/) {
  if (p.x < 0 || p.y < 0 || p.x >= s.width || p.y >= s.height) return vec4(0.f, 0.f, 0.f, 0.f);
  return fetch_texel(s, p.x, p.y);
}
inline ivec2 textureSize(const sampler2D& s, int /*lod*/) { return ivec2(s.width, s.height); }

// ---------------------------------------------------------------- extra operators / built-ins
inline vec2 operator/(float s, const ivec2& v) { return vec2(s / (float)v.x, s / (float)v.y); }
inline vec2 operator-(const vec2& a, const vec2& b) { return vec2(a.x - b.x, a.y - b.y); }
inline vec2& operator*=(vec2& a, const vec2& b) { a.x = a.x * b.x; a.y = a.y * b.y; return a; }
inline vec2& operator+=(vec2& a, float s) { a.x = a.x + s; a.y = a.y + s; return a; }
inline ivec2 operator+(const ivec2& a, const ivec2& b) { return ivec2(a.x + b.x, a.y + b.y); }
inline ivec2 to_ivec2(const vec4& v) { return ivec2((int)v.x, (int)v.y); }      // ivec2(gl_FragCoord)
inline float atan(float y, float x) { return std::atan2(y, x); }
inline float asin(float x) { return std::asin(x); }
inline vec3 clamp(const vec3& x, const vec3& lo, const vec3& hi) { return min(max(x, lo), hi); }

extern thread_local vec4 gl_FragCoord;

}  // namespace glslpost
