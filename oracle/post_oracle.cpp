// post_oracle.cpp — TEST INFRASTRUCTURE ("port"). Not part of the product.
//
// CPU restatement of the stages either side of the ray-tracing pass, one function
// per reference shader / host routine, FP32, unfused, in the source's order of operations
// (compile with -ffp-contract=off). Pinned against oracle/_ref/libref_post.so — the reference's
// own fragment shaders compiled as C++ — by tests/test_post_oracle.py; post_generate_aabbs is
// pinned against the golden decodes of the shipped scenes and post_decode_rgbe restates the
// published algorithm of stb_image (absent from /root/reference: "parity unpinned" for it).
// Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load this library.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <omp.h>
#include "post_api.h"
#include "../include/rt_abi.h"

namespace {

struct V3 { float x, y, z; };
inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 rgb(const float* img, int w, int x, int y) { const float* t = img + ((size_t)y * w + x) * 4; return v3(t[0], t[1], t[2]); }
inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
inline int wrapi(int i, int n) { int m = i % n; return m < 0 ? m + n : m; }
inline float fmin_(float a, float b) { return (a != a) ? b : ((b != b) ? a : (b < a ? b : a)); }   // NaN operand ignored
inline float fmax_(float a, float b) { return (a != a) ? b : ((b != b) ? a : (a < b ? b : a)); }

// float -> nearest-even half -> float
float to_half(float f) {
  uint32_t x; std::memcpy(&x, &f, 4);
  const uint32_t sign = x & 0x80000000u;
  uint32_t a = x & 0x7fffffffu;
  if (a >= 0x7f800000u) return f;
  if (a >= 0x477ff000u) { const uint32_t inf = sign | 0x7f800000u; float r; std::memcpy(&r, &inf, 4); return r; }
  if (a < 0x38800000u) {
    const float r = std::nearbyint(std::fabs(f) * 16777216.0f) * (1.0f / 16777216.0f);
    return sign ? -r : r;
  }
  a += 0xfffu + ((a >> 13) & 1u);
  a &= ~0x1fffu; a |= sign;
  float r; std::memcpy(&r, &a, 4); return r;
}
inline void put_half(float* dst, V3 c) { dst[0] = to_half(c.x); dst[1] = to_half(c.y); dst[2] = to_half(c.z); dst[3] = 1.0f; }

// GL_LINEAR blend with the fractional position snapped to 1/256
inline float snapw(float a) { return std::nearbyint(a * 256.0f) * (1.0f / 256.0f); }
inline float tap(float t0, float t1, float a) { return a == 0.0f ? t0 : (a == 1.0f ? t1 : t0 * (1.0f - a) + t1 * a); }
inline V3 tap(V3 t0, V3 t1, float a) { return v3(tap(t0.x, t1.x, a), tap(t0.y, t1.y, a), tap(t0.z, t1.z, a)); }

const float kW[5] = {0.227027f, 0.1945946f, 0.1216216f, 0.054054f, 0.016216f};     // gaussian_blurFs.glsl:8

}  // namespace

extern "C" const char* post_oracle_kind(void) { return "port"; }

// brightness_extractFS.glsl:10-18
extern "C" int post_bloom_extract(const float* color, int w, int h, float threshold, float* bright) {
#pragma omp parallel for schedule(static)
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      const V3 c = rgb(color, w, x, y);
      const float brightness = (c.x * 0.2126f + c.y * 0.7152f) + c.z * 0.0722f;
      put_half(bright + ((size_t)y * w + x) * 4, brightness > threshold ? c : v3(0.f, 0.f, 0.f));
    }
  return 0;
}

// gaussian_blurFs.glsl:10-27
extern "C" int post_gaussian_blur(const float* in, int w, int h, int horizontal, float* out) {
#pragma omp parallel for schedule(static)
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      const V3 c = rgb(in, w, x, y);
      V3 r = v3(c.x * kW[0], c.y * kW[0], c.z * kW[0]);
      for (int i = 1; i < 5; ++i) {
        const V3 a = horizontal ? rgb(in, w, clampi(x + i, 0, w - 1), y) : rgb(in, w, x, clampi(y + i, 0, h - 1));
        const V3 b = horizontal ? rgb(in, w, clampi(x - i, 0, w - 1), y) : rgb(in, w, x, clampi(y - i, 0, h - 1));
        r = v3(r.x + a.x * kW[i], r.y + a.y * kW[i], r.z + a.z * kW[i]);
        r = v3(r.x + b.x * kW[i], r.y + b.y * kW[i], r.z + b.z * kW[i]);
      }
      put_half(out + ((size_t)y * w + x) * 4, r);
    }
  return 0;
}

// bloom_combineFs.glsl:10-14
extern "C" int post_bloom_combine(const float* color, const float* bloom, int w, int h, float strength, float* display) {
#pragma omp parallel for schedule(static)
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      const V3 s = rgb(color, w, x, y), b = rgb(bloom, w, x, y);
      float* d = display + ((size_t)y * w + x) * 4;
      d[0] = s.x + b.x * strength; d[1] = s.y + b.y * strength; d[2] = s.z + b.z * strength; d[3] = 1.0f;
    }
  return 0;
}

// taaFs.glsl:21-52
extern "C" int post_taa(const float* color, const float* history, const float* normal, int w, int h,
                        float blend_factor, float jx, float jy, float* out) {
  const float W = (float)w, H = (float)h;
#pragma omp parallel for schedule(static)
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      const float tu = ((float)x + 0.5f) / W, tv = ((float)y + 0.5f) / H;
      const float ju = tu + jx, jv = tv + jy;                                            // :23
      const float u = ju * W - 0.5f, v = jv * H - 0.5f;                                  // :24, LINEAR / REPEAT
      const float fu = std::floor(u), fv = std::floor(v);
      const float a = snapw(u - fu), b = snapw(v - fv);
      const int i0 = wrapi((int)fu, w), i1 = wrapi((int)fu + 1, w), j0 = wrapi((int)fv, h), j1 = wrapi((int)fv + 1, h);
      const V3 current = tap(tap(rgb(color, w, i0, j0), rgb(color, w, i1, j0), a), tap(rgb(color, w, i0, j1), rgb(color, w, i1, j1), a), b);
      V3 hist = rgb(history, w, x, y);                                                   // :27 (texel centre)
      V3 lo = current, hi = current;                                                     // :30-37
      for (int dx = -1; dx <= 1; ++dx)
        for (int dy = -1; dy <= 1; ++dy) {
          const int nx = x + dx, ny = y + dy;
          const V3 nb = (nx < 0 || ny < 0 || nx >= w || ny >= h) ? v3(0.f, 0.f, 0.f) : rgb(color, w, nx, ny);
          lo = v3(fmin_(lo.x, nb.x), fmin_(lo.y, nb.y), fmin_(lo.z, nb.z));
          hi = v3(fmax_(hi.x, nb.x), fmax_(hi.y, nb.y), fmax_(hi.z, nb.z));
        }
      float blend = 0.0f;                                                                // :40-45
      const V3 prev = rgb(normal, w, x, y);
      const V3 curr = rgb(normal, w, wrapi((int)std::floor(ju * W), w), wrapi((int)std::floor(jv * H), h));
      if ((prev.x * curr.x + prev.y * curr.y) + prev.z * curr.z < 0.9f) blend = blend_factor * 0.2f;
      const V3 center = v3(0.5f * (hi.x + lo.x), 0.5f * (hi.y + lo.y), 0.5f * (hi.z + lo.z));   // clipAABB :13-19
      const V3 ext = v3(0.5f * (hi.x - lo.x), 0.5f * (hi.y - lo.y), 0.5f * (hi.z - lo.z));
      V3 clip = v3(hist.x - center.x, hist.y - center.y, hist.z - center.z);
      clip = v3(fmin_(fmax_(clip.x, -ext.x), ext.x), fmin_(fmax_(clip.y, -ext.y), ext.y), fmin_(fmax_(clip.z, -ext.z), ext.z));
      hist = v3(center.x + clip.x, center.y + clip.y, center.z + clip.z);
      const float ia = 1.0f - blend;                                                     // :51 mix
      float* d = out + ((size_t)y * w + x) * 4;
      d[0] = hist.x * ia + current.x * blend; d[1] = hist.y * ia + current.y * blend; d[2] = hist.z * ia + current.z * blend; d[3] = 1.0f;
    }
  return 0;
}

// skyboxFs.glsl:6-19 over the capture views of src/TextureLoader.cpp:150-157
extern "C" int post_equirect_to_cubemap(const float* src, int w, int h, int S, float* faces) {
  auto texel = [&](int x, int y) { const float* t = src + ((size_t)y * w + x) * 3; return v3(to_half(t[0]), to_half(t[1]), to_half(t[2])); };
  for (int face = 0; face < 6; ++face) {
#pragma omp parallel for schedule(static)
    for (int j = 0; j < S; ++j)
      for (int i = 0; i < S; ++i) {
        const float s = 2.0f * (((float)i + 0.5f) / (float)S) - 1.0f;
        const float t = 2.0f * (((float)j + 0.5f) / (float)S) - 1.0f;
        V3 p;
        switch (face) {
          case 0: p = v3(1.0f, -t, -s); break;
          case 1: p = v3(-1.0f, -t, s); break;
          case 2: p = v3(s, 1.0f, t); break;
          case 3: p = v3(s, -1.0f, -t); break;
          case 4: p = v3(s, -t, 1.0f); break;
          default: p = v3(-s, -t, -1.0f); break;
        }
        const float inv = 1.0f / std::sqrt((p.x * p.x + p.y * p.y) + p.z * p.z);
        const V3 d = v3(p.x * inv, p.y * inv, p.z * inv);
        const float uu = std::atan2(d.z, d.x) * 0.1591f + 0.5f, vv = std::asin(d.y) * 0.3183f + 0.5f;
        const float u = uu * (float)w - 0.5f, v = vv * (float)h - 0.5f;
        const float fu = std::floor(u), fv = std::floor(v);
        const float a = snapw(u - fu), b = snapw(v - fv);
        const int i0 = clampi((int)fu, 0, w - 1), i1 = clampi((int)fu + 1, 0, w - 1);
        const int j0 = clampi((int)fv, 0, h - 1), j1 = clampi((int)fv + 1, 0, h - 1);
        put_half(faces + (((size_t)face * S + j) * S + i) * 4, tap(tap(texel(i0, j0), texel(i1, j0), a), tap(texel(i0, j1), texel(i1, j1), a), b));
      }
  }
  return 0;
}

// src/SceneIO.h:75-104 (glm FP32)
extern "C" int post_generate_aabbs(void* objects, int n) {
  rt_object* o = static_cast<rt_object*>(objects);
  for (int k = 0; k < n; ++k) {
    rt_object& ob = o[k];
    const V3 pos = v3(ob.position[0], ob.position[1], ob.position[2]);
    V3 lo, hi;
    if (ob.type == 0) {
      const float r = ob.radius;
      lo = v3(pos.x - r, pos.y - r, pos.z - r); hi = v3(pos.x + r, pos.y + r, pos.z + r);
    } else if (ob.type == 1) {
      const V3 nrm = v3(ob.normal[0], ob.normal[1], ob.normal[2]);
      auto cross = [](V3 a, V3 b) { return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); };
      auto norm = [](V3 a) { const float inv = 1.0f / std::sqrt((a.x * a.x + a.y * a.y) + a.z * a.z); return v3(a.x * inv, a.y * inv, a.z * inv); };
      V3 right, forward;
      if (std::fabs((double)nrm.y) > 0.9) { right = v3(1.f, 0.f, 0.f); forward = v3(0.f, 0.f, 1.f); }
      else { right = norm(cross(nrm, v3(0.f, 1.f, 0.f))); forward = norm(cross(right, nrm)); }
      const float hx = ob.size[0] / 2.0f, hy = ob.size[1] / 2.0f;
      const V3 ax = v3(right.x * hx, right.y * hx, right.z * hx), ay = v3(forward.x * hy, forward.y * hy, forward.z * hy);
      const V3 sh = v3(nrm.x * 0.01f, nrm.y * 0.01f, nrm.z * 0.01f);
      lo = v3(((pos.x - ax.x) - ay.x) + sh.x, ((pos.y - ax.y) - ay.y) + sh.y, ((pos.z - ax.z) - ay.z) + sh.z);
      hi = v3(((pos.x + ax.x) + ay.x) + sh.x, ((pos.y + ax.y) + ay.y) + sh.y, ((pos.z + ax.z) + ay.z) + sh.z);
    } else {
      continue;
    }
    ob.bounds.min[0] = lo.x; ob.bounds.min[1] = lo.y; ob.bounds.min[2] = lo.z;
    ob.bounds.max[0] = hi.x; ob.bounds.max[1] = hi.y; ob.bounds.max[2] = hi.z;
  }
  return 0;
}

// stb_image.h (v2.x) stbi__hdr_convert: rgb = mantissa * 2^(e - 136), e == 0 -> 0
extern "C" int post_decode_rgbe(const unsigned char* rgbe, size_t texels, float* out) {
  for (size_t i = 0; i < texels; ++i) {
    const unsigned char* q = rgbe + 4 * i;
    if (q[3] != 0) {
      const float f = std::ldexp(1.0f, (int)q[3] - 136);
      out[3 * i] = (float)q[0] * f; out[3 * i + 1] = (float)q[1] * f; out[3 * i + 2] = (float)q[2] * f;
    } else {
      out[3 * i] = out[3 * i + 1] = out[3 * i + 2] = 0.0f;
    }
  }
  return 0;
}
