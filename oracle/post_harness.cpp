// post_harness.cpp — TEST INFRASTRUCTURE (oracle/_ref build). Not part of the product.
//
// Wraps the reference's own fragment shaders (shader/brightness_extractFS.glsl,
// gaussian_blurFs.glsl, bloom_combineFs.glsl, taaFs.glsl, skyboxFs.glsl), mechanically rewritten
// to C++ by make_ref.py into temp files named by the *_INC macros, behind oracle/post_api.h. Each
// shader sits in its own namespace (they share global names). The host side mirrors what
// src/ForwardShadingPipeline.cpp:185-254 and src/TextureLoader.cpp:118-194 bind before each draw:
// texture objects with their filter / wrap state, uniforms, then main() once per fragment with
// TexCoords = (fragment centre) / size from the full-screen quad (shader/outputVs.glsl).
#include <omp.h>
#include "post_shim.h"
#include "post_api.h"

namespace glslpost {
thread_local vec4 gl_FragCoord;
namespace extract {
#include EXTRACT_INC
}
namespace blur {
#include BLUR_INC
}
namespace combine {
#include COMBINE_INC
}
namespace taa {
#include TAA_INC
}
namespace skybox {
#include SKYBOX_INC
}

// (the helpers and entry points stay inside the namespace: at global scope `sampler2D` would be
// ambiguous between this shim's texture object and glsl_shim.h's R8 noise sampler)

static sampler2D tex(const float* p, int w, int h, int ch, Filter f, Wrap wr, bool half_storage = false) {
  sampler2D s;
  s.texels = p; s.width = w; s.height = h; s.channels = ch; s.filter = f; s.wrap = wr; s.half_storage = half_storage;
  return s;
}
static void store_half(float* dst, const vec4& v) {
  dst[0] = round_to_half(v.x); dst[1] = round_to_half(v.y); dst[2] = round_to_half(v.z); dst[3] = round_to_half(v.w);
}
static void store(float* dst, const vec4& v) { dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w; }
static vec2 quad_uv(int x, int y, int w, int h) { return vec2(((float)x + 0.5f) / (float)w, ((float)y + 0.5f) / (float)h); }

extern "C" const char* post_oracle_kind(void) { return "reference"; }

extern "C" int post_bloom_extract(const float* color, int w, int h, float threshold, float* bright) {
  // outputTex: RGBA32F, LINEAR, no wrap mode set (REPEAT)   (src/ForwardShadingPipeline.cpp:57-65)
  extract::hdrTexture = tex(color, w, h, 4, LINEAR, REPEAT);
  extract::threshold = threshold;
#pragma omp parallel for schedule(static)
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      extract::TexCoords = quad_uv(x, y, w, h);
      extract::shader_main();
      store_half(bright + ((size_t)y * w + x) * 4, extract::brightColor);
    }
  return 0;
}

extern "C" int post_gaussian_blur(const float* in, int w, int h, int horizontal, float* out) {
  blur::image = tex(in, w, h, 4, LINEAR, CLAMP_TO_EDGE);      // bloomTextures (:72-79)
  blur::horizontal = horizontal != 0;
#pragma omp parallel for schedule(static)
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      blur::TexCoords = quad_uv(x, y, w, h);
      blur::shader_main();
      store_half(out + ((size_t)y * w + x) * 4, blur::FragColor);
    }
  return 0;
}

extern "C" int post_bloom_combine(const float* color, const float* bloom, int w, int h, float strength, float* display) {
  combine::scene = tex(color, w, h, 4, LINEAR, REPEAT);
  combine::bloomBlur = tex(bloom, w, h, 4, LINEAR, CLAMP_TO_EDGE);
  combine::bloomStrength = strength;
#pragma omp parallel for schedule(static)
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      combine::TexCoords = quad_uv(x, y, w, h);
      combine::shader_main();
      store(display + ((size_t)y * w + x) * 4, combine::FragColor);
    }
  return 0;
}

extern "C" int post_taa(const float* color, const float* history, const float* normal, int w, int h,
                        float blend, float jitter_x, float jitter_y, float* out) {
  taa::uCurrentFrame = tex(color, w, h, 4, LINEAR, REPEAT);
  taa::uHistory = tex(history, w, h, 4, LINEAR, CLAMP_TO_EDGE);      // historyTex (:92-99)
  taa::gNormal = tex(normal, w, h, 4, NEAREST, REPEAT);              // gNormalTex (:121-126)
  taa::uBlendFactor = blend; taa::uJitterX = jitter_x; taa::uJitterY = jitter_y;
#pragma omp parallel for schedule(static)
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      taa::TexCoords = quad_uv(x, y, w, h);
      gl_FragCoord = vec4((float)x + 0.5f, (float)y + 0.5f, 0.5f, 1.0f);
      taa::shader_main();
      store(out + ((size_t)y * w + x) * 4, taa::fragColor);
    }
  return 0;
}

extern "C" int post_equirect_to_cubemap(const float* equirect_rgb, int w, int h, int S, float* faces) {
  skybox::equirectangularMap = tex(equirect_rgb, w, h, 3, LINEAR, CLAMP_TO_EDGE, true);   // GL_RGB16F (:127-131)
  for (int face = 0; face < 6; ++face) {
#pragma omp parallel for schedule(static)
    for (int j = 0; j < S; ++j)
      for (int i = 0; i < S; ++i) {
        // localPos: the cube-surface point under the fragment centre for captureViews[face]
        // (perspective-correct interpolation of aPos over the face at distance 1, 90 degree fov)
        const float s = 2.0f * (((float)i + 0.5f) / (float)S) - 1.0f;
        const float t = 2.0f * (((float)j + 0.5f) / (float)S) - 1.0f;
        vec3 p;
        switch (face) {
          case 0: p = vec3(1.0f, -t, -s); break;
          case 1: p = vec3(-1.0f, -t, s); break;
          case 2: p = vec3(s, 1.0f, t); break;
          case 3: p = vec3(s, -1.0f, -t); break;
          case 4: p = vec3(s, -t, 1.0f); break;
          default: p = vec3(-s, -t, -1.0f); break;
        }
        skybox::localPos = p;
        skybox::shader_main();
        store_half(faces + (((size_t)face * S + j) * S + i) * 4, skybox::FragColor);
      }
  }
  return 0;
}

extern "C" int post_generate_aabbs(void*, int) { return -1; }
extern "C" int post_decode_rgbe(const unsigned char*, size_t, float*) { return -1; }

}  // namespace glslpost
