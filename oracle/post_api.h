// post_api.h — TEST INFRASTRUCTURE. The C API both CPU oracles of the f1-f3 stages export
// (include/rt_post.h is the product's side):
//   oracle/post_oracle.cpp        ("port": this repo's CPU restatement)
//   oracle/_ref/libref_post.so    ("reference": the reference's own fragment-shader sources
//                                  compiled as C++ by make_ref.py; f1 and the RGBE decode have no
//                                  compilable reference source here and exist in the port only)
// Images are float arrays, row 0 = GL y 0, 4 floats per texel unless stated; an RGBA16F image is
// a float array whose values are already rounded to half precision.
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* post_oracle_kind(void);                      // "port" or "reference"

// shader/brightness_extractFS.glsl -> RGBA16F
int post_bloom_extract(const float* color, int w, int h, float threshold, float* bright);
// shader/gaussian_blurFs.glsl, one pass, RGBA16F -> RGBA16F (CLAMP_TO_EDGE, LINEAR)
int post_gaussian_blur(const float* in, int w, int h, int horizontal, float* out);
// shader/bloom_combineFs.glsl -> RGBA32F
int post_bloom_combine(const float* color, const float* bloom, int w, int h, float strength, float* display);
// shader/taaFs.glsl -> RGBA32F. color: LINEAR/REPEAT; history: LINEAR/CLAMP; normal: NEAREST/REPEAT, RGBA16F values
int post_taa(const float* color, const float* history, const float* normal, int w, int h,
             float blend, float jitter_x, float jitter_y, float* out);
// shader/skyboxFs.glsl over the 6 capture views of src/TextureLoader.cpp:150-157.
// equirect: w*h*3 floats (stored as GL_RGB16F: rounded on fetch); faces: 6*S*S*4 floats, RGBA16F values
int post_equirect_to_cubemap(const float* equirect_rgb, int w, int h, int S, float* faces);

// port only (return -1 from the reference build):
int post_generate_aabbs(void* objects, int n);                                   // src/SceneIO.h:75-104
int post_decode_rgbe(const unsigned char* rgbe, size_t texels, float* rgb);      // stb_image stbi__hdr_convert

#ifdef __cplusplus
}
#endif
