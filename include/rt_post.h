/*
 * rt_post.h — C ABI of the stages either side of the ray-tracing pass ("next" rows of the scope
 * table), in the same library (librt_b200.so) and on the same rt_context:
 *
 *   f1  rt_generate_aabbs        GenerateAABBForObject per object   src/SceneIO.h:75-104
 *   f2  rt_equirect_to_cubemap   ConvertHDRToCubemap                src/TextureLoader.cpp:118-194,
 *                                                                   shader/skyboxFs.glsl:6-19
 *       rt_decode_rgbe           the Radiance .hdr decode stb_image does for the reference
 *   f3  the post chain of ForwardShadingPipline::Render             src/ForwardShadingPipeline.cpp:185-254
 *         rt_bloom_extract       shader/brightness_extractFS.glsl   (:192-201)
 *         rt_gaussian_blur       shader/gaussian_blurFs.glsl        (:206-216, one of the 10 passes)
 *         rt_bloom_combine       shader/bloom_combineFs.glsl        (:220-230)
 *         rt_taa_resolve         shader/taaFs.glsl                  (:233-254)
 *         rt_post_chain          all of :185-254 in one call; results identical to the four
 *                                calls above issued in the reference's order
 *
 * Images are device pointers, row 0 = GL y 0, tightly packed:
 *   RGBA32F  float[4] per texel (outputTex, historyTex, the display image)
 *   RGBA16F  IEEE half[4] per texel (gNormalTex, bloomTextures)
 * All arithmetic is FP32, unfused, in the shader's order of operations; conversions to RGBA16F
 * round to nearest even. Where GL leaves the result to the implementation, the choice made is
 * written beside the entry point (and is what oracle/ checks).
 */
#ifndef RT_POST_H_
#define RT_POST_H_

#include "rt_abi.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- f1 ---------------------------------------------------------------------------------- */
/* Fills `bounds` of n Object records in place (host memory, 176-B stride): H2D, one thread per
 * object, D2H of the bounds. Bit-identical to the host GenerateAABBForObject (FP32, glm's order:
 * sphere = centre -/+ radius; plane = rectangle corners, both shifted by +0.01 * normal).
 * Padding words of `bounds` are left untouched. Objects of any other type keep their bounds. */
rt_status rt_generate_aabbs(rt_context* ctx, void* objects, int n);
/* The same on n records already in device memory (no copies). */
rt_status rt_generate_aabbs_device(rt_context* ctx, void* d_objects, int n, void* cuda_stream);

/* ---- f2 ---------------------------------------------------------------------------------- */
/* Radiance RGBE texels (4 bytes each: 8-bit mantissas + shared exponent, already de-RLE'd) ->
 * RGB32F, as stb_image's stbi__hdr_convert does: e == 0 -> 0, else m * 2^(e - 136). */
rt_status rt_decode_rgbe(rt_context* ctx, const unsigned char* d_rgbe, float* d_rgb32f, size_t texels, void* cuda_stream);
/* Equirectangular RGB32F (w x h, row 0 = the file's first row after stb's vertical flip, i.e.
 * v = 0 at the bottom) -> 6 cube faces of face_size^2 RGBA16F texels, order +X,-X,+Y,-Y,+Z,-Z.
 * Per texel: the direction through the texel centre for that face's lookAt view (:150-157),
 * uv = (atan(d.z, d.x) * 0.1591 + 0.5, asin(d.y) * 0.3183 + 0.5), GL_LINEAR fetch with
 * CLAMP_TO_EDGE (:128-131), alpha 1. The bilinear weights are FP32 (GL hardware quantises them:
 * parity with a GL driver is "unpinned" for this fetch, as for the cubemap fetch of the hot path). */
rt_status rt_equirect_to_cubemap(rt_context* ctx, const float* d_equirect_rgb32f, int w, int h,
                                 void* d_faces_rgba16f, int face_size, void* cuda_stream);

/* ---- f3 ---------------------------------------------------------------------------------- */
typedef struct rt_post_params {
  int32_t width, height;
  float   bloom_threshold;   /* extractShader "threshold"        1.0  (:198)                    */
  float   bloom_strength;    /* bloomCombineShader "bloomStrength" 0.5 (:223)                   */
  int32_t blur_passes;       /* 10, alternating, the first one horizontal (:208-216)            */
  int32_t taa_enabled;       /* imguiManager.IsTAAEnabled() (:233)                              */
  float   taa_blend;         /* uBlendFactor (:238)                                             */
  float   jitter_x, jitter_y;/* uJitterX / uJitterY (:242-243): halton(frame % 8, 2|3) * 0.5 / size */
  uint32_t flags;            /* RT_POST_*                                                       */
} rt_post_params;

#define RT_POST_LITERAL_PASSES 0x01u /* rt_post_chain issues the per-pass kernels instead of the fused ones */

/* brightColor = dot(rgb, (0.2126, 0.7152, 0.0722)) > threshold ? (rgb, 1) : (0, 0, 0, 1), RGBA16F.
 * The colour image is sampled at texel centres (GL_LINEAR at a centre = the texel). */
rt_status rt_bloom_extract(rt_context* ctx, const rt_post_params* p, const float* d_color_rgba32f,
                           void* d_bright_rgba16f, void* cuda_stream);
/* One 9-tap pass, weights (0.227027, 0.1945946, 0.1216216, 0.054054, 0.016216), sum in the
 * shader's order (centre, then +i, -i for i = 1..4), CLAMP_TO_EDGE, RGBA16F in and out.
 * Taps land on texel centres (offset = i / size). in and out must not alias. */
rt_status rt_gaussian_blur(rt_context* ctx, const rt_post_params* p, const void* d_in_rgba16f,
                           void* d_out_rgba16f, int horizontal, void* cuda_stream);
/* display = (scene.rgb + bloom.rgb * strength, 1), RGBA32F (the reference draws this into the
 * window's framebuffer; the 8-bit conversion of that framebuffer is not applied here). */
rt_status rt_bloom_combine(rt_context* ctx, const rt_post_params* p, const float* d_color_rgba32f,
                           const void* d_bloom_rgba16f, float* d_display_rgba32f, void* cuda_stream);
/* taaFs.glsl. current = GL_LINEAR fetch of the colour image at the jittered coordinate, wrap
 * REPEAT (outputTex sets no wrap mode), FP32 weights; history at the texel centre; 3x3
 * neighbourhood by texelFetch — a fetch outside the image is undefined in GL and reads zeros
 * here (robust-access behaviour); normals by GL_NEAREST, wrap REPEAT; history clipped to the
 * neighbourhood box; result = mix(history, current, blend) with blend = taa_blend * 0.2 where
 * the two normals differ (dot < 0.9), else 0. history_prev and history_out must not alias. */
rt_status rt_taa_resolve(rt_context* ctx, const rt_post_params* p, const float* d_color_rgba32f,
                         const float* d_history_prev_rgba32f, const void* d_normal_rgba16f,
                         float* d_history_out_rgba32f, void* cuda_stream);
/* extract -> blur_passes x blur -> combine -> (taa_enabled ? taa) on one stream. The bloom
 * ping-pong images live in the context. d_history_* may be NULL when taa_enabled == 0. */
rt_status rt_post_chain(rt_context* ctx, const rt_post_params* p, const float* d_color_rgba32f,
                        const void* d_normal_rgba16f, const float* d_history_prev_rgba32f,
                        float* d_history_out_rgba32f, float* d_display_rgba32f, void* cuda_stream);
/* Kernels launched by the rt_post / f1 / f2 entry points on this context so far. */
uint64_t  rt_post_launch_count(rt_context* ctx);

#ifdef __cplusplus
}
#endif
#endif /* RT_POST_H_ */
