/*
 * rt_frame.h — C ABI of one whole iteration of ForwardShadingPipline::Render
 * (/root/reference/src/ForwardShadingPipeline.cpp:139-272) without the window and the UI:
 *
 *     ray-tracing dispatch (:153-180)  ->  bloom extract / 10 blurs / combine into the window's
 *     framebuffer (:185-230)  ->  TAA resolve into historyTex[frameCount % 2] (:232-254)
 *
 * In the reference the three images the ray-tracing pass writes (outputTex, gPositionTex,
 * gNormalTex) are GL textures that never leave the GPU: the post chain consumes them in place and
 * only the combined image reaches the screen. rt_frame keeps that data flow: an rt_frame owns the
 * device images of one pipeline instance (what ForwardShadingPipline::Init creates, :51-137),
 * rt_frame_render issues rt_render + rt_post_chain on them back to back on one stream with
 * nothing crossing PCIe, and rt_frame_read_display is the presentation step (the equivalent of
 * the swap / a glReadPixels of the default framebuffer).
 *
 * Same library (librt_b200.so), same rt_context, same rules: plain pointers and sizes, every call
 * returns a status, no CPU fallback. Results are those of rt_render followed by rt_post_chain
 * on the same inputs, bit for bit (tests/test_frame_gpu.py).
 */
#ifndef RT_FRAME_H_
#define RT_FRAME_H_

#include "rt_abi.h"
#include "rt_post.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rt_frame rt_frame;   /* opaque; belongs to one rt_context; NOT thread-safe */

/* Formats of rt_frame_read_display. */
#define RT_DISPLAY_RGBA32F 0  /* the combined image as computed (16 B per pixel)                     */
#define RT_DISPLAY_RGBA8   1  /* what the reference's window holds: its default framebuffer is 8-bit
                                 UNORM (glfw default, no sRGB enable anywhere in src/), so each
                                 channel is clamp(x, 0, 1) * 255 rounded to nearest (ties to even; GL
                                 leaves the rounding to the implementation), NaN -> 0; 4 B per pixel */

/* ForwardShadingPipline::Init (:51-137) for a width x height pipeline: outputTex RGBA32F,
 * gPositionTex RGBA32F, gNormalTex RGBA16F, the display image RGBA32F, historyTex[2] RGBA32F.
 * The history images start as zeros (the reference allocates them with a NULL pointer, i.e.
 * undefined contents). The bloom ping-pong pair lives in the context (rt_post_chain). */
rt_status rt_frame_create(rt_context* ctx, int width, int height, rt_frame** out);
void      rt_frame_destroy(rt_frame* frame);

/* One iteration of Render on `cuda_stream` (asynchronous): rt_render(rp) into the frame's images,
 * then rt_post_chain(pp) on them. rp->width/height and pp->width/height must be the frame's, and
 * rp must cover the whole image (row_begin 0, row_end height, no row bands): the post chain needs
 * every pixel, so a frame is not row-sharded (multi-GPU = independent frames per GPU).
 * History ping-pong as in the reference (:141, :236, :249): with pp->taa_enabled the resolve reads
 * historyTex[1 - rp->frame_count % 2] and writes historyTex[rp->frame_count % 2]; the CALLER
 * advances frame_count (the reference does so only while TAA is enabled, :252). */
rt_status rt_frame_render(rt_frame* frame, const rt_frame_params* rp, const rt_post_params* pp, void* cuda_stream);

/* The device images (any pointer argument may be NULL). d_history = the image written by the last
 * rt_frame_render with TAA enabled (historyTex[frame_count % 2]). Valid until rt_frame_destroy. */
rt_status rt_frame_images(rt_frame* frame, float** d_color_rgba32f, float** d_position_rgba32f, void** d_normal_rgba16f,
                          float** d_display_rgba32f, float** d_history_rgba32f);

/* Presentation: copies the display image into host memory (pinned memory from rt_host_alloc for
 * full PCIe speed) in the given format, ordered after everything queued on `cuda_stream`, and
 * returns when the copy has landed. */
rt_status rt_frame_read_display(rt_frame* frame, void* host_dst, int format, void* cuda_stream);

/* End to end, host buffers in and out: scene upload (ssbo.update() + lightSSBO.update(), every
 * frame as the reference does; objects == NULL && lights == NULL keeps the uploaded scene) +
 * rt_frame_render + rt_frame_read_display on the frame's own stream. */
rt_status rt_frame_render_host(rt_frame* frame, const rt_frame_params* rp, const rt_post_params* pp,
                               const void* objects, int n, const void* lights, int m, void* host_display, int format);

/* Kernels launched by rt_frame entry points themselves (the RGBA8 conversion); the ray-tracing and
 * post-chain launches are counted by rt_launch_count / rt_post_launch_count of the context. */
uint64_t  rt_frame_launch_count(rt_frame* frame);

#ifdef __cplusplus
}
#endif
#endif /* RT_FRAME_H_ */
