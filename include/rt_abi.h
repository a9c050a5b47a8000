/*
 * rt_abi.h — C ABI of librt_b200.so: the sm_100a replacement for the reference's
 * per-pixel ray-tracing compute pass (shader/raytracingCs.glsl).
 *
 * The reference has no FFI; its "operator surface" for this pass is the GL binding
 * state set up at ONE call site, src/ForwardShadingPipeline.cpp:155-180:
 *   - Object[]  SSBO binding 0      (src/SSBO.h:16-23,      stride 176 B)
 *   - Light[]   SSBO binding 1      (src/LightSSBO.h:16-25, stride  96 B)
 *   - 13 uniforms                   (src/ForwardShadingPipeline.cpp:156-166,
 *                                    shader/raytracingCs.glsl:72-89)
 *   - 3 output images               (src/ForwardShadingPipeline.cpp:57-65,115-126)
 *   - cubemap + noise samplers      (src/ForwardShadingPipeline.cpp:39-49,167-170)
 *   - glDispatchCompute + barrier   (src/ForwardShadingPipeline.cpp:175-180)
 * Every entry point below names the reference interface it replaces.
 *
 * Conventions: plain C, no torch / CUDA types in signatures; every call returns an
 * rt_status and never throws or exits; the library copies what it is given and never
 * frees caller memory; a context is bound to one CUDA device and one caller thread at
 * a time. There is NO CPU fallback: without a CUDA device rt_create fails.
 */
#ifndef RT_ABI_H_
#define RT_ABI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 3

/* ------------------------------------------------------------------------- */
/* Data contracts: byte-exact std430 layouts (== the alignas() C++ structs).   */
/* ------------------------------------------------------------------------- */

/* src/Material.h:11-23  <->  shader/raytracingCs.glsl:20-32   (80 B, align 16) */
typedef struct rt_material {
  int32_t type;              /*  0  never read by the shader                    */
  int32_t _pad0[3];          /*  4                                             */
  float   albedo[3];         /* 16                                             */
  float   metallic;          /* 28                                             */
  float   roughness;         /* 32                                             */
  float   diffuseStrength;   /* 36  >0 selects the diffuse-bounce branch        */
  float   ior;               /* 40                                             */
  float   transparency;      /* 44                                             */
  float   specular;          /* 48  never read by the shader                    */
  float   subsurfaceScatter; /* 52                                             */
  int32_t _pad1[2];          /* 56                                             */
  float   subsurfaceColor[3];/* 64                                             */
  float   scatterDistance;   /* 76                                             */
} rt_material;

/* src/Object.h:8-11  <->  shader/raytracingCs.glsl:8-11       (32 B, align 16) */
typedef struct rt_aabb {
  float   min[3]; int32_t _pad0;   /*  0 */
  float   max[3]; int32_t _pad1;   /* 16 */
} rt_aabb;

/* src/Object.h:13-21  <->  shader/raytracingCs.glsl:34-42    (176 B, align 16) */
typedef struct rt_object {
  int32_t type;              /*   0  0 = sphere, 1 = plane                      */
  int32_t _pad0[3];          /*   4                                            */
  float   position[3];       /*  16                                            */
  float   radius;            /*  28                                            */
  float   normal[3];         /*  32                                            */
  int32_t _pad1;             /*  44                                            */
  float   size[2];           /*  48                                            */
  int32_t _pad2[2];          /*  56                                            */
  rt_material material;      /*  64                                            */
  rt_aabb bounds;            /* 144                                            */
} rt_object;

/* src/Light.h:7-20  <->  shader/raytracingCs.glsl:44-58       (96 B, align 16) */
typedef struct rt_light {
  int32_t type;              /*  0  0 = point, 1 = directional, 2 = area        */
  int32_t _pad0[3];          /*  4                                             */
  float   position[3];       /* 16                                             */
  int32_t _pad1;             /* 28                                             */
  float   direction[3];      /* 32                                             */
  int32_t _pad2;             /* 44                                             */
  float   color[3];          /* 48                                             */
  float   intensity;         /* 60                                             */
  float   radius;            /* 64  never read by the shader                    */
  int32_t samples;           /* 68  never read by the shader                    */
  float   shadowSoftness;    /* 72                                             */
  int32_t shadowType;        /* 76  0 none, 1 PCF, 2 PCSS, other -> shadow 0    */
  int32_t pcfSamples;        /* 80                                             */
  float   lightSize;         /* 84                                             */
  float   angularRadius;     /* 88  dead in the shader (:352 overwritten :364)  */
  int32_t _pad3;             /* 92                                             */
} rt_light;

#define RT_OBJECT_STRIDE   176
#define RT_LIGHT_STRIDE     96
#define RT_MATERIAL_STRIDE  80

#if defined(__cplusplus)
static_assert(sizeof(rt_material) == 80 && sizeof(rt_aabb) == 32, "std430 layout");
static_assert(sizeof(rt_object) == 176 && sizeof(rt_light) == 96, "std430 layout");
static_assert(offsetof(rt_object, material) == 64 && offsetof(rt_object, bounds) == 144, "std430 layout");
static_assert(offsetof(rt_material, subsurfaceColor) == 64 && offsetof(rt_light, shadowSoftness) == 72, "std430 layout");
#else
_Static_assert(sizeof(rt_material) == 80 && sizeof(rt_aabb) == 32, "std430 layout");
_Static_assert(sizeof(rt_object) == 176 && sizeof(rt_light) == 96, "std430 layout");
#endif

/* ------------------------------------------------------------------------- */
/* Status codes (the reference propagates no errors: src/Shader.h:132-149)     */
/* ------------------------------------------------------------------------- */
typedef enum rt_status {
  RT_OK = 0,
  RT_ERR_INVALID_ARG = 1,
  RT_ERR_CUDA = 2,
  RT_ERR_NCCL = 3,
  RT_ERR_OOM = 4,
  RT_ERR_UNAVAILABLE = 5
} rt_status;

/* rt_frame_params.flags */
#define RT_FLAG_NORMAL_FP32     0x01u /* normal buffer is RGBA32F instead of the reference's RGBA16F   */
#define RT_FLAG_PARITY_MATH     0x02u /* IEEE div/sqrt everywhere, no FMA contraction (-fmad=false)    */
#define RT_FLAG_SKYBOX_HW       0x04u /* miss path through the CUDA cubemap texture object (HW filter) */
#define RT_FLAG_COUNTERS        0x08u /* accumulate the reference-literal op counters (rt_get_counters)*/
#define RT_FLAG_ACCUMULATE      0x10u /* colour += instead of = (sample-range partial sums) */
#define RT_FLAG_SUM_SAMPLES     0x40u /* colour = plain sum over num_samples (no 1/n): one partial per rank */
#define RT_FLAG_LITERAL_KERNEL  0x20u /* one-ray-per-lane literal kernel instead of the packet kernel   */

/* == the shader's uniforms (shader/raytracingCs.glsl:72-89) + the compile-time constants
 *    promoted to run time (:4 MAX_RAY_DEPTH, image size) + the sharding window.          */
typedef struct rt_frame_params {
  int32_t  width, height;        /* imageSize(outputImage)                                    */
  int32_t  row_begin, row_end;   /* rows to render; full frame = 0,height. Row 0 = gl y 0      */
  int32_t  num_objects;          /* numObjects  (src/ForwardShadingPipeline.cpp:156)          */
  int32_t  num_lights;           /* numLights   (:157)                                        */
  float    camera_pos[3];        /* cameraPos   (:158)                                        */
  float    camera_dir[3];        /* cameraDir   (:159)                                        */
  float    camera_up[3];         /* cameraUp    (:160)                                        */
  float    camera_right[3];      /* cameraRight (:161)                                        */
  float    fov_deg;              /* fov         (:162)  shader default 60, Camera::FOV 45     */
  float    focal_length;         /* focalLength — never set by the host: 1.0 (glsl:80)        */
  float    max_ray_distance;     /* maxRayDistance — never set by the host: 114514 (glsl:85)  */
  int32_t  use_skybox;           /* useSkybox   (:166)                                        */
  float    noise_scale[2];       /* noiseScale  (:164)  1/1024                                */
  int32_t  frame_count;          /* frameCount  (:163)  first sample index                    */
  int32_t  num_samples;          /* 1 = the reference; n = mean over frame_count..+n-1        */
  int32_t  max_ray_depth;        /* #define MAX_RAY_DEPTH 3 (glsl:4)                          */
  uint32_t flags;                /* RT_FLAG_*                                                 */
  /* Interleaved row bands (multi-GPU load balance): row_band = 0 renders the whole
   * window [row_begin,row_end); otherwise only the rows r of the window with
   * ((r - row_begin) / row_band) % row_period == row_phase. row_band must be a multiple of 4. */
  int32_t  row_band, row_period, row_phase;
} rt_frame_params;

/* Reference-literal operation counters (flop model v0). Index into
 * the uint64 array returned by rt_get_counters.                                  */
enum {
  RT_CNT_RAYS = 0,          /* intersectObjects calls (glsl:155)                   */
  RT_CNT_AABB,              /* intersectAABB calls    (glsl:163)                   */
  RT_CNT_SPHERE,            /* intersectSphere calls  (glsl:171)                   */
  RT_CNT_PLANE_EARLY,       /* intersectPlane returning before the frame (glsl:122-124) */
  RT_CNT_PLANE_FULL,        /* intersectPlane reaching the frame build (glsl:126-150)   */
  RT_CNT_HIT_UPDATE_SPHERE, /* closest-hit updates on spheres (glsl:178-188)       */
  RT_CNT_HIT_UPDATE_PLANE,  /* closest-hit updates on planes  (glsl:178-190)       */
  RT_CNT_LIGHT_POINT,       /* computeLighting iterations by light type (glsl:468-489) */
  RT_CNT_LIGHT_DIR,
  RT_CNT_LIGHT_AREA,
  RT_CNT_PBR,               /* computePBR calls (glsl:499)                         */
  RT_CNT_PCF_SETUP,         /* pcfShadow calls (glsl:342)                          */
  RT_CNT_PCF_SAMPLE,        /* pcfShadow loop iterations (glsl:362)                */
  RT_CNT_PCSS_PROBE,        /* pcssShadow loop iterations (glsl:406)               */
  RT_CNT_SSS_SAMPLE,        /* computeSubsurfaceScattering iterations (glsl:318)   */
  RT_CNT_BOUNCE_MIRROR,     /* next-ray branch taken (glsl:572-576)                */
  RT_CNT_BOUNCE_REFRACT,    /* (glsl:568-571)                                      */
  RT_CNT_BOUNCE_DIFFUSE,    /* (glsl:555-567)                                      */
  RT_CNT_RR,                /* Russian-roulette evaluations (glsl:544-549)         */
  RT_CNT_CAMERA,            /* generateCameraRay calls (glsl:198)                  */
  RT_CNT_MISS,              /* miss path taken (glsl:530-534)                      */
  RT_CNT_COUNT
};

typedef struct rt_context rt_context;   /* opaque; one per CUDA device; NOT thread-safe */

/* Library / device info. */
int         rt_abi_version(void);
const char* rt_status_string(rt_status s);

/* Replaces ForwardShadingPipline::Init's GL setup for this pass
 * (src/ForwardShadingPipeline.cpp:6-20: shader compile, SSBO gen, texture gen).  */
rt_status rt_create(rt_context** out, int cuda_device);
void      rt_destroy(rt_context* ctx);
const char* rt_last_error(rt_context* ctx);

/* Replaces SSBO::update (src/SSBO.h:16-23) + LightSSBO::update (src/LightSSBO.h:16-25):
 * copies n Object (176-B stride) and m Light (96-B stride) records host->device and
 * re-stages the kernel's SoA scene. The caller keeps ownership. n, m may be 0.
 * async != 0: the copy is enqueued on `cuda_stream` from an internal pinned staging
 * buffer and the call returns without synchronising.                               */
rt_status rt_upload_scene(rt_context* ctx, const void* objects, int n,
                          const void* lights, int m, void* cuda_stream, int async);

/* Replaces glBindTexture(GL_TEXTURE_CUBE_MAP, ...) (src/ForwardShadingPipeline.cpp:167-170)
 * and the finished faces ConvertHDRToCubemap produces (src/TextureLoader.cpp:118-194).
 * faces: 6 x face_size x face_size x channels, order +X,-X,+Y,-Y,+Z,-Z, row 0 = t 0;
 * channels 3 or 4; is_fp16 selects half vs float texels (stored as RGBA16F, like the
 * reference's GL_RGB16F faces). NULL clears the skybox.                              */
rt_status rt_set_skybox(rt_context* ctx, const void* faces, int face_size,
                        int channels, int is_fp16);

/* Replaces InitBlueNoiseTex (src/ForwardShadingPipeline.cpp:39-49): GL_R8, REPEAT,
 * bilinear. NULL => every fetch returns (0,0,0,1) ("as shipped": the texture is
 * mip-incomplete and its sampler unit is never assigned).               */
rt_status rt_set_noise(rt_context* ctx, const unsigned char* r8, int w, int h);

/* Replaces glDispatchCompute + glMemoryBarrier (src/ForwardShadingPipeline.cpp:175-180).
 * Outputs are DEVICE pointers to full-frame (width x height) row-major buffers; only
 * rows [row_begin,row_end) are written. d_color / d_position: RGBA32F (outputImage /
 * gPosition, glsl:61-62). d_normal: RGBA16F (gNormal, glsl:63) or RGBA32F with
 * RT_FLAG_NORMAL_FP32. d_object_id: optional int32 per pixel (last-hit object index,
 * -1 on a primary miss), may be NULL. Any of the three image pointers may be NULL to
 * skip that store. The pointers may be peer-mapped (NVLink) addresses.
 * Asynchronous on `cuda_stream` (a cudaStream_t, NULL = the legacy default stream). Launches of
 * one context may overlap on different streams (each takes its own work counter); the CALLS
 * themselves must come from one thread at a time.                                            */
rt_status rt_render(rt_context* ctx, const rt_frame_params* params,
                    float* d_color_rgba32f, float* d_position_rgba32f,
                    void* d_normal, int32_t* d_object_id, void* cuda_stream);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Same arguments, same images as rt_render
 * (bit for bit), but the call allocates NOTHING: it runs the wavefront pipeline with its ray queues in a file-scope
 * `static __device__` array of RT_STATIC_SCRATCH_BYTES (64 MiB) that every GPU loading the library holds until the process
 * exits, cutting the frame into chunks of tile rows that fit (so large frames take more, smaller launches than rt_render).
 * Limits, rejected with RT_ERR_INVALID_ARG and a message in rt_last_error: num_samples must be 1; no COUNTERS /
 * LITERAL_KERNEL flags; the scene must be one the wavefront pipeline takes (staged in shared memory, <= 32 lights,
 * depth <= 16); one tile row of queues must fit: (ceil(width / 8) * 32) * (160 + 36 * max(lights, 1)) bytes <= 64 MiB - 4 608
 * (width <= 62592 at 3 lights, 12784 at 32). The array is SHARED by every call on the device and keeps the previous call's
 * contents (the call clears the queue counters itself, nothing else needs to start at zero): two calls that overlap on the
 * GPU — different streams, different contexts of one device, different threads — race; order them on one stream. */
#define RT_STATIC_SCRATCH_BYTES (64u << 20)
rt_status rt_render_static(rt_context* ctx, const rt_frame_params* params,
                           float* d_color_rgba32f, float* d_position_rgba32f,
                           void* d_normal, int32_t* d_object_id, void* cuda_stream);

/* One whole reference frame through HOST buffers, as ForwardShadingPipline::Render does
 * it (SSBO upload every frame, dispatch, read-back): scene H2D + render + D2H of the
 * rendered rows into pinned or pageable host buffers. objects/lights may be NULL to
 * reuse the last upload. Synchronous.                                              */
rt_status rt_render_host(rt_context* ctx, const rt_frame_params* params,
                         const void* objects, int n, const void* lights, int m,
                         float* h_color_rgba32f, float* h_position_rgba32f,
                         void* h_normal, int32_t* h_object_id);

/* How rt_render_host moves the three images to the host (a tuning knob, not semantics):
 * AUTO = ZEROCOPY when every host buffer is pinned and device-mapped (rt_host_alloc /
 * cudaHostRegister: the kernel stores straight into host memory over PCIe), else BANDED. */
enum { RT_HOST_PATH_AUTO = 0, RT_HOST_PATH_STAGED = 1, RT_HOST_PATH_BANDED = 2, RT_HOST_PATH_ZEROCOPY = 3 };
rt_status rt_set_host_path(rt_context* ctx, int path);

/* Replaces the profiler's glFinish (src/PerformanceProfiler.cpp:49-65). */
rt_status rt_synchronize(rt_context* ctx);

/* CUDA-GL interop variant writing into the app's GL textures. No GL stack exists on
 * the build or GPU machines: always RT_ERR_UNAVAILABLE here.                       */
rt_status rt_render_gl(rt_context* ctx, const rt_frame_params* params,
                       unsigned gl_tex_color, unsigned gl_tex_position, unsigned gl_tex_normal);

/* Counters accumulated by renders that carried RT_FLAG_COUNTERS since the last reset.
 * out must hold RT_CNT_COUNT uint64 values. Synchronises the context's device.     */
rt_status rt_get_counters(rt_context* ctx, uint64_t* out, int reset);

/* Device memory shared between the per-GPU processes of one box (cudaMalloc + CUDA IPC): the
 * root rank allocates the frame, exports a 64-byte handle, every other rank opens it and passes
 * the mapped address to rt_render — its stores then travel over NVLink into the root's frame
 * (the "fused gather" of SURVEY 8e). rt_ipc_open returns a pointer valid in the caller's process. */
rt_status rt_device_alloc(rt_context* ctx, void** out, size_t bytes);
rt_status rt_device_free(rt_context* ctx, void* p);
rt_status rt_ipc_export(rt_context* ctx, void* device_ptr, unsigned char handle[64]);
rt_status rt_ipc_open(rt_context* ctx, const unsigned char handle[64], void** out);
rt_status rt_ipc_close(rt_context* ctx, void* p);

/* Pinned host memory helpers for the e2e path (cudaHostAlloc / cudaFreeHost). */
rt_status rt_host_alloc(void** out, size_t bytes);
rt_status rt_host_free(void* p);

/* Measurement helper: runs an FMA-only micro-kernel and returns the measured FP32
 * peak in TFLOP/s (the roofline denominator MEASURED_PEAKS.json lacks) and the SM
 * count. Not on the render path.                                                  */
rt_status rt_measure_fp32_peak(rt_context* ctx, double* tflops, int* sm_count, double* ms);

/* Device kernel launch counter (every kernel this library launched on ctx). */
uint64_t  rt_launch_count(rt_context* ctx);

/* ------------------------------------------------------------------------- */
/* Kernel variants and their tuning.                                           */
/* The pass exists in several builds that produce IDENTICAL images (same per-pixel arithmetic,
 * different scheduling): the persistent megakernel (rt_packet.cuh) at two register budgets, with
 * independent warps or CTA-coherent phases, the same megakernel with the rays of a CTA regrouped
 * between bounces and shadow jobs through shared-memory queues (rt_regroup.cuh), and the wavefront
 * pipeline (rt_wavefront.cuh: bounce / shadow-job launches over compacted ray queues in HBM).
 * Which is fastest depends on the
 * scene. The reference has no counterpart (one GLSL program, src/ForwardShadingPipeline.cpp:153);
 * a caller that edits the scene every frame (src/ImGUIManager.cpp:197-202) uses rt_tune once per
 * scene shape, or rt_set_tuning to pin a variant, and never renders a trial frame.            */
/* ------------------------------------------------------------------------- */
enum { RT_KERNEL_AUTO = -1,
       RT_KERNEL_MEGAKERNEL = 0,   /* rt_packet.cuh: persistent, one pixel per lane for the whole path                  */
       RT_KERNEL_WAVEFRONT  = 1,   /* rt_wavefront.cuh: bounce / shadow-job launches over compacted queues in HBM       */
       RT_KERNEL_LITERAL    = 2,   /* rt_kernel.cuh: the reference-literal loops (counters, A/B baseline); not tunable   */
       RT_KERNEL_REGROUP    = 3 }; /* rt_regroup.cuh: persistent, rays regrouped per CTA through shared-memory queues   */
enum { RT_TUNING_DEFAULT = 0,    /* nothing measured for this shape: the built-in default variant  */
       RT_TUNING_TRIALS  = 1,    /* autotune is timing variants on the caller's launches            */
       RT_TUNING_CHOSEN  = 2,    /* measured (rt_tune or autotune); the fastest variant is in use   */
       RT_TUNING_PINNED  = 3 };  /* rt_set_tuning                                                   */
typedef struct rt_tuning {
  int32_t kernel;          /* RT_KERNEL_*                                                            */
  int32_t registers;       /* megakernel, large scenes: 0 = 128 registers (2 CTAs/SM), 1 = 80 (3)    */
  int32_t phase_level;     /* megakernel: 0 independent warps, 1..3 CTA-coherent phases              */
  int32_t state;           /* RT_TUNING_* (output only)                                              */
  int32_t num_variants;    /* how many variants apply to this scene / launch shape (output only)     */
  float   variant_ms[8];   /* best measured time of each, 0 = not measured (output only)            */
  int32_t variant_kernel[8], variant_registers[8], variant_phase_level[8];
} rt_tuning;

/* Times every applicable variant for `params` on the uploaded scene (2 timed launches each, every one enqueued
 * behind an untimed launch so that only device time is measured; internal scratch images; synchronous) and keeps
 * the fastest for launches of that shape. */
rt_status rt_tune(rt_context* ctx, const rt_frame_params* params, rt_tuning* out /* nullable */);
/* What rt_render would run for `params` now, and what is known about the variants. params == NULL:
 * the variant the LAST launch of this context ran (kernel = RT_KERNEL_LITERAL when the scene or the
 * flags sent it to the literal kernel).                                                          */
rt_status rt_get_tuning(rt_context* ctx, const rt_frame_params* params, rt_tuning* out);
/* Pins kernel / registers / phase_level for every later launch (fields < 0 = leave to the library);
 * NULL removes the pin.                                                                          */
rt_status rt_set_tuning(rt_context* ctx, const rt_tuning* tuning);
/* on = 0: rt_render never times variants on the caller's launches (untuned shapes run the default
 * variant until rt_tune is called). Default: on.                                                */
rt_status rt_set_autotune(rt_context* ctx, int on);
/* Sizes every later render grid for (SM count - sms) multiprocessors, so that a kernel of another stream — the
 * peer reduce of the previous frame (rt_multi.h), a post-processing chain — finds free CTA slots while a
 * persistent render kernel is running instead of waiting for its last CTA. 0 (default) = the whole device;
 * sms must be in [0, SM count). No counterpart in the reference (one GL queue, one dispatch at a time). */
rt_status rt_set_sm_reserve(rt_context* ctx, int sms);

/* Introspection of the host-side scene staging (no CUDA call, no context): builds the light grids
 * rt_upload_scene would build for these records (large scenes only) into
 * `out` (16-byte units) and returns how many units the blob has (0 = no grid for this scene); when
 * `capacity_f4` is too small nothing is written. Layout: 4 units of header per light, then the u32
 * cell offsets at unit *off_f4, then the u16 sphere-slot entries at unit *ent_f4; `slot_to_object`
 * (n ints, may be NULL) receives the original object index of every sphere slot. Used by the CPU
 * tests to check that the grids are conservative.                                              */
size_t    rt_debug_light_grids(const void* objects, int n, const void* lights, int m, void* out, size_t capacity_f4,
                               int* off_f4, int* ent_f4, int* slot_to_object);

#ifdef __cplusplus
}
#endif
#endif /* RT_ABI_H_ */
