/*
 * rt_multi.h — multi-GPU entry points of librt_b200.so: the ray-tracing pass of
 * /root/reference/shader/raytracingCs.glsl sharded over the GPUs of one box.
 *
 * The reference has one GPU and one dispatch (src/ForwardShadingPipeline.cpp:175-180); its C++ host reaches
 * more than one B200 through this header. Two shardings, both with identical per-pixel arithmetic:
 *   RT_MULTI_ROWS     GPU g renders the interleaved 32-row bands g, g+G, ... of the frame; its kernel stores
 *                     float4 / half4 results straight into the ROOT GPU's images over NVLink (peer access):
 *                     the gather is the kernel's own epilogue, there is no collective. Bit-identical to one GPU.
 *   RT_MULTI_SAMPLES  GPU g renders its range of the num_samples frame_count indices into a local FP32 partial
 *                     sum; one peer pull-reduce kernel per GPU then sums slice g of all partials IN RANK ORDER
 *                     (deterministic), scales by 1/num_samples and stores into the root's colour image.
 *                     Position / normal / object id are those of sample 0 (GPU 0 renders them).
 *
 *   rt_multi          ONE host process drives all GPUs (what the reference's single-threaded render loop is):
 *                     N rt_contexts, peer access enabled, cross-device ordering by CUDA events.
 *   rt_peer_reduce    the same reduce for ONE-PROCESS-PER-GPU deployments (torch.distributed / MPI ranks): the
 *                     partial frames are mapped into every rank with rt_ipc_export / rt_ipc_open, and the ranks
 *                     synchronise through flag words in peer memory inside the kernel — no NCCL call.
 *
 * Same conventions as rt_abi.h: plain C, every call returns an rt_status, no CPU fallback.
 */
#ifndef RT_MULTI_H_
#define RT_MULTI_H_

#include "rt_abi.h"

#ifdef __cplusplus
extern "C" {
#endif

#define RT_MAX_PEERS 16

typedef struct rt_multi rt_multi;   /* opaque; owns one rt_context per device; NOT thread-safe */

enum { RT_MULTI_ROWS = 0, RT_MULTI_SAMPLES = 1 };

/* devices[0] is the root: the three output images live there. A device may be listed more than once (several
 * contexts time-share it: functional testing on a one-GPU box). Enables peer access between distinct devices;
 * RT_ERR_UNAVAILABLE when a pair cannot reach each other. */
rt_status   rt_multi_create(rt_multi** out, const int* devices, int n);
void        rt_multi_destroy(rt_multi* m);
const char* rt_multi_last_error(rt_multi* m);
int         rt_multi_size(rt_multi* m);
rt_context* rt_multi_context(rt_multi* m, int i);      /* the i-th device's context (scene / texture setters, tuning) */

/* Replicated inputs: SSBO::update + LightSSBO::update (src/SSBO.h:16-23, src/LightSSBO.h:16-25), the cube map
 * (src/ForwardShadingPipeline.cpp:167-170) and the noise texture (:39-49) on every device. */
rt_status rt_multi_upload_scene(rt_multi* m, const void* objects, int n, const void* lights, int nl);
rt_status rt_multi_set_skybox(rt_multi* m, const void* faces, int face_size, int channels, int is_fp16);
rt_status rt_multi_set_noise(rt_multi* m, const unsigned char* r8, int w, int h);

/* One frame over all devices. The output pointers are DEVICE pointers on the root device (devices[0]) to
 * full-frame images, as for rt_render; any may be NULL. params describes the whole frame (row window and row
 * bands must be unset); RT_FLAG_ACCUMULATE / RT_FLAG_SUM_SAMPLES are not accepted (the sharding uses them).
 * Asynchronous: returns when everything is enqueued; rt_multi_synchronize waits for the frame. */
rt_status rt_multi_render(rt_multi* m, const rt_frame_params* params, int sharding,
                          float* d_color_rgba32f, float* d_position_rgba32f, void* d_normal, int32_t* d_object_id);
rt_status rt_multi_synchronize(rt_multi* m);

/* The same through HOST buffers (what rt_render_host is for one GPU): scene upload to every device (objects /
 * lights may be NULL to reuse the last upload), the frame, the three images copied into host memory. Synchronous. */
rt_status rt_multi_render_host(rt_multi* m, const rt_frame_params* params, int sharding,
                               const void* objects, int n, const void* lights, int nl,
                               float* h_color_rgba32f, float* h_position_rgba32f, void* h_normal, int32_t* h_object_id);

/* Kernels launched through m (render launches of the member contexts + the reduce kernels). */
uint64_t  rt_multi_launch_count(rt_multi* m);

/* ---- one process per GPU: the reduce step alone -------------------------------------------------------- */
#define RT_PEER_FLAG_WORDS 64      /* per rank: [0,16) arrived[src], [16,32) done[src] (rank 0's block), [32] CTA counter, [33] status */
typedef struct rt_peer_reduce_desc {
  int32_t      rank, world;                 /* this process's rank; world <= RT_MAX_PEERS                               */
  const float* partial[RT_MAX_PEERS];       /* RGBA32F partial sums of THIS epoch on every rank (own + rt_ipc_open'ed)  */
  float*       result;                      /* rank 0's RGBA32F frame (peer-mapped on the other ranks)                  */
  uint32_t*    flags[RT_MAX_PEERS];         /* every rank's flag block: RT_PEER_FLAG_WORDS zero-initialised uint32      */
  uint64_t     pixels;                      /* width * height                                                           */
  float        scale;                       /* 1 / total samples                                                        */
  uint32_t     epoch;                       /* 1, 2, 3, ...: the same sequence on every rank, one per call              */
} rt_peer_reduce_desc;

/* Enqueues ONE kernel on `cuda_stream` that (1) tells every peer that this rank's partial of `epoch` is complete
 * (everything enqueued on the stream before the call), (2) waits until every peer said so, (3) sums slice `rank`
 * of all partials in rank order, scales, and stores RGB + alpha 1 into `result`, (4) tells rank 0 the slice is
 * there; on rank 0 the kernel ends only when every slice is. Callers rotate the partial buffers between
 * consecutive epochs (a rank may render epoch e+1 while a peer still reads its partial of epoch e): TWO when the
 * render and the reduce share a stream; THREE when the reduce of epoch e runs on its own stream beside the render
 * of epoch e+1 (rt_set_sm_reserve leaves it CTA slots) — then order the render of epoch e+2 behind this rank's
 * reduce of epoch e: that kernel got past step (2) only after every peer had finished its reduce of epoch e-1,
 * the last reader of the buffer epoch e+2 overwrites. Waits are bounded (about 10 s): a peer that never arrives
 * sets the status word and rt_peer_reduce_status reports it. */
rt_status rt_peer_reduce(rt_context* ctx, const rt_peer_reduce_desc* desc, void* cuda_stream);
/* After synchronising: RT_OK, or RT_ERR_UNAVAILABLE when a wait of an earlier reduce timed out. own_flags = this rank's block. */
rt_status rt_peer_reduce_status(rt_context* ctx, const uint32_t* own_flags);

#ifdef __cplusplus
}
#endif
#endif /* RT_MULTI_H_ */
