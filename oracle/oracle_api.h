// oracle_api.h — TEST INFRASTRUCTURE. The C API both CPU oracles export:
//   oracle/rt_oracle.cpp       ("port": this repo's CPU restatement of the shader)
//   oracle/_ref/libref_shader  ("reference": the shader's own source compiled as C++)
// Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load them.
#pragma once
#include <stdint.h>
#include "../include/rt_abi.h"

#ifdef __cplusplus
extern "C" {
#endif

// "port" or "reference".
const char* oracle_kind(void);

// Renders the nx*ny pixels (x0 + i*sx, y0 + j*sy) of the width x height frame described by
// *p (row_begin/row_end are ignored) into COMPACT nx*ny RGBA32F arrays.
//   sky_rgba32f : 6 x face_size^2 x 4 floats (values already rounded to FP16), or NULL
//   noise_r8    : noise_w x noise_h bytes, or NULL => texel (0,0,0,1)
//   colour      : mean over p->num_samples frames (sequential FP32 sum, then * 1/n);
//                 position/normal come from the first sample.
//   object_id   : nx*ny int32, index of the last-hit object (-1 = primary miss), or NULL.
//                 (_ref writes -2: the shader has no such output.)
//   counters    : RT_CNT_COUNT reference-literal op counters summed over all rendered
//                 pixels and samples, or NULL. (_ref returns zeros: it is uninstrumented.)
//   nthreads    : OpenMP threads, <= 0 => all.
// Returns 0 on success.
int oracle_render(const rt_frame_params* p, const void* objects, const void* lights,
                  const float* sky_rgba32f, int face_size,
                  const unsigned char* noise_r8, int noise_w, int noise_h,
                  int x0, int y0, int sx, int sy, int nx, int ny,
                  float* color, float* position, float* normal, int32_t* object_id,
                  uint64_t* counters, int nthreads);

#ifdef __cplusplus
}
#endif
