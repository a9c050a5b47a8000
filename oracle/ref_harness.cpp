// ref_harness.cpp — TEST INFRASTRUCTURE (oracle/_ref build). Not part of the product.
//
// Wraps the reference's own shader source — /root/reference/shader/raytracingCs.glsl,
// mechanically rewritten to C++ by make_ref.py into REF_SHADER_INC (a temp file outside
// the repo; reference sources are never copied into the tree) — behind the oracle C API
// (oracle_api.h). The shader text is included inside namespace glsl so that its calls
// resolve to the shim's built-ins. The host side mirrors what
// src/ForwardShadingPipeline.cpp:155-180 does before glDispatchCompute: set the uniforms,
// bind the two SSBOs, three images and two samplers, then run main() once per pixel.
#include <vector>
#include <omp.h>
#include "glsl_shim.h"
#include "oracle_api.h"

namespace glsl {
thread_local uvec3 gl_GlobalInvocationID;
thread_local float* g_store[3] = {nullptr, nullptr, nullptr};
#include REF_SHADER_INC
}  // namespace glsl

using namespace glsl;

static vec3 v3(const float* p) { return vec3(p[0], p[1], p[2]); }

extern "C" const char* oracle_kind(void) { return "reference"; }

extern "C" int oracle_render(const rt_frame_params* p, const void* objects_, const void* lights_,
                             const float* sky_rgba32f, int face_size,
                             const unsigned char* noise_r8, int noise_w, int noise_h,
                             int x0, int y0, int sx, int sy, int nx, int ny,
                             float* color, float* position, float* normal, int32_t* object_id,
                             uint64_t* counters, int nthreads) {
  if (!p || !color || !position || !normal) return 1;
  const rt_object* ro = (const rt_object*)objects_;
  const rt_light* rl = (const rt_light*)lights_;
  std::vector<Object> objs((size_t)(p->num_objects > 0 ? p->num_objects : 0));
  std::vector<Light> lts((size_t)(p->num_lights > 0 ? p->num_lights : 0));
  for (size_t i = 0; i < objs.size(); ++i) {           // std430 bytes -> the shader's structs
    Object& o = objs[i]; const rt_object& s = ro[i];
    o.type = s.type; o.position = v3(s.position); o.radius = s.radius; o.normal = v3(s.normal);
    o.size = vec2(s.size[0], s.size[1]);
    o.material.type = s.material.type; o.material.albedo = v3(s.material.albedo);
    o.material.metallic = s.material.metallic; o.material.roughness = s.material.roughness;
    o.material.diffuseStrength = s.material.diffuseStrength; o.material.ior = s.material.ior;
    o.material.transparency = s.material.transparency; o.material.specular = s.material.specular;
    o.material.subsurfaceScatter = s.material.subsurfaceScatter;
    o.material.subsurfaceColor = v3(s.material.subsurfaceColor);
    o.material.scatterDistance = s.material.scatterDistance;
    o.bounds.min = v3(s.bounds.min); o.bounds.max = v3(s.bounds.max);
  }
  for (size_t i = 0; i < lts.size(); ++i) {
    Light& l = lts[i]; const rt_light& s = rl[i];
    l.type = s.type; l.position = v3(s.position); l.direction = v3(s.direction); l.color = v3(s.color);
    l.intensity = s.intensity; l.radius = s.radius; l.samples = s.samples;
    l.shadowSoftness = s.shadowSoftness; l.shadowType = s.shadowType; l.pcfSamples = s.pcfSamples;
    l.lightSize = s.lightSize; l.angularRadius = s.angularRadius;
  }
  // uniforms (src/ForwardShadingPipeline.cpp:156-166) + the never-set defaults (glsl:79-85)
  glsl::objects = objs.data(); glsl::lights = lts.data();
  numObjects = p->num_objects; numLights = p->num_lights;
  cameraPos = v3(p->camera_pos); cameraDir = v3(p->camera_dir);
  cameraUp = v3(p->camera_up); cameraRight = v3(p->camera_right);
  fov = p->fov_deg; focalLength = p->focal_length; maxRayDistance = p->max_ray_distance;
  useSkybox = p->use_skybox != 0;
  noiseScale = vec2(p->noise_scale[0], p->noise_scale[1]);
  MAX_RAY_DEPTH = p->max_ray_depth;
  outputImage.slot = 0; gPosition.slot = 1; gNormal.slot = 2;
  outputImage.width = gPosition.width = gNormal.width = p->width;
  outputImage.height = gPosition.height = gNormal.height = p->height;
  skybox.faces = sky_rgba32f; skybox.size = face_size;
  blueNoiseTex.r8 = noise_r8; blueNoiseTex.width = noise_w; blueNoiseTex.height = noise_h;
  if (counters) for (int k = 0; k < RT_CNT_COUNT; ++k) counters[k] = 0;   // _ref carries no counters

  if (object_id) for (size_t o = 0; o < (size_t)nx * ny; ++o) object_id[o] = -2;
  const int ns = p->num_samples > 0 ? p->num_samples : 1;
  if (nthreads <= 0) nthreads = omp_get_max_threads();
  for (int s = 0; s < ns; ++s) {
    frameCount = p->frame_count + s;
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
    for (int j = 0; j < ny; ++j) {
      for (int i = 0; i < nx; ++i) {
        float c[4], P[4], N[4];
        g_store[0] = c; g_store[1] = P; g_store[2] = N;
        gl_GlobalInvocationID = uvec3((unsigned)(x0 + i * sx), (unsigned)(y0 + j * sy), 0u);
        shader_main();
        size_t o = ((size_t)j * nx + i) * 4;
        if (s == 0) {
          for (int k = 0; k < 4; ++k) { color[o + k] = c[k]; position[o + k] = P[k]; normal[o + k] = N[k]; }
        } else {
          for (int k = 0; k < 3; ++k) color[o + k] = color[o + k] + c[k];
        }
      }
    }
  }
  if (ns > 1) {
    const float inv = 1.0f / (float)ns;
    for (size_t o = 0; o < (size_t)nx * ny; ++o)
      for (int k = 0; k < 3; ++k) color[o * 4 + k] = color[o * 4 + k] * inv;
  }
  return 0;
}
