// rt_host.hpp — header-only C++17 host layer over the C ABI (include/rt_abi.h), in the reference's
// own vocabulary: what a maintainer of the reference links instead of GL for the ray-tracing pass.
//
//   rtb200::Object / Light / Material     == the std430 structs (src/Object.h:8-21, src/Material.h:11-23,
//                                            src/Light.h:7-20) with the reference's default initialisers
//   rtb200::GenerateAABBForObject          src/SceneIO.h:75-104
//   rtb200::SceneIO::Load / Save           src/SceneIO.h:108-186 (the .scene text format)
//   rtb200::SceneIO::SaveV2                the lossless OBJECT2 / LIGHT2 records (Load reads both kinds)
//   rtb200::Camera                         src/Camera.h:7-42
//   rtb200::SSBO / LightSSBO               src/SSBO.h, src/LightSSBO.h  (vector + update())
//   rtb200::RaytracingShader               src/Shader.h:106-128 (use / set* by uniform name)
//   rtb200::ForwardShadingPipeline         the :155-182 slice of ForwardShadingPipline::Render,
//                                          headless: three output images in pinned host memory
//
// No GL, no window, no CPU fallback: every failure of the library throws std::runtime_error.
// Compile with -ffp-contract=off (GenerateAABBForObject and Camera must round like glm's FP32).
#pragma once

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <initializer_list>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "rt_abi.h"
#include "rt_frame.h"
#include "rt_multi.h"

namespace rtb200 {

// ---------------------------------------------------------------- FP32 vector algebra (glm order of operations)
struct vec3 {
  float x, y, z;
};
inline vec3 v3(float x, float y, float z) { return vec3{x, y, z}; }
inline vec3 v3(const float* p) { return vec3{p[0], p[1], p[2]}; }
inline vec3 operator+(vec3 a, vec3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(vec3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
inline float dot(vec3 a, vec3 b) {
  const float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z;
  return (tx + ty) + tz;
}
inline vec3 cross(vec3 a, vec3 b) {
  return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline vec3 normalize(vec3 a) { return a * (1.0f / std::sqrt(dot(a, a))); }   // glm::inversesqrt
inline void store(float* p, vec3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }

// ---------------------------------------------------------------- data contracts + defaults
using Material = rt_material;
using Object = rt_object;
using Light = rt_light;
enum ObjectType { SPHERE = 0, PLANE = 1 };                       // src/Object.h:6
enum LightType { POINT = 0, DIRECTIONAL = 1, AREA = 2 };         // src/Light.h:5
enum MaterialType { MATERIAL_METALLIC = 0, MATERIAL_DIELECTRIC = 1, MATERIAL_PLASTIC = 2 };
enum ShadowType { SHADOW_NONE = 0, SHADOW_PCF = 1, SHADOW_PCSS = 2 };

// Default member initialisers of src/Object.h:13-21 + src/Material.h:11-23. diffuseStrength has
// none in the reference (indeterminate); it is pinned to 0 here, padding is zero.
inline Object MakeObject() {
  Object o;
  std::memset(&o, 0, sizeof(o));
  o.type = SPHERE;
  o.radius = 1.0f;
  o.normal[1] = 1.0f;
  o.size[0] = o.size[1] = 1.0f;
  Material& m = o.material;
  m.type = MATERIAL_PLASTIC;
  m.albedo[0] = m.albedo[1] = m.albedo[2] = 1.0f;
  m.roughness = 0.5f;
  m.ior = 1.0f;
  m.specular = 0.5f;
  m.subsurfaceColor[0] = m.subsurfaceColor[1] = m.subsurfaceColor[2] = 1.0f;
  m.scatterDistance = 0.1f;
  return o;
}
// src/Light.h:7-20; angularRadius (no initialiser, dead in the shader) pinned to 0.
inline Light MakeLight() {
  Light l;
  std::memset(&l, 0, sizeof(l));
  l.type = POINT;
  l.direction[1] = -1.0f;
  l.color[0] = l.color[1] = l.color[2] = 1.0f;
  l.intensity = 1.0f;
  l.radius = 0.5f;
  l.samples = 4;
  l.shadowSoftness = 1.0f;
  l.shadowType = SHADOW_PCF;
  l.pcfSamples = 4;
  l.lightSize = 1.0f;
  return l;
}

// src/SceneIO.h:75-104. Sphere: centre -/+ radius. Plane: the in-plane rectangle, both corners
// then shifted by +0.01 * normal (a zero-thickness box, min > max on some axes for some
// orientations; the shader's slab test accepts both).
inline void GenerateAABBForObject(Object& obj) {
  const vec3 pos = v3(obj.position);
  if (obj.type == SPHERE) {
    const float r = obj.radius;
    store(obj.bounds.min, v3(pos.x - r, pos.y - r, pos.z - r));
    store(obj.bounds.max, v3(pos.x + r, pos.y + r, pos.z + r));
  } else if (obj.type == PLANE) {
    const vec3 n = v3(obj.normal);
    vec3 right, forward;
    if (std::fabs((double)n.y) > 0.9) {
      right = v3(1, 0, 0);
      forward = v3(0, 0, 1);
    } else {
      right = normalize(cross(n, v3(0, 1, 0)));
      forward = normalize(cross(right, n));
    }
    const vec3 hx = right * (obj.size[0] / 2.0f), hy = forward * (obj.size[1] / 2.0f);
    const vec3 shift = n * 0.01f;
    store(obj.bounds.min, ((pos - hx) - hy) + shift);
    store(obj.bounds.max, ((pos + hx) + hy) + shift);
  }
}

// ---------------------------------------------------------------- SceneIO (.scene text format)
struct SceneNames {
  std::vector<std::string> objects, lights;
};

class SceneIO {
 public:
  // One line = one record, whitespace-separated, read with operator>> exactly as the reference
  // does: a failed extraction stores 0 and sets failbit, after which every later field keeps its
  // default (the shipped performance_test.scene relies on this for its short lines).
  static void Parse(std::istream& in, std::vector<Object>& objects, std::vector<Light>& lights, SceneNames* names = nullptr) {
    objects.clear();
    lights.clear();
    if (names) { names->objects.clear(); names->lights.clear(); }
    std::string line;
    while (std::getline(in, line)) {
      std::istringstream iss(line);
      std::string kind, type, name;
      iss >> kind;
      if (kind == "OBJECT") {
        Object o = MakeObject();
        int matType = 0;                       // uninitialised in the reference; pinned 0
        iss >> type >> name;
        iss >> o.position[0] >> o.position[1] >> o.position[2];
        o.type = type == "PLANE" ? PLANE : SPHERE;
        iss >> o.radius >> o.normal[0] >> o.normal[1] >> o.normal[2] >> o.size[0] >> o.size[1];
        iss >> matType;
        o.material.type = matType;
        Material& m = o.material;
        iss >> m.albedo[0] >> m.albedo[1] >> m.albedo[2] >> m.metallic >> m.roughness >> m.ior >> m.transparency >> m.specular;
        GenerateAABBForObject(o);
        objects.push_back(o);
        if (names) names->objects.push_back(name);
      } else if (kind == "LIGHT") {
        Light l = MakeLight();
        iss >> type >> name;
        l.type = type == "DIRECTIONAL" ? DIRECTIONAL : type == "AREA" ? AREA : POINT;
        iss >> l.position[0] >> l.position[1] >> l.position[2] >> l.direction[0] >> l.direction[1] >> l.direction[2];
        iss >> l.color[0] >> l.color[1] >> l.color[2] >> l.intensity >> l.radius >> l.samples;
        lights.push_back(l);
        if (names) names->lights.push_back(name);
      } else if (kind == "OBJECT2") {          // the lossless record (WriteV2): every field the shader reads
        Object o = MakeObject();
        Material& m = o.material;
        iss >> type >> name;
        o.type = TypeV2(type, {"SPHERE", "PLANE"}, SPHERE);
        iss >> o.position[0] >> o.position[1] >> o.position[2] >> o.radius >> o.normal[0] >> o.normal[1] >> o.normal[2] >> o.size[0] >> o.size[1];
        iss >> m.type >> m.albedo[0] >> m.albedo[1] >> m.albedo[2] >> m.metallic >> m.roughness >> m.diffuseStrength >> m.ior >> m.transparency >>
            m.specular >> m.subsurfaceScatter >> m.subsurfaceColor[0] >> m.subsurfaceColor[1] >> m.subsurfaceColor[2] >> m.scatterDistance;
        GenerateAABBForObject(o);
        objects.push_back(o);
        if (names) names->objects.push_back(name);
      } else if (kind == "LIGHT2") {
        Light l = MakeLight();
        iss >> type >> name;
        l.type = TypeV2(type, {"POINT", "DIRECTIONAL", "AREA"}, POINT);
        iss >> l.position[0] >> l.position[1] >> l.position[2] >> l.direction[0] >> l.direction[1] >> l.direction[2];
        iss >> l.color[0] >> l.color[1] >> l.color[2] >> l.intensity >> l.radius >> l.samples;
        iss >> l.shadowSoftness >> l.shadowType >> l.pcfSamples >> l.lightSize >> l.angularRadius;
        lights.push_back(l);
        if (names) names->lights.push_back(name);
      }
    }
  }
  // Type token of a v2 record: one of `known` (its position is the tag), or TYPE<n> for a tag the
  // reference's enums do not name (never hit / black in the shader, but it must survive a round trip).
  static int TypeV2(const std::string& word, std::initializer_list<const char*> known, int fallback) {
    int tag = 0;
    for (const char* k : known) { if (word == k) return tag; ++tag; }
    if (word.compare(0, 4, "TYPE") == 0 && word.size() > 4) {
      char* end = nullptr;
      const long v = std::strtol(word.c_str() + 4, &end, 10);
      if (end && *end == 0) return (int)v;
    }
    return fallback;
  }
  // The lossless text format: the reference's columns in their order plus
  // diffuseStrength, the subsurface fields and the five shadow fields its own format drops
  // (src/SceneIO.h:124-142). 9 significant digits round-trip every finite FP32 value. Bounds are
  // recomputed on load. The reference's loader skips record kinds it does not know (:114-119).
  static void WriteV2(std::ostream& out, const std::vector<Object>& objects, const std::vector<Light>& lights, const SceneNames* names = nullptr) {
    auto num = [&](float v) {
      if (!std::isfinite(v)) throw std::runtime_error("SceneIO::WriteV2: non-finite value in a scene record");
      char buf[32];
      std::snprintf(buf, sizeof(buf), " %.9g", (double)v);
      out << buf;
    };
    auto tag = [&](int t, std::initializer_list<const char*> known) {
      int i = 0;
      for (const char* k : known) { if (i == t) { out << k; return; } ++i; }
      out << "TYPE" << t;
    };
    out << "# rt-scene 2\n";
    for (size_t i = 0; i < objects.size(); ++i) {
      const Object& o = objects[i];
      const Material& m = o.material;
      out << "OBJECT2 ";
      tag(o.type, {"SPHERE", "PLANE"});
      out << ' ' << (names && i < names->objects.size() ? names->objects[i] : "Object" + std::to_string(i));
      for (float v : {o.position[0], o.position[1], o.position[2], o.radius, o.normal[0], o.normal[1], o.normal[2], o.size[0], o.size[1]}) num(v);
      out << ' ' << m.type;
      for (float v : {m.albedo[0], m.albedo[1], m.albedo[2], m.metallic, m.roughness, m.diffuseStrength, m.ior, m.transparency, m.specular,
                      m.subsurfaceScatter, m.subsurfaceColor[0], m.subsurfaceColor[1], m.subsurfaceColor[2], m.scatterDistance}) num(v);
      out << '\n';
    }
    for (size_t i = 0; i < lights.size(); ++i) {
      const Light& l = lights[i];
      out << "LIGHT2 ";
      tag(l.type, {"POINT", "DIRECTIONAL", "AREA"});
      out << ' ' << (names && i < names->lights.size() ? names->lights[i] : "Light" + std::to_string(i));
      for (float v : {l.position[0], l.position[1], l.position[2], l.direction[0], l.direction[1], l.direction[2], l.color[0], l.color[1], l.color[2], l.intensity, l.radius}) num(v);
      out << ' ' << l.samples;
      num(l.shadowSoftness);
      out << ' ' << l.shadowType << ' ' << l.pcfSamples;
      num(l.lightSize);
      num(l.angularRadius);
      out << '\n';
    }
  }
  static bool SaveV2(const std::string& path, const std::vector<Object>& objects, const std::vector<Light>& lights, const SceneNames* names = nullptr) {
    std::ofstream f(path);
    if (!f.is_open()) return false;
    WriteV2(f, objects, lights, names);
    return (bool)f;
  }
  static bool Load(const std::string& path, std::vector<Object>& objects, std::vector<Light>& lights, SceneNames* names = nullptr) {
    std::ifstream f(path);
    if (!f.is_open()) return false;
    Parse(f, objects, lights, names);
    return true;
  }
  // The reference's (lossy) text format: no diffuseStrength, subsurface or shadow fields.
  static void Write(std::ostream& out, const std::vector<Object>& objects, const std::vector<Light>& lights, const SceneNames* names = nullptr) {
    for (size_t i = 0; i < objects.size(); ++i) {
      const Object& o = objects[i];
      const Material& m = o.material;
      out << "OBJECT " << (o.type == SPHERE ? "SPHERE" : o.type == PLANE ? "PLANE" : "UNKNOWN") << ' '
          << (names && i < names->objects.size() ? names->objects[i] : "Object" + std::to_string(i));
      for (float v : {o.position[0], o.position[1], o.position[2], o.radius, o.normal[0], o.normal[1], o.normal[2], o.size[0], o.size[1]}) out << ' ' << v;
      out << ' ' << m.type;
      for (float v : {m.albedo[0], m.albedo[1], m.albedo[2], m.metallic, m.roughness, m.ior, m.transparency, m.specular}) out << ' ' << v;
      out << '\n';
    }
    for (size_t i = 0; i < lights.size(); ++i) {
      const Light& l = lights[i];
      out << "LIGHT " << (l.type == DIRECTIONAL ? "DIRECTIONAL" : l.type == AREA ? "AREA" : "POINT") << ' '
          << (names && i < names->lights.size() ? names->lights[i] : "Light" + std::to_string(i));
      for (float v : {l.position[0], l.position[1], l.position[2], l.direction[0], l.direction[1], l.direction[2], l.color[0], l.color[1], l.color[2], l.intensity, l.radius}) out << ' ' << v;
      out << ' ' << l.samples << '\n';
    }
  }
  static bool Save(const std::string& path, const std::vector<Object>& objects, const std::vector<Light>& lights, const SceneNames* names = nullptr) {
    std::ofstream f(path);
    if (!f.is_open()) return false;
    Write(f, objects, lights, names);
    return true;
  }
};

// ---------------------------------------------------------------- Camera (src/Camera.h:7-42)
struct Camera {
  vec3 Position{0.0f, 0.0f, 0.0f};
  vec3 Front{0.0f, 0.0f, -1.0f};
  vec3 Up{0.0f, 1.0f, 0.0f};
  vec3 Right{1.0f, 0.0f, 0.0f};
  float FOV = 45.0f;
  float Yaw = -90.0f, Pitch = 0.0f;
  float FocalLength = 1.0f;

  static float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }   // glm::radians<float>
  void UpdateVectors() {                                                                  // :26-34
    const float ry = radians(Yaw), rp = radians(Pitch);
    Front = normalize(v3(std::cos(ry) * std::cos(rp), std::sin(rp), std::sin(ry) * std::cos(rp)));
    Right = normalize(cross(Front, v3(0, 1, 0)));
    Up = normalize(cross(Right, Front));
  }
};

// ---------------------------------------------------------------- library handle
inline void check(rt_context* ctx, rt_status st, const char* what) {
  if (st == RT_OK) return;
  std::string msg = std::string(what) + ": " + rt_status_string(st);
  if (ctx) { msg += " ("; msg += rt_last_error(ctx); msg += ")"; }
  throw std::runtime_error(msg);
}

// Pinned, device-mapped host image (rt_host_alloc): with it rt_render_host takes the zero-copy path.
template <typename T>
class HostImage {
 public:
  HostImage() = default;
  HostImage(const HostImage&) = delete;
  HostImage& operator=(const HostImage&) = delete;
  ~HostImage() { reset(0); }
  void reset(size_t count) {
    if (p_) rt_host_free(p_);
    p_ = nullptr; n_ = 0;
    if (count) {
      void* q = nullptr;
      check(nullptr, rt_host_alloc(&q, count * sizeof(T)), "rt_host_alloc");
      p_ = static_cast<T*>(q); n_ = count;
      std::memset(p_, 0, count * sizeof(T));
    }
  }
  T* data() { return p_; }
  const T* data() const { return p_; }
  size_t size() const { return n_; }
 private:
  T* p_ = nullptr;
  size_t n_ = 0;
};

class ForwardShadingPipeline;

// src/SSBO.h / src/LightSSBO.h: a std::vector the UI edits + update() (the reference re-uploads
// the whole buffer; here update() marks the scene for upload before the next dispatch).
struct SSBO {
  std::vector<Object> objects;
  bool dirty = true;
  void update() { dirty = true; }
};
struct LightSSBO {
  std::vector<Light> lights;
  bool dirty = true;
  void update() { dirty = true; }
};

// src/Shader.h:106-128 for raytracingCs.glsl's uniforms (glsl:72-89). Unknown names are ignored,
// as a glGetUniformLocation of -1 is. Shader-side defaults: fov 60, focalLength 1,
// maxRayDistance 114514, noiseScale unset (0).
class RaytracingShader {
 public:
  RaytracingShader() {
    std::memset(&params, 0, sizeof(params));
    params.fov_deg = 60.0f;
    params.focal_length = 1.0f;
    params.max_ray_distance = 114514.0f;
    params.camera_dir[2] = -1.0f; params.camera_up[1] = 1.0f; params.camera_right[0] = 1.0f;
    params.num_samples = 1;
    params.max_ray_depth = 3;                 // #define MAX_RAY_DEPTH 3 (glsl:4)
    params.row_period = 1;
  }
  void use() const {}
  void setInt(const std::string& name, int v) {
    if (name == "numObjects") params.num_objects = v;
    else if (name == "numLights") params.num_lights = v;
    else if (name == "frameCount") params.frame_count = v;
  }
  void setBool(const std::string& name, bool v) { if (name == "useSkybox") params.use_skybox = v ? 1 : 0; }
  void setFloat(const std::string& name, float v) {
    if (name == "fov") params.fov_deg = v;
    else if (name == "focalLength") params.focal_length = v;
    else if (name == "maxRayDistance") params.max_ray_distance = v;
  }
  void setVec3(const std::string& name, vec3 v) {
    float* dst = name == "cameraPos" ? params.camera_pos : name == "cameraDir" ? params.camera_dir
               : name == "cameraUp" ? params.camera_up : name == "cameraRight" ? params.camera_right : nullptr;
    if (dst) store(dst, v);
  }
  void setVec2(const std::string& name, float x, float y) {
    if (name == "noiseScale") { params.noise_scale[0] = x; params.noise_scale[1] = y; }
  }
  rt_frame_params params;
};

// The ray-tracing stage of ForwardShadingPipline (src/ForwardShadingPipeline.{h,cpp}): Init creates
// the three output images (outputTex RGBA32F :57-65, gPositionTex RGBA32F :115-120, gNormalTex
// RGBA16F :121-126); Render performs the :155-182 slice of one frame through rt_render_host.
// RenderFrame performs the whole iteration (ray tracing + post chain) through include/rt_frame.h.
// Window and UI are not part of this layer.
class ForwardShadingPipeline {
 public:
  ForwardShadingPipeline(int width = 800, int height = 800, int device = 0) : WIDTH(width), HEIGHT(height) {
    if (rt_abi_version() != RT_ABI_VERSION) throw std::runtime_error("librt_b200.so: ABI version mismatch");
    check(nullptr, rt_create(&ctx_, device), "rt_create");
    Init();
  }
  // Several GPUs of the box behind the same Render(): include/rt_multi.h (devices[0] holds the images;
  // sharding = RT_MULTI_ROWS or RT_MULTI_SAMPLES). RenderFrame (the post chain) stays on one GPU.
  ForwardShadingPipeline(int width, int height, const std::vector<int>& devices, int sharding) : WIDTH(width), HEIGHT(height), sharding_(sharding) {
    if (rt_abi_version() != RT_ABI_VERSION) throw std::runtime_error("librt_b200.so: ABI version mismatch");
    const rt_status st = rt_multi_create(&multi_, devices.data(), (int)devices.size());
    if (st != RT_OK) throw std::runtime_error(std::string("rt_multi_create: ") + rt_status_string(st));
    ctx_ = rt_multi_context(multi_, 0);
    Init();
  }
  ForwardShadingPipeline(const ForwardShadingPipeline&) = delete;
  ForwardShadingPipeline& operator=(const ForwardShadingPipeline&) = delete;
  ~ForwardShadingPipeline() {
    if (frame_) rt_frame_destroy(frame_);  // a frame goes before its context
    if (multi_) rt_multi_destroy(multi_);  // owns its contexts
    else if (ctx_) rt_destroy(ctx_);
  }

  void Init() {
    const size_t px = (size_t)WIDTH * HEIGHT;
    outputTex.reset(px * 4);
    gPositionTex.reset(px * 4);
    gNormalTex.reset(px * 4);             // 4 x fp16 per pixel
  }
  // InitBlueNoiseTex (:39-49): R8, REPEAT wrap. The reference decodes a PNG; here raw bytes.
  void SetNoise(const unsigned char* r8, int w, int h) {
    if (multi_) mcheck(rt_multi_set_noise(multi_, r8, w, h), "rt_multi_set_noise");
    else check(ctx_, rt_set_noise(ctx_, r8, w, h), "rt_set_noise");
  }
  // glBindTexture(GL_TEXTURE_CUBE_MAP, ...) (:167-170): 6 faces (+X,-X,+Y,-Y,+Z,-Z) of size^2 texels.
  void SetSkybox(const void* faces, int face_size, int channels, bool fp16) {
    if (multi_) mcheck(rt_multi_set_skybox(multi_, faces, face_size, channels, fp16 ? 1 : 0), "rt_multi_set_skybox");
    else check(ctx_, rt_set_skybox(ctx_, faces, face_size, channels, fp16 ? 1 : 0), "rt_set_skybox");
    skyboxEnabled = faces != nullptr;
  }
  bool LoadScene(const std::string& path) {
    if (!SceneIO::Load(path, ssbo.objects, lightSSBO.lights, &names)) return false;
    ssbo.update(); lightSSBO.update();
    return true;
  }

  void SetUniforms() {                     // :153-166
    RaytracingShader& sh = raytracingShader;
    sh.use();
    sh.setInt("numObjects", (int)ssbo.objects.size());
    sh.setInt("numLights", (int)lightSSBO.lights.size());
    sh.setVec3("cameraPos", camera.Position);
    sh.setVec3("cameraDir", camera.Front);
    sh.setVec3("cameraUp", camera.Up);
    sh.setVec3("cameraRight", camera.Right);
    sh.setFloat("fov", camera.FOV);
    sh.setInt("frameCount", frameCount);
    sh.setVec2("noiseScale", 1.0f / 1024.0f, 1.0f / 1024.0f);
    sh.setBool("useSkybox", skyboxEnabled);
    rt_frame_params& p = sh.params;
    p.width = WIDTH; p.height = HEIGHT; p.row_begin = 0; p.row_end = HEIGHT;
    p.max_ray_depth = maxRayDepth; p.num_samples = numSamples; p.flags = flags;
  }
  // One frame's ray-tracing stage. Outputs land in outputTex / gPositionTex / gNormalTex.
  void Render() {
    const bool upload = ssbo.dirty || lightSSBO.dirty;
    SetUniforms();
    rt_frame_params& p = raytracingShader.params;
    if (multi_) {
      mcheck(rt_multi_render_host(multi_, &p, sharding_, upload ? ssbo.objects.data() : nullptr, (int)ssbo.objects.size(),
                                  upload ? lightSSBO.lights.data() : nullptr, (int)lightSSBO.lights.size(),
                                  outputTex.data(), gPositionTex.data(), gNormalTex.data(), nullptr), "rt_multi_render_host");
      ssbo.dirty = lightSSBO.dirty = false;
      if (taaEnabled) ++frameCount;
      return;
    }
    if (upload)                            // ssbo.update() + lightSSBO.update() (src/SSBO.h:16-23)
      check(ctx_, rt_upload_scene(ctx_, ssbo.objects.data(), (int)ssbo.objects.size(), lightSSBO.lights.data(),
                                  (int)lightSSBO.lights.size(), nullptr, 0), "rt_upload_scene");
    check(ctx_, rt_render_host(ctx_, &p, nullptr, 0, nullptr, 0, outputTex.data(), gPositionTex.data(), gNormalTex.data(), nullptr),
          "rt_render_host");
    ssbo.dirty = lightSSBO.dirty = false;
    if (taaEnabled) ++frameCount;          // the reference advances frameCount inside the TAA branch (:232-254)
  }

  // Kernel variants (identical images; include/rt_abi.h rt_tune / rt_set_tuning / rt_set_autotune / rt_set_sm_reserve).
  // No counterpart in the reference (one GLSL program). One GPU only: an rt_multi owns and tunes its contexts itself.
  rt_tuning Tune() {                       // times every variant on the CURRENT scene and frame shape, keeps the fastest
    if (multi_) throw std::runtime_error("Tune: one GPU only");
    if (ssbo.dirty || lightSSBO.dirty) {
      check(ctx_, rt_upload_scene(ctx_, ssbo.objects.data(), (int)ssbo.objects.size(), lightSSBO.lights.data(),
                                  (int)lightSSBO.lights.size(), nullptr, 0), "rt_upload_scene");
      ssbo.dirty = lightSSBO.dirty = false;
    }
    SetUniforms();
    rt_tuning t;
    check(ctx_, rt_tune(ctx_, &raytracingShader.params, &t), "rt_tune");
    return t;
  }
  void PinKernel(int kernel) {             // RT_KERNEL_MEGAKERNEL / _WAVEFRONT / _REGROUP; -1 removes the pin
    if (multi_) throw std::runtime_error("PinKernel: one GPU only");
    if (kernel < 0) { check(ctx_, rt_set_tuning(ctx_, nullptr), "rt_set_tuning"); return; }
    rt_tuning t;
    std::memset(&t, 0, sizeof(t));
    t.kernel = kernel; t.registers = -1; t.phase_level = -1;
    check(ctx_, rt_set_tuning(ctx_, &t), "rt_set_tuning");
  }
  rt_tuning LastKernel() {                 // what the last Render() ran
    rt_tuning t;
    check(ctx_, rt_get_tuning(ctx_, nullptr, &t), "rt_get_tuning");
    return t;
  }
  void SetAutotune(bool on) { check(ctx_, rt_set_autotune(ctx_, on ? 1 : 0), "rt_set_autotune"); }
  void SetSmReserve(int sms) { check(ctx_, rt_set_sm_reserve(ctx_, sms), "rt_set_sm_reserve"); }

  // The host's haltonSequence (src/global.cpp), float arithmetic, for the TAA jitter (:242-243).
  static float haltonSequence(int index, int base) {
    float f = 1.0f, r = 0.0f;
    while (index > 0) {
      f = f / (float)base;
      r = r + f * (float)(index % base);
      index = index / base;
    }
    return r;
  }

  // One WHOLE iteration of Render (:139-272) through rt_frame: ray tracing, bloom (threshold 1.0,
  // 10 blurs, strength 0.5 — the values the reference hard-codes, :198/:208/:223), combine, TAA
  // resolve into historyTex[frameCount % 2]; the three intermediate images stay on the device as
  // the reference's GL textures do. The presented image lands in displayTex (8-bit UNORM, the
  // window's framebuffer) — and, on request, in displayTexF (RGBA32F, before that conversion).
  void RenderFrame(bool readFloat = false) {
    if (multi_) throw std::runtime_error("RenderFrame: the post chain reads the whole image; a frame is not sharded (use one GPU)");
    if (!frame_) {
      check(ctx_, rt_frame_create(ctx_, WIDTH, HEIGHT, &frame_), "rt_frame_create");
      displayTex.reset((size_t)WIDTH * HEIGHT * 4);
    }
    SetUniforms();
    rt_post_params pp;
    std::memset(&pp, 0, sizeof(pp));
    pp.width = WIDTH; pp.height = HEIGHT;
    pp.bloom_threshold = 1.0f; pp.bloom_strength = 0.5f; pp.blur_passes = 10;
    pp.taa_enabled = taaEnabled ? 1 : 0;
    pp.taa_blend = taaBlendFactor;
    pp.jitter_x = (float)(haltonSequence(frameCount % 8, 2) * 0.5 / WIDTH);      // double arithmetic, then setFloat
    pp.jitter_y = (float)(haltonSequence(frameCount % 8, 3) * 0.5 / HEIGHT);
    const bool upload = ssbo.dirty || lightSSBO.dirty;
    check(ctx_, rt_frame_render_host(frame_, &raytracingShader.params, &pp, upload ? ssbo.objects.data() : nullptr, (int)ssbo.objects.size(),
                                     upload ? lightSSBO.lights.data() : nullptr, (int)lightSSBO.lights.size(), displayTex.data(), RT_DISPLAY_RGBA8),
          "rt_frame_render_host");
    if (readFloat) {
      if (displayTexF.size() == 0) displayTexF.reset((size_t)WIDTH * HEIGHT * 4);
      check(ctx_, rt_frame_read_display(frame_, displayTexF.data(), RT_DISPLAY_RGBA32F, nullptr), "rt_frame_read_display");
    }
    ssbo.dirty = lightSSBO.dirty = false;
    if (taaEnabled) ++frameCount;
  }

  rt_context* context() { return ctx_; }
  rt_frame* frame() { return frame_; }
  rt_multi* multi() { return multi_; }
  uint64_t launches() { return multi_ ? rt_multi_launch_count(multi_) : rt_launch_count(ctx_); }

  const int WIDTH, HEIGHT;                 // src/global.cpp:4-5
  Camera camera;
  SSBO ssbo;
  LightSSBO lightSSBO;
  SceneNames names;
  RaytracingShader raytracingShader;
  HostImage<float> outputTex, gPositionTex;
  HostImage<uint16_t> gNormalTex;
  HostImage<unsigned char> displayTex;     // RenderFrame: the window's 8-bit framebuffer
  HostImage<float> displayTexF;            // RenderFrame(true): the same image before the 8-bit conversion
  float taaBlendFactor = 0.1f;             // ImGUIManager's TAA blend slider default
  int frameCount = 0;
  int maxRayDepth = 3;
  int numSamples = 1;
  uint32_t flags = 0;
  bool skyboxEnabled = false;
  bool taaEnabled = true;

 private:
  void mcheck(rt_status st, const char* what) {
    if (st != RT_OK) throw std::runtime_error(std::string(what) + ": " + rt_multi_last_error(multi_));
  }
  rt_context* ctx_ = nullptr;
  rt_frame* frame_ = nullptr;
  rt_multi* multi_ = nullptr;
  int sharding_ = RT_MULTI_ROWS;
};

// ---------------------------------------------------------------- image writers (headless "display")
// Portable float map, RGB, bottom row first — the same row order as the GL image (row 0 = y 0).
inline bool WritePFM(const std::string& path, const float* rgba, int w, int h) {
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) return false;
  std::fprintf(f, "PF\n%d %d\n-1.0\n", w, h);
  std::vector<float> row((size_t)w * 3);
  for (int y = 0; y < h; ++y) {
    for (int x = 0; x < w; ++x)
      for (int c = 0; c < 3; ++c) row[(size_t)x * 3 + c] = rgba[((size_t)y * w + x) * 4 + c];
    std::fwrite(row.data(), sizeof(float), row.size(), f);
  }
  return std::fclose(f) == 0;
}
// Binary PPM of an RGBA8 image, top row first (the GL image has row 0 at the bottom).
inline bool WritePPM(const std::string& path, const unsigned char* rgba8, int w, int h) {
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) return false;
  std::fprintf(f, "P6\n%d %d\n255\n", w, h);
  std::vector<unsigned char> row((size_t)w * 3);
  for (int y = h - 1; y >= 0; --y) {
    for (int x = 0; x < w; ++x)
      for (int c = 0; c < 3; ++c) row[(size_t)x * 3 + c] = rgba8[((size_t)y * w + x) * 4 + c];
    std::fwrite(row.data(), 1, row.size(), f);
  }
  return std::fclose(f) == 0;
}
inline bool WriteRaw(const std::string& path, const void* p, size_t bytes) {
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) return false;
  const bool ok = std::fwrite(p, 1, bytes, f) == bytes;
  return (std::fclose(f) == 0) && ok;
}
inline std::vector<unsigned char> ReadRaw(const std::string& path) {
  std::ifstream f(path, std::ios::binary);
  if (!f.is_open()) throw std::runtime_error("cannot open " + path);
  return std::vector<unsigned char>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

}  // namespace rtb200
