// rt_headless — the reference's frame loop without a window (src/main.cpp + the ray-tracing slice
// of ForwardShadingPipline::Render), on top of host/rt_host.hpp and librt_b200.so.
//
//   rt_headless --scene default.scene --size 1920 1080 --frames 4 --out /tmp/frame
//   rt_headless --dump-scene in.scene out.bin        (SceneIO only; needs no GPU)
//   rt_headless --resave in.scene out.scene          (Load + Save round trip; needs no GPU)
//   rt_headless --resave-v2 in.scene out.scene       (Load + SaveV2: the lossless OBJECT2 / LIGHT2 records)
//
//   rt_headless --scene default.scene --gpus 0,1,2,3 --shard rows|samples ...   (several GPUs: include/rt_multi.h)
//
//   rt_headless --scene default.scene --post --frames 8 --out /tmp/frame   (whole iterations: ray tracing +
//                                        bloom + TAA on device-resident images; writes <out>_display.ppm)
//
// Outputs (row 0 = GL y 0): <out>_color.pfm, and with --raw also <out>_color.f32, <out>_position.f32
// (RGBA32F) and <out>_normal.f16 (RGBA16F) — the three images the reference's post chain consumes.
#include <chrono>
#include <cstdlib>
#include <iostream>
#include <memory>
#include <sstream>

#include "rt_host.hpp"

using namespace rtb200;

static int usage() {
  std::fprintf(stderr,
               "usage: rt_headless --scene FILE [--size W H] [--frames N] [--depth D] [--samples S] [--flags F]\n"
               "                   [--camera px py pz yaw pitch fov] [--noise FILE W H] [--skybox FILE SIZE CH f16|f32]\n"
               "                   [--device N | --gpus A,B,... [--shard rows|samples]] [--out PREFIX] [--raw] [--post]\n"
               "                   [--kernel megakernel|wavefront|regroup | --tune]   (one GPU: pin a variant / time them all once)\n"
               "       rt_headless --dump-scene IN.scene OUT.bin\n"
               "       rt_headless --resave IN.scene OUT.scene\n"
               "       rt_headless --resave-v2 IN.scene OUT.scene\n");
  return 2;
}

static int dump_scene(const char* in, const char* out) {
  std::vector<Object> objects;
  std::vector<Light> lights;
  SceneNames names;
  if (!SceneIO::Load(in, objects, lights, &names)) { std::fprintf(stderr, "cannot open %s\n", in); return 1; }
  // int32 n_objects, int32 n_lights, Object[n], Light[m]
  std::vector<unsigned char> blob(8 + objects.size() * sizeof(Object) + lights.size() * sizeof(Light));
  const int32_t n = (int32_t)objects.size(), m = (int32_t)lights.size();
  std::memcpy(blob.data(), &n, 4);
  std::memcpy(blob.data() + 4, &m, 4);
  if (n) std::memcpy(blob.data() + 8, objects.data(), objects.size() * sizeof(Object));
  if (m) std::memcpy(blob.data() + 8 + objects.size() * sizeof(Object), lights.data(), lights.size() * sizeof(Light));
  if (!WriteRaw(out, blob.data(), blob.size())) { std::fprintf(stderr, "cannot write %s\n", out); return 1; }
  for (const auto& s : names.objects) std::printf("OBJECT %s\n", s.c_str());
  for (const auto& s : names.lights) std::printf("LIGHT %s\n", s.c_str());
  return 0;
}

static int resave(const char* in, const char* out, bool v2) {
  std::vector<Object> objects;
  std::vector<Light> lights;
  SceneNames names;
  if (!SceneIO::Load(in, objects, lights, &names)) { std::fprintf(stderr, "cannot open %s\n", in); return 1; }
  try {
    return (v2 ? SceneIO::SaveV2(out, objects, lights, &names) : SceneIO::Save(out, objects, lights, &names)) ? 0 : 1;
  } catch (const std::exception& e) {
    std::fprintf(stderr, "rt_headless: %s\n", e.what());
    return 1;
  }
}

int main(int argc, char** argv) {
  if (argc >= 4 && std::string(argv[1]) == "--dump-scene") return dump_scene(argv[2], argv[3]);
  if (argc >= 4 && std::string(argv[1]) == "--resave") return resave(argv[2], argv[3], false);
  if (argc >= 4 && std::string(argv[1]) == "--resave-v2") return resave(argv[2], argv[3], true);

  std::string scene, out = "frame", noise_file, sky_file;
  int W = 800, H = 800, frames = 1, depth = 3, samples = 1, device = 0;   // src/global.cpp:4-5: 800 x 800
  int noise_w = 0, noise_h = 0, sky_size = 0, sky_ch = 4;
  bool sky_f16 = true, raw = false, have_cam = false, post = false, tune = false;
  int kernel = -1;
  std::vector<int> gpus;
  int sharding = RT_MULTI_ROWS;
  unsigned flags = 0;
  float cam[6] = {0, 0, 0, -90, 0, 45};
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto need = [&](int k) { if (i + k >= argc) { usage(); std::exit(2); } };
    if (a == "--scene") { need(1); scene = argv[++i]; }
    else if (a == "--size") { need(2); W = std::atoi(argv[++i]); H = std::atoi(argv[++i]); }
    else if (a == "--frames") { need(1); frames = std::atoi(argv[++i]); }
    else if (a == "--depth") { need(1); depth = std::atoi(argv[++i]); }
    else if (a == "--samples") { need(1); samples = std::atoi(argv[++i]); }
    else if (a == "--flags") { need(1); flags = (unsigned)std::strtoul(argv[++i], nullptr, 0); }
    else if (a == "--device") { need(1); device = std::atoi(argv[++i]); }
    else if (a == "--gpus") { need(1); std::stringstream ss(argv[++i]); std::string tok; while (std::getline(ss, tok, ',')) gpus.push_back(std::atoi(tok.c_str())); }
    else if (a == "--shard") { need(1); sharding = std::string(argv[++i]) == "samples" ? RT_MULTI_SAMPLES : RT_MULTI_ROWS; }
    else if (a == "--out") { need(1); out = argv[++i]; }
    else if (a == "--raw") raw = true;
    else if (a == "--post") post = true;
    else if (a == "--tune") tune = true;
    else if (a == "--kernel") { need(1); const std::string k = argv[++i];
      kernel = k == "megakernel" ? RT_KERNEL_MEGAKERNEL : k == "wavefront" ? RT_KERNEL_WAVEFRONT : k == "regroup" ? RT_KERNEL_REGROUP : -2;
      if (kernel == -2) return usage(); }
    else if (a == "--camera") { need(6); for (int k = 0; k < 6; ++k) cam[k] = (float)std::atof(argv[++i]); have_cam = true; }
    else if (a == "--noise") { need(3); noise_file = argv[++i]; noise_w = std::atoi(argv[++i]); noise_h = std::atoi(argv[++i]); }
    else if (a == "--skybox") { need(4); sky_file = argv[++i]; sky_size = std::atoi(argv[++i]); sky_ch = std::atoi(argv[++i]); sky_f16 = std::string(argv[++i]) == "f16"; }
    else return usage();
  }
  if (scene.empty()) return usage();

  try {
    std::unique_ptr<ForwardShadingPipeline> owned(gpus.empty() ? new ForwardShadingPipeline(W, H, device)
                                                               : new ForwardShadingPipeline(W, H, gpus, sharding));
    ForwardShadingPipeline& pipeline = *owned;
    if (!pipeline.LoadScene(scene)) { std::fprintf(stderr, "cannot open %s\n", scene.c_str()); return 1; }
    if (!noise_file.empty()) {
      const std::vector<unsigned char> n = ReadRaw(noise_file);
      if (n.size() != (size_t)noise_w * noise_h) throw std::runtime_error("noise file size != W*H");
      pipeline.SetNoise(n.data(), noise_w, noise_h);
    }
    if (!sky_file.empty()) {
      const std::vector<unsigned char> s = ReadRaw(sky_file);
      if (s.size() != (size_t)6 * sky_size * sky_size * sky_ch * (sky_f16 ? 2 : 4)) throw std::runtime_error("skybox file size");
      pipeline.SetSkybox(s.data(), sky_size, sky_ch, sky_f16);
    }
    if (have_cam) {
      pipeline.camera.Position = v3(cam[0], cam[1], cam[2]);
      pipeline.camera.Yaw = cam[3]; pipeline.camera.Pitch = cam[4]; pipeline.camera.FOV = cam[5];
      pipeline.camera.UpdateVectors();
    }
    pipeline.maxRayDepth = depth;
    pipeline.numSamples = samples;
    pipeline.flags = flags;
    static const char* const kKernelNames[] = {"megakernel", "wavefront", "literal", "regroup"};
    if (gpus.empty()) {
      pipeline.SetAutotune(false);                                // no trial frames inside the loop below
      if (kernel >= 0) pipeline.PinKernel(kernel);
      else if (tune) {
        const rt_tuning t = pipeline.Tune();
        std::fprintf(stderr, "rt_tune: %s (%d variants:", kKernelNames[t.kernel & 3], t.num_variants);
        for (int i = 0; i < t.num_variants; ++i) std::fprintf(stderr, " %s %.4f ms", kKernelNames[t.variant_kernel[i] & 3], t.variant_ms[i]);
        std::fprintf(stderr, ")\n");
      }
    }
    const auto t0 = std::chrono::steady_clock::now();
    for (int f = 0; f < frames; ++f) {
      if (post) pipeline.RenderFrame(raw && f == frames - 1);    // the whole iteration: ray tracing + bloom + TAA
      else pipeline.Render();
    }
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    std::printf("{\"scene\": \"%s\", \"objects\": %zu, \"lights\": %zu, \"width\": %d, \"height\": %d, \"frames\": %d, "
                "\"ms_per_frame_host\": %.4f, \"launches\": %llu, \"kernel\": \"%s\"}\n",
                scene.c_str(), pipeline.ssbo.objects.size(), pipeline.lightSSBO.lights.size(), W, H, frames, ms / frames,
                (unsigned long long)pipeline.launches(), gpus.empty() ? kKernelNames[pipeline.LastKernel().kernel & 3] : "per device");
    const size_t px = (size_t)W * H;
    if (post) {                                                  // what the window would show
      if (!WritePPM(out + "_display.ppm", pipeline.displayTex.data(), W, H)) throw std::runtime_error("cannot write " + out + "_display.ppm");
      if (raw && (!WriteRaw(out + "_display.rgba8", pipeline.displayTex.data(), px * 4) ||
                  !WriteRaw(out + "_display.f32", pipeline.displayTexF.data(), px * 16)))
        throw std::runtime_error("cannot write raw display outputs");
      return 0;
    }
    if (!WritePFM(out + "_color.pfm", pipeline.outputTex.data(), W, H)) throw std::runtime_error("cannot write " + out + "_color.pfm");
    if (raw) {
      const bool n32 = (flags & RT_FLAG_NORMAL_FP32) != 0;
      if (n32) throw std::runtime_error("--raw with RT_FLAG_NORMAL_FP32: the pipeline's normal image is RGBA16F");
      if (!WriteRaw(out + "_color.f32", pipeline.outputTex.data(), px * 16) ||
          !WriteRaw(out + "_position.f32", pipeline.gPositionTex.data(), px * 16) ||
          !WriteRaw(out + "_normal.f16", pipeline.gNormalTex.data(), px * 8))
        throw std::runtime_error("cannot write raw outputs");
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "rt_headless: %s\n", e.what());
    return 1;
  }
  return 0;
}
