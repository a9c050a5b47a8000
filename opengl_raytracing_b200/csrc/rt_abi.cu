// rt_abi.cu — host side of the C ABI (include/rt_abi.h): context, scene staging, texture
// setup, launch, host-buffer (e2e) path. The reference-side call sites each entry point
// replaces are cited in the header. No CPU rendering path exists here.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <utility>
#include <string>
#include <vector>
#include <cuda_fp16.h>
#include <nvtx3/nvToolsExt.h>
#include "rt_device.h"

namespace rt {
size_t scene_smem_bytes(int num_objects, int num_planes, int num_lights) {
  return sizeof(float4) * ((size_t)3 * num_objects + (size_t)3 * num_planes + (size_t)6 * num_lights);
}
cudaError_t measure_fp32_peak(int sm_count, cudaStream_t stream, double* tflops, double* ms);   // rt_peak.cu
}  // namespace rt

// ---- rt_render_static: the wavefront queues in a file-scope device array instead of a cudaMalloc'ed buffer, for callers that
// may not allocate (include/rt_abi.h). Every GPU that loads the library holds these 64 MiB until the process exits. In HOST
// code the name below is only a stand-in for the symbol: its device address comes from cudaGetSymbolAddress (rt_create, once
// per context, with the context's device current) and is what the kernels are given.
constexpr size_t kStaticScratchBytes = (size_t)64 << 20;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

struct rt_context {
  int device = 0;
  int sm_count = 0;
  int sm_reserve = 0;            // rt_set_sm_reserve: render grids are sized for sm_count - sm_reserve SMs
  int max_smem_optin = 0;
  std::string err;
  // scene blob (device) + pinned staging (host)
  float4* d_scene = nullptr; size_t d_scene_cap = 0;
  float4* h_stage = nullptr; size_t h_stage_cap = 0;
  cudaEvent_t stage_done = nullptr; bool stage_pending = false;
  int n_obj = 0, n_planes = 0, n_lights = 0;
  size_t off_plane = 0, off_material = 0, off_lights = 0;   // in float4 units
  // packet-kernel blob (appended to the same device buffer)
  size_t off_pk = 0, pk_float4s = 0; int pk_spheres = 0, pk_planes = 0, pk_lights_offset = 0, pk_clusters = 0, pk_fan_ok = 0, pk_material_offset = -1;
  // textures
  uint2* d_sky = nullptr; cudaArray_t sky_array = nullptr; cudaTextureObject_t sky_tex = 0; int sky_size = 0;
  unsigned char* d_noise = nullptr; int noise_w = 0, noise_h = 0;
  // work distribution + counters. Every launch takes the next of kCounterSlots work counters (each
  // monotonically increasing, 128 B apart); a slot is reused only behind the event its previous
  // launch recorded, so launches that overlap on different streams never share a counter.
  static constexpr int kCounterSlots = 16;
  unsigned* d_tile_counter = nullptr;
  struct CounterSlot { unsigned base = 0; cudaEvent_t done = nullptr; bool used = false; } slots[kCounterSlots];
  int next_slot = 0;
  unsigned long long* d_counters = nullptr;
  uint64_t launches = 0;
  // e2e path
  int host_path = RT_HOST_PATH_AUTO;
  int coherent_override = -1;   // RT_PK_COHERENT=0..3 pins the phase level (both kernels)
  // Kernel variants with identical images (rt_packet.cuh): the large-scene kernel exists at 128 registers
  // (2 CTAs/SM) and 80 registers (3 CTAs/SM), each with independent warps or CTA-coherent phases; which is
  // fastest depends on the scene (B200: C3 coherent, C5 independent). The choice per (scene shape, launch
  // shape) is made once — by rt_tune, or (autotune on, the default) on the launches after the first of a new
  // shape, each variant timed kTuneTrials times with events on the caller's stream — and kept in a small
  // table, so launches of different shapes (the bands of the BANDED host path) do not evict each other.
  int dense_override = -1;      // RT_PK_DENSE=0/1 pins the register budget
  int fan_override = -1;        // dev A/B: RT_PK_FAN=0 disables the analytic PCSS probe classification of the fast build
  int grid_override = -1;       // dev A/B: RT_PK_GRID=0 disables the light grids of large scenes
  // light grids (build_light_grids): cached while the uploaded bytes do not change (the reference re-uploads every frame)
  std::vector<unsigned char> grid_key; std::vector<float4> grid_data; int grid_off_f4 = 0, grid_ent_f4 = 0;
  size_t off_grid = 0, grid_f4s = 0;
  static constexpr int kTuneTrials = 2, kTuneVariants = 8, kTuneEntries = 8;
  struct Variant { int kernel, dense, coherent; };          // kernel: RT_KERNEL_*
  struct TuneEntry {
    long long key[10] = {-1, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    int nv = 0; Variant v[kTuneVariants] = {};
    int state = 0;                 // 0 new, 1..nv*kTuneTrials trials issued, then chosen
    int choice = 0; bool chosen = false;
    float ms[kTuneVariants] = {};  // best trial of each variant (0 = not measured)
    cudaEvent_t ev[kTuneVariants][kTuneTrials][2] = {};
    uint64_t last_use = 0;
  };
  std::vector<TuneEntry> tune;
  bool autotune = true;
  bool pinned = false; Variant pin = {0, 0, 0};
  int last_kernel = -1, last_dense = 0, last_coherent = 0;
  uint64_t tune_clock = 0;
  int regroup_override = -1;    // RT_PK_REGROUP=0 removes the regrouping megakernel from the variants, 1 pins it
  int wavefront_override = -1;  // RT_PK_WAVEFRONT=0 removes the wavefront pipeline from the variants, 1 pins it
  void* wf_state = nullptr;     // wavefront queues (rt_wavefront.cuh), sized on first use
  size_t wf_bytes = 0;
  cudaEvent_t wf_done = nullptr;
  void* static_scratch = nullptr;          // device address of g_static_scratch on this context's device (rt_create)
  std::string static_scratch_err;          // why it could not be resolved
  // RT_WF_CHUNK: pixels per wavefront chunk (queue memory ~ 300 B per pixel: 2.6 GB at the default, a whole 4K frame).
  // Measured on B200 (C3 4K / C5 8K, ms): 0.5 M 18.3 / 13.3, 1 M 15.5 / 10.4, 2 M 14.4 / 8.9, 4 M 13.9 / 8.1, 9 M 13.4 / 7.7 —
  // every chunk pays the ramp-up and tail of 2 D + 1 launches.
  size_t wf_chunk_pixels = (size_t)1 << 23;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t band_done[16] = {};
  cudaStream_t stream = nullptr;
  float4* o_color = nullptr; float4* o_position = nullptr; void* o_normal = nullptr; int* o_id = nullptr;
  size_t o_pixels = 0;
  bool tables_uploaded = false;
  void* post_state = nullptr;   // rt_post.cu
};

namespace {

rt_status fail(rt_context* c, rt_status s, const std::string& msg) {
  if (c) c->err = msg;
  return s;
}
rt_status cuda_fail(rt_context* c, cudaError_t e, const char* what) {
  rt_status s = (e == cudaErrorMemoryAllocation) ? RT_ERR_OOM : RT_ERR_CUDA;
  return fail(c, s, std::string(what) + ": " + cudaGetErrorString(e));
}
void release_tune_events(rt_context::TuneEntry& t) {
  for (auto& v : t.ev) for (auto& k : v) for (auto& e : k) if (e) { cudaEventDestroy(e); e = nullptr; }
}
}  // namespace

namespace rt {
int ctx_device(rt_context* c) { return c->device; }
rt_status ctx_fail(rt_context* c, rt_status s, const std::string& msg) { return fail(c, s, msg); }
void** ctx_post_slot(rt_context* c) { return &c->post_state; }
void ctx_count_launch(rt_context* c) { c->launches += 1; }
}  // namespace rt

namespace {
#define RT_CUDA(ctx, call)                                         \
  do {                                                             \
    cudaError_t e__ = (call);                                      \
    if (e__ != cudaSuccess) return cuda_fail((ctx), e__, #call);   \
  } while (0)

// ---- FP32 helpers evaluated exactly as the shader would (this TU is compiled with
// -ffp-contract=off for the host compiler, so no FMA forms).
struct h3 { float x, y, z; };
inline h3 hcross(h3 a, h3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline float hdot(h3 a, h3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline h3 hnormalize(h3 a) { float inv = 1.0f / std::sqrt(hdot(a, a)); return {a.x * inv, a.y * inv, a.z * inv}; }

float halton_host(int index, int base) {           // glsl:278-288, IEEE FP32
  float result = 0.0f, f = 1.0f / (float)base;
  for (int i = index; i > 0; i /= base) { result = result + f * (float)(i % base); f = f / (float)base; }
  return result;
}

// Object[]/Light[] (std430 bytes) -> the kernel's SoA blob. Layout in rt_device.h.
void stage_scene(const rt_object* obj, int n, const rt_light* lights, int m, int n_planes, float4* out,
                 size_t off_plane, size_t off_material, size_t off_lights) {
  float4* isect = out; float4* plane = out + off_plane; float4* mat = out + off_material; float4* lt = out + off_lights;
  int slot = 0;
  for (int i = 0; i < n; ++i) {
    const rt_object& o = obj[i];
    int this_slot = 0;
    if (o.type == 1) {
      this_slot = slot++;
      const h3 nrm = {o.normal[0], o.normal[1], o.normal[2]};
      h3 right, forward;                             // frame of glsl:130-138, hoisted per object
      if (std::fabs(nrm.y) > 0.9f) right = hnormalize(hcross(nrm, h3{0.f, 0.f, 1.f}));
      else right = hnormalize(hcross(nrm, h3{0.f, 1.f, 0.f}));
      forward = hnormalize(hcross(right, nrm));
      plane[3 * this_slot + 0] = make_float4(nrm.x, nrm.y, nrm.z, o.size[0] / 2.0f);
      plane[3 * this_slot + 1] = make_float4(right.x, right.y, right.z, o.size[1] / 2.0f);
      plane[3 * this_slot + 2] = make_float4(forward.x, forward.y, forward.z, 0.f);
    }
    float type_bits, slot_bits;
    std::memcpy(&type_bits, &o.type, 4);
    std::memcpy(&slot_bits, &this_slot, 4);
    isect[3 * i + 0] = make_float4(o.bounds.min[0], o.bounds.min[1], o.bounds.min[2], type_bits);
    isect[3 * i + 1] = make_float4(o.bounds.max[0], o.bounds.max[1], o.bounds.max[2], slot_bits);
    isect[3 * i + 2] = make_float4(o.position[0], o.position[1], o.position[2], o.type == 0 ? o.radius * o.radius : 0.f);
    const rt_material& a = o.material;
    mat[4 * i + 0] = make_float4(a.albedo[0], a.albedo[1], a.albedo[2], a.metallic);
    mat[4 * i + 1] = make_float4(a.roughness, a.diffuseStrength, a.ior, a.transparency);
    mat[4 * i + 2] = make_float4(a.subsurfaceColor[0], a.subsurfaceColor[1], a.subsurfaceColor[2], a.subsurfaceScatter);
    mat[4 * i + 3] = make_float4(a.scatterDistance, 0.f, 0.f, 0.f);
  }
  (void)n_planes;
  if (m > 0) std::memcpy(lt, lights, (size_t)m * sizeof(rt_light));
}

// Object[]/Light[] -> the packet kernel's shared-memory image (layout in rt_packet.cuh).
// Returns the size in float4 units, or 0 when the scene cannot use the sign-select loops
// (a non-finite bound). `out` may be null to size only.
struct PkLayout { size_t float4s; int spheres, planes, lights_offset, clusters, fan_ok, material_offset; };

inline unsigned morton10(unsigned v) {                   // spread 10 bits to every third position
  v &= 1023u;
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}

// Slot order of the spheres in the packet blob: array order, or Morton order of the centres when clustered.
std::vector<int> sphere_slot_order(const rt_object* obj, int n, bool clustered) {
  std::vector<int> order;
  for (int i = 0; i < n; ++i) if (obj[i].type == 0) order.push_back(i);
  if (clustered) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i : order) for (int k = 0; k < 3; ++k) {
      const float v = obj[i].position[k];
      if (v < lo[k]) lo[k] = v;
      if (v > hi[k]) hi[k] = v;
    }
    std::vector<std::pair<unsigned, int>> keyed;
    keyed.reserve(order.size());
    for (int i : order) {
      unsigned code = 0;
      for (int k = 0; k < 3; ++k) {
        const float ext = hi[k] - lo[k];
        float q = ext > 0.f ? (obj[i].position[k] - lo[k]) / ext * 1023.f : 0.f;
        if (!(q >= 0.f)) q = 0.f;                      // NaN centres sort first
        if (q > 1023.f) q = 1023.f;
        code |= morton10((unsigned)q) << k;
      }
      keyed.emplace_back(code, i);
    }
    std::sort(keyed.begin(), keyed.end());
    for (size_t j = 0; j < keyed.size(); ++j) order[j] = keyed[j].second;
  }
  return order;
}

// ---- light grids (large scenes): per light with a NARROW shadow cone (PCF only, sin <= 0.05), the sphere
// slots binned by where they can occlude, so that a shadow job looks up a handful of spheres instead of
// scanning every cluster (rt_packet.cuh scan_candidates). Conservative by construction; the job still
// runs its exact cone test and the exact per-ray tests on every entry, so results do not change.
//   point / area light at Lp: every ray of a job stays within 1.1*s*t of the axis o + t*L, t <= dist, and
//     every axis point is seen from the light in direction -L. A sphere (c, r) can be touched only if
//     the line from Lp along -L passes within R = r + 1.15*s*Dmax of c, i.e. if -L lies within
//     asin(R / |c - Lp|) of (c - Lp): binned over a cube map of directions around the light (8 x 8 cells
//     per face); spheres with |c - Lp| <= R go to a list every job reads.
//   directional light: the same bound around the axis line o + t*L, binned over a 32 x 32 grid of the plane
//     orthogonal to L (the cell of the origin's projection).
// Dmax bounds t: jobs farther from the light (point), or starting outside the scene's bounding ball
// (directional), take the cluster scan. Layout (float4 units from the start of the grid blob):
//   header, 4 float4 per light: (mode bits, s_grid, bound, first cell bits) (Lp | U, G | umin) (V, vmin) (centre, cells per unit)
//   offsets: u32 per cell (+ the shared cell + end), all lights back to back;  entries: u16 sphere slots.
inline float __int_as_float_host(int v) { float f; std::memcpy(&f, &v, 4); return f; }
struct GridBlob { std::vector<float4> data; int off_f4 = 0, ent_f4 = 0; };

inline int cube_cell(double x, double y, double z, int G) {    // == rt_packet.cuh grid_cube_cell
  const double ax = std::fabs(x), ay = std::fabs(y), az = std::fabs(z);
  int face; double a, b;
  if (ax >= ay && ax >= az) { face = x >= 0 ? 0 : 1; a = y / ax; b = z / ax; }
  else if (ay >= az)        { face = y >= 0 ? 2 : 3; a = x / ay; b = z / ay; }
  else                      { face = z >= 0 ? 4 : 5; a = x / az; b = y / az; }
  int ix = (int)std::floor((a * 0.5 + 0.5) * G), iy = (int)std::floor((b * 0.5 + 0.5) * G);
  ix = ix < 0 ? 0 : (ix > G - 1 ? G - 1 : ix); iy = iy < 0 ? 0 : (iy > G - 1 ? G - 1 : iy);
  return (face * G + iy) * G + ix;
}
inline void cube_dir(int face, double a, double b, double d[3]) {
  switch (face) {
    case 0: d[0] = 1; d[1] = a; d[2] = b; break;
    case 1: d[0] = -1; d[1] = a; d[2] = b; break;
    case 2: d[0] = a; d[1] = 1; d[2] = b; break;
    case 3: d[0] = a; d[1] = -1; d[2] = b; break;
    case 4: d[0] = a; d[1] = b; d[2] = 1; break;
    default: d[0] = a; d[1] = b; d[2] = -1; break;
  }
  const double inv = 1.0 / std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  d[0] *= inv; d[1] *= inv; d[2] *= inv;
}

constexpr int kGridCube = 8, kGridOrtho = 32;
constexpr float kGridMaxSin = 0.05f;

GridBlob build_light_grids(const rt_object* obj, int n, const rt_light* lights, int m) {
  GridBlob g;
  int spheres = 0;
  for (int i = 0; i < n; ++i) spheres += obj[i].type == 0;
  if (spheres < rt::kPkClusterMinSpheres || spheres > 65535 || m <= 0 || m > 64) return g;
  const std::vector<int> order = sphere_slot_order(obj, n, true);
  // bounding ball of everything a shading point can lie on (spheres and plane rectangles)
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (int i = 0; i < n; ++i) {
    if (obj[i].type != 0 && obj[i].type != 1) continue;
    for (int k = 0; k < 3; ++k) {
      if (!std::isfinite(obj[i].position[k])) return g;
      lo[k] = std::min(lo[k], (double)obj[i].position[k]); hi[k] = std::max(hi[k], (double)obj[i].position[k]);
    }
  }
  const double ctr[3] = {0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1]), 0.5 * (lo[2] + hi[2])};
  double Rall = 0.0;
  for (int i = 0; i < n; ++i) {
    if (obj[i].type != 0 && obj[i].type != 1) continue;
    const double ext = obj[i].type == 0 ? std::fabs((double)obj[i].radius)
                                        : 0.5 * std::sqrt((double)obj[i].size[0] * obj[i].size[0] + (double)obj[i].size[1] * obj[i].size[1]);
    if (!std::isfinite(ext)) return g;
    const double dx = obj[i].position[0] - ctr[0], dy = obj[i].position[1] - ctr[1], dz = obj[i].position[2] - ctr[2];
    Rall = std::max(Rall, std::sqrt(dx * dx + dy * dy + dz * dz) + ext);
  }
  Rall = Rall * 1.001 + 0.01;                            // + the 0.001 * N offset of a shadow origin
  // static tables of the cube map: cell centre directions and angular radii (initialised once, thread-safe)
  struct CubeTables { double dir[6 * kGridCube * kGridCube][3]; double rad[6 * kGridCube * kGridCube]; };
  static const CubeTables* const tables = [] {
    CubeTables* t = new CubeTables;
    const int G = kGridCube;
    for (int f = 0; f < 6; ++f) for (int iy = 0; iy < G; ++iy) for (int ix = 0; ix < G; ++ix) {
      const int cell = (f * G + iy) * G + ix;
      const double a = (ix + 0.5) / G * 2 - 1, b = (iy + 0.5) / G * 2 - 1, h = 1.0 / G;
      cube_dir(f, a, b, t->dir[cell]);
      double worst = 1.0;
      for (int c4 = 0; c4 < 4; ++c4) {
        double d[3];
        cube_dir(f, a + ((c4 & 1) ? h : -h), b + ((c4 & 2) ? h : -h), d);
        worst = std::min(worst, d[0] * t->dir[cell][0] + d[1] * t->dir[cell][1] + d[2] * t->dir[cell][2]);
      }
      t->rad[cell] = std::acos(std::max(-1.0, std::min(1.0, worst)));
    }
    return t;
  }();
  const double (*cdir)[3] = tables->dir;
  const double* crad = tables->rad;
  std::vector<float4> hdr((size_t)4 * m, make_float4(0.f, 0.f, 0.f, 0.f));
  std::vector<unsigned> offsets;
  std::vector<unsigned short> entries;
  bool any = false;
  for (int j = 0; j < m; ++j) {
    const rt_light& L = lights[j];
    if (L.shadowType != 1) continue;                     // PCF only: PCSS probes open the cone to tens of degrees
    // Cone::s of such a job (rt_packet.cuh shadow_packets), a hair wider so that the device's `cone.s <= s_grid` holds whatever the last ulp
    const float s_grid = (2.0f * std::fabs(L.shadowSoftness * 0.005f) * 1.01f + 1.0e-5f) * 1.0001f;
    if (!(s_grid <= kGridMaxSin)) continue;
    std::vector<std::vector<unsigned short>> cells;
    int mode = 0;
    float4 h1 = make_float4(0, 0, 0, 0), h2 = h1;
    float bound = 0.f, per_unit = 0.f;
    if (L.type == 0 || L.type == 2) {
      const double Lp[3] = {L.position[0], L.position[1], L.position[2]};
      if (!std::isfinite(Lp[0]) || !std::isfinite(Lp[1]) || !std::isfinite(Lp[2])) continue;
      const double dl = std::sqrt((Lp[0] - ctr[0]) * (Lp[0] - ctr[0]) + (Lp[1] - ctr[1]) * (Lp[1] - ctr[1]) + (Lp[2] - ctr[2]) * (Lp[2] - ctr[2]));
      const double Dmax = dl + Rall;
      const int ncell = 6 * kGridCube * kGridCube;
      cells.assign(ncell + 1, {});
      for (int ks = 0; ks < (int)order.size(); ++ks) {
        const rt_object& o = obj[order[ks]];
        const double u[3] = {o.position[0] - Lp[0], o.position[1] - Lp[1], o.position[2] - Lp[2]};
        const double len = std::sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
        const double R = std::fabs((double)o.radius) * 1.001 + 1.15 * s_grid * Dmax + 1e-4 * (len + 1.0);
        if (!(len > R * 1.01) || !std::isfinite(R)) { cells[ncell].push_back((unsigned short)ks); continue; }
        const double ang = std::asin(R / len) + 2e-3;
        for (int f = 0; f < 6; ++f) {
          // a face spans 54.74 degrees around its axis: skip the faces the sphere's disc cannot reach
          const double axis_dot = (f & 1 ? -1.0 : 1.0) * u[f >> 1];
          if (axis_dot < std::cos(std::min(3.14159265358979, ang + 0.9554)) * len) continue;
          for (int c = f * kGridCube * kGridCube; c < (f + 1) * kGridCube * kGridCube; ++c) {
            const double lim = std::min(3.14159265358979, ang + crad[c]);
            if ((u[0] * cdir[c][0] + u[1] * cdir[c][1] + u[2] * cdir[c][2]) >= std::cos(lim) * len) cells[c].push_back((unsigned short)ks);
          }
        }
      }
      mode = 1; bound = (float)(Dmax * 0.999);
      h1 = make_float4((float)Lp[0], (float)Lp[1], (float)Lp[2], (float)kGridCube);
    } else if (L.type == 1) {
      double D[3] = {-(double)L.direction[0], -(double)L.direction[1], -(double)L.direction[2]};   // lightDir = normalize(-direction)
      const double dn = std::sqrt(D[0] * D[0] + D[1] * D[1] + D[2] * D[2]);
      if (!(dn > 1e-20) || !std::isfinite(dn)) continue;
      for (double& v : D) v /= dn;
      double T[3] = {0, 0, 0};
      T[std::fabs(D[0]) < 0.6 ? 0 : (std::fabs(D[1]) < 0.6 ? 1 : 2)] = 1.0;
      double U[3] = {D[1] * T[2] - D[2] * T[1], D[2] * T[0] - D[0] * T[2], D[0] * T[1] - D[1] * T[0]};
      const double un = std::sqrt(U[0] * U[0] + U[1] * U[1] + U[2] * U[2]);
      for (double& v : U) v /= un;
      const double V[3] = {D[1] * U[2] - D[2] * U[1], D[2] * U[0] - D[0] * U[2], D[0] * U[1] - D[1] * U[0]};
      const double Dmax = 2.1 * Rall, infl = 1.15 * s_grid * Dmax;
      const double cu = ctr[0] * U[0] + ctr[1] * U[1] + ctr[2] * U[2], cv = ctr[0] * V[0] + ctr[1] * V[1] + ctr[2] * V[2];
      const double umin = cu - Rall, vmin = cv - Rall, cell = 2.0 * Rall / kGridOrtho;      // origins lie inside the ball
      const int ncell = kGridOrtho * kGridOrtho;
      cells.assign(ncell + 1, {});
      for (int ks = 0; ks < (int)order.size(); ++ks) {
        const rt_object& o = obj[order[ks]];
        const double pu = o.position[0] * U[0] + o.position[1] * U[1] + o.position[2] * U[2];
        const double pv = o.position[0] * V[0] + o.position[1] * V[1] + o.position[2] * V[2];
        const double R = std::fabs((double)o.radius) * 1.001 + infl + 1e-4 * (Rall + 1.0) + 1e-3 * cell;
        if (!std::isfinite(R) || !std::isfinite(pu) || !std::isfinite(pv)) { cells[ncell].push_back((unsigned short)ks); continue; }
        int x0 = (int)std::floor((pu - R - umin) / cell), x1 = (int)std::floor((pu + R - umin) / cell);
        int y0 = (int)std::floor((pv - R - vmin) / cell), y1 = (int)std::floor((pv + R - vmin) / cell);
        x0 = std::max(x0, 0); y0 = std::max(y0, 0); x1 = std::min(x1, kGridOrtho - 1); y1 = std::min(y1, kGridOrtho - 1);
        for (int y = y0; y <= y1; ++y) for (int x = x0; x <= x1; ++x) cells[(size_t)y * kGridOrtho + x].push_back((unsigned short)ks);
      }
      mode = 2; bound = (float)(Rall * Rall * 0.998);
      h1 = make_float4((float)U[0], (float)U[1], (float)U[2], (float)umin);
      h2 = make_float4((float)V[0], (float)V[1], (float)V[2], (float)vmin);
      per_unit = (float)(1.0 / cell);
    } else continue;
    float4 h0;
    h0.x = __int_as_float_host(mode); h0.y = s_grid; h0.z = bound; h0.w = __int_as_float_host((int)offsets.size());
    hdr[4 * j + 0] = h0; hdr[4 * j + 1] = h1; hdr[4 * j + 2] = h2;
    hdr[4 * j + 3] = make_float4((float)ctr[0], (float)ctr[1], (float)ctr[2], per_unit);
    for (const auto& cl : cells) { offsets.push_back((unsigned)entries.size()); entries.insert(entries.end(), cl.begin(), cl.end()); }
    offsets.push_back((unsigned)entries.size());
    any = true;
  }
  if (!any) return g;
  g.off_f4 = (int)hdr.size();
  const size_t off_f4s = (offsets.size() * 4 + 15) / 16, ent_f4s = (entries.size() * 2 + 15) / 16 + 1;
  g.ent_f4 = g.off_f4 + (int)off_f4s;
  g.data.assign(hdr.size() + off_f4s + ent_f4s, make_float4(0.f, 0.f, 0.f, 0.f));
  std::memcpy(g.data.data(), hdr.data(), hdr.size() * 16);
  std::memcpy(g.data.data() + g.off_f4, offsets.data(), offsets.size() * 4);
  if (!entries.empty()) std::memcpy(g.data.data() + g.ent_f4, entries.data(), entries.size() * 2);
  return g;
}
PkLayout stage_scene_packet(const rt_object* obj, int n, const rt_light* lights, int m, float4* out) {
  PkLayout L = {0, 0, 0, 0, 0, 1, -1};
  for (int i = 0; i < n; ++i) {
    if (obj[i].type == 0) L.spheres++; else if (obj[i].type == 1) L.planes++;
    if (obj[i].type == 0 || obj[i].type == 1)
      for (int k = 0; k < 3; ++k)
        if (!std::isfinite(obj[i].bounds.min[k]) || !std::isfinite(obj[i].bounds.max[k])) return PkLayout{0, 0, 0, 0, 0, 0, -1};
    // fan_ok: every sphere's box contains the sphere (what GenerateAABBForObject produces). The
    // fast build's analytic PCSS probe classification (rt_packet.cuh probe_fan) relies on "a ray
    // that hits the sphere with margin also passes its box"; any other box sends probes down the
    // exact per-ray path.
    if (obj[i].type == 0) {
      const float r = std::fabs(obj[i].radius);
      if (!(r > 0.0f) || !std::isfinite(r)) L.fan_ok = 0;
      for (int k = 0; k < 3; ++k) {
        const float lo = std::fmin(obj[i].bounds.min[k], obj[i].bounds.max[k]), hi = std::fmax(obj[i].bounds.min[k], obj[i].bounds.max[k]);
        const float c = obj[i].position[k];
        if (!(lo <= c - r && hi >= c + r)) L.fan_ok = 0;
      }
    }
  }
  // Sphere clusters (rt_packet.cuh "two-level boxes"): from kPkClusterMinSpheres spheres on, the
  // sphere slots are ordered along a Morton curve of their centres and every kPkClusterSize
  // consecutive slots get a cluster record (union box + bounding sphere).
  L.clusters = L.spheres >= rt::kPkClusterMinSpheres ? (L.spheres + rt::kPkClusterSize - 1) / rt::kPkClusterSize : 0;
  size_t bytes = (size_t)L.clusters * 64 + (size_t)L.spheres * 64 + (size_t)L.planes * 128 + (size_t)L.spheres * 4;
  bytes = (bytes + 15) / 16 * 16;
  L.lights_offset = (int)bytes;
  bytes += (size_t)m * sizeof(rt_light);
  // small scenes: the material records (4 float4 per ORIGINAL object index, as stage_scene lays them out)
  // ride in the shared-memory blob too, so a closest hit reads them with LDS instead of 4 global loads
  if (L.spheres + L.planes <= 48 && n <= 64) { L.material_offset = (int)bytes; bytes += (size_t)n * 64; }
  L.float4s = bytes / 16;
  if (L.float4s == 0) L.float4s = 1;
  if (!out) return L;
  std::memset(out, 0, L.float4s * 16);
  char* base = (char*)out;
  float* clu = (float*)base;
  float* sph = (float*)(base + (size_t)L.clusters * 64);
  float* pln = sph + (size_t)L.spheres * 16;
  int* sph_index = (int*)(pln + (size_t)L.planes * 32);
  auto box = [](float* r, const rt_aabb& b) {          // per axis (lo,hi,hi,lo), lo <= hi
    for (int k = 0; k < 3; ++k) {
      const float lo = b.max[k] < b.min[k] ? b.max[k] : b.min[k];
      const float hi = b.max[k] < b.min[k] ? b.min[k] : b.max[k];
      r[4 * k + 0] = lo; r[4 * k + 1] = hi; r[4 * k + 2] = hi; r[4 * k + 3] = lo;
    }
  };
  const std::vector<int> order = sphere_slot_order(obj, n, L.clusters > 0);
  for (int ks = 0; ks < L.spheres; ++ks) {
    const rt_object& o = obj[order[ks]];
    float* r = sph + (size_t)ks * 16;
    box(r, o.bounds);
    r[12] = o.position[0]; r[13] = o.position[1]; r[14] = o.position[2]; r[15] = o.radius;
    sph_index[ks] = order[ks];
  }
  for (int c = 0; c < L.clusters; ++c) {
    const int first = c * rt::kPkClusterSize, last = std::min(L.spheres, first + rt::kPkClusterSize);
    float* r = clu + (size_t)c * 16;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int ks = first; ks < last; ++ks) for (int k = 0; k < 3; ++k) {
      const float* mr = sph + (size_t)ks * 16 + 4 * k;   // member (lo,hi,...)
      if (mr[0] < lo[k]) lo[k] = mr[0];
      if (mr[1] > hi[k]) hi[k] = mr[1];
    }
    for (int k = 0; k < 3; ++k) { r[4 * k + 0] = lo[k]; r[4 * k + 1] = hi[k]; r[4 * k + 2] = hi[k]; r[4 * k + 3] = lo[k]; }
    // bounding sphere of the member spheres (double precision, then inflated) for the cone test
    double cx = 0.5 * ((double)lo[0] + hi[0]), cy = 0.5 * ((double)lo[1] + hi[1]), cz = 0.5 * ((double)lo[2] + hi[2]), R = 0.0;
    for (int ks = first; ks < last; ++ks) {
      const float* mr = sph + (size_t)ks * 16;
      const double dx = mr[12] - cx, dy = mr[13] - cy, dz = mr[14] - cz;
      const double d = std::sqrt(dx * dx + dy * dy + dz * dz) + std::fabs((double)mr[15]);
      if (!(d <= R)) R = d;                              // NaN radius/centre -> NaN bound -> never culled
    }
    r[12] = (float)cx; r[13] = (float)cy; r[14] = (float)cz; r[15] = (float)(R * 1.001 + 1e-6);
  }
  int kp = 0;
  for (int i = 0; i < n; ++i) {
    const rt_object& o = obj[i];
    if (o.type != 1) continue;
    float* r = pln + (size_t)kp * 32;
    box(r, o.bounds);
    const h3 nrm = {o.normal[0], o.normal[1], o.normal[2]};
    h3 right, forward;                               // frame of glsl:130-138, hoisted per object
    if (std::fabs(nrm.y) > 0.9f) right = hnormalize(hcross(nrm, h3{0.f, 0.f, 1.f}));
    else right = hnormalize(hcross(nrm, h3{0.f, 1.f, 0.f}));
    forward = hnormalize(hcross(right, nrm));
    r[12] = nrm.x; r[13] = nrm.y; r[14] = nrm.z; r[15] = o.size[0] / 2.0f;
    r[16] = right.x; r[17] = right.y; r[18] = right.z; r[19] = o.size[1] / 2.0f;
    r[20] = forward.x; r[21] = forward.y; r[22] = forward.z; r[23] = 0.f;
    r[24] = o.position[0]; r[25] = o.position[1]; r[26] = o.position[2];
    std::memcpy(&r[27], &i, 4);
    // conservative bounds for the shadow-job cone test: rectangle bounding radius, |normal|
    r[28] = std::sqrt(r[15] * r[15] + r[19] * r[19]) * 1.0001f; r[29] = std::sqrt(hdot(nrm, nrm)) * 1.0001f;
    kp++;
  }
  if (m > 0) std::memcpy(base + L.lights_offset, lights, (size_t)m * sizeof(rt_light));
  if (L.material_offset >= 0) {
    float4* mat = (float4*)(base + L.material_offset);
    for (int i = 0; i < n; ++i) {
      const rt_material& a = obj[i].material;
      mat[4 * i + 0] = make_float4(a.albedo[0], a.albedo[1], a.albedo[2], a.metallic);
      mat[4 * i + 1] = make_float4(a.roughness, a.diffuseStrength, a.ior, a.transparency);
      mat[4 * i + 2] = make_float4(a.subsurfaceColor[0], a.subsurfaceColor[1], a.subsurfaceColor[2], a.subsurfaceScatter);
      mat[4 * i + 3] = make_float4(a.scatterDistance, 0.f, 0.f, 0.f);
    }
  }
  return L;
}

rt_status ensure_outputs(rt_context* c, size_t pixels) {
  if (pixels <= c->o_pixels) return RT_OK;
  cudaFree(c->o_color); cudaFree(c->o_position); cudaFree(c->o_normal); cudaFree(c->o_id);
  c->o_color = c->o_position = nullptr; c->o_normal = nullptr; c->o_id = nullptr; c->o_pixels = 0;
  RT_CUDA(c, cudaMalloc(&c->o_color, pixels * sizeof(float4)));
  RT_CUDA(c, cudaMalloc(&c->o_position, pixels * sizeof(float4)));
  RT_CUDA(c, cudaMalloc(&c->o_normal, pixels * sizeof(float4)));
  RT_CUDA(c, cudaMalloc(&c->o_id, pixels * sizeof(int)));
  c->o_pixels = pixels;
  return RT_OK;
}

}  // namespace

extern "C" {

int rt_abi_version(void) { return RT_ABI_VERSION; }

const char* rt_status_string(rt_status s) {
  switch (s) {
    case RT_OK: return "RT_OK";
    case RT_ERR_INVALID_ARG: return "RT_ERR_INVALID_ARG";
    case RT_ERR_CUDA: return "RT_ERR_CUDA";
    case RT_ERR_NCCL: return "RT_ERR_NCCL";
    case RT_ERR_OOM: return "RT_ERR_OOM";
    case RT_ERR_UNAVAILABLE: return "RT_ERR_UNAVAILABLE";
  }
  return "RT_ERR_?";
}

const char* rt_last_error(rt_context* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

rt_status rt_create(rt_context** out, int cuda_device) {
  if (!out) return RT_ERR_INVALID_ARG;
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) return RT_ERR_UNAVAILABLE;   // no GPU: there is no CPU fallback
  if (cuda_device < 0 || cuda_device >= count) return RT_ERR_INVALID_ARG;
  rt_context* c = new rt_context();
  c->device = cuda_device;
  auto bail = [&](cudaError_t err) { rt_status s = cuda_fail(c, err, "rt_create"); fprintf(stderr, "rt_create: %s\n", c->err.c_str()); delete c; return s; };
  if ((e = cudaSetDevice(cuda_device)) != cudaSuccess) return bail(e);
  if ((e = cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, cuda_device)) != cudaSuccess) return bail(e);
  if ((e = cudaDeviceGetAttribute(&c->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, cuda_device)) != cudaSuccess) return bail(e);
  if ((e = cudaMalloc(&c->d_tile_counter, sizeof(unsigned) * 32 * rt_context::kCounterSlots)) != cudaSuccess) return bail(e);
  if ((e = cudaMemset(c->d_tile_counter, 0, sizeof(unsigned) * 32 * rt_context::kCounterSlots)) != cudaSuccess) return bail(e);
  for (auto& sl : c->slots)
    if ((e = cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming)) != cudaSuccess) return bail(e);
  if ((e = cudaMalloc(&c->d_counters, sizeof(unsigned long long) * RT_CNT_COUNT)) != cudaSuccess) return bail(e);
  if ((e = cudaMemset(c->d_counters, 0, sizeof(unsigned long long) * RT_CNT_COUNT)) != cudaSuccess) return bail(e);
  if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e);
  if ((e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e);
  {                                                      // set-up of rt_render_static: resolve the scratch array on THIS device
    void* sp = nullptr;
    const cudaError_t es = cudaGetSymbolAddress(&sp, g_static_scratch);
    if (es == cudaSuccess) c->static_scratch = sp;
    else { c->static_scratch_err = cudaGetErrorString(es); cudaGetLastError(); }
  }
  for (int i = 0; i < 16; ++i)
    if ((e = cudaEventCreateWithFlags(&c->band_done[i], cudaEventDisableTiming)) != cudaSuccess) return bail(e);
  if ((e = cudaEventCreateWithFlags(&c->stage_done, cudaEventDisableTiming)) != cudaSuccess) return bail(e);
  float h2[64], h3t[64];
  for (int i = 0; i < 64; ++i) { h2[i] = halton_host(i, 2); h3t[i] = halton_host(i, 3); }
  int order[16];                                       // PCSS probes sorted by offset (rt_packet.cuh)
  for (int i = 0; i < 16; ++i) order[i] = i;
  for (int i = 1; i < 16; ++i)
    for (int j = i; j > 0 && h3t[order[j]] < h3t[order[j - 1]]; --j) { int t = order[j]; order[j] = order[j - 1]; order[j - 1] = t; }
  if ((e = rt::upload_tables_fast(h2, h3t, order)) != cudaSuccess) return bail(e);
  if ((e = rt::upload_tables_parity(h2, h3t, order)) != cudaSuccess) return bail(e);
  if ((e = rt::upload_tables_fast_small(h2, h3t, order)) != cudaSuccess || (e = rt::upload_tables_fast_big(h2, h3t, order)) != cudaSuccess ||
      (e = rt::upload_tables_parity_small(h2, h3t, order)) != cudaSuccess || (e = rt::upload_tables_parity_big(h2, h3t, order)) != cudaSuccess) return bail(e);
  if (const char* e2 = getenv("RT_PK_COHERENT")) c->coherent_override = atoi(e2);
  if (const char* e3 = getenv("RT_PK_DENSE")) c->dense_override = atoi(e3);
  if (const char* e4 = getenv("RT_PK_FAN")) c->fan_override = atoi(e4);
  if (const char* e5 = getenv("RT_PK_GRID")) c->grid_override = atoi(e5);
  if (const char* e9 = getenv("RT_PK_REGROUP")) c->regroup_override = atoi(e9);
  if (const char* e6 = getenv("RT_PK_WAVEFRONT")) c->wavefront_override = atoi(e6);
  if (const char* e8 = getenv("RT_WF_CHUNK")) c->wf_chunk_pixels = (size_t)atoll(e8) > 0 ? (size_t)atoll(e8) : c->wf_chunk_pixels;
  if (const char* e7 = getenv("RT_AUTOTUNE")) c->autotune = atoi(e7) != 0;
  *out = c;
  return RT_OK;
}

void rt_destroy(rt_context* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  rt::post_release(c->post_state);
  for (auto& t : c->tune) release_tune_events(t);
  for (auto& sl : c->slots) if (sl.done) cudaEventDestroy(sl.done);
  if (c->sky_tex) cudaDestroyTextureObject(c->sky_tex);
  if (c->sky_array) cudaFreeArray(c->sky_array);
  cudaFree(c->d_sky); cudaFree(c->d_noise); cudaFree(c->d_scene);
  cudaFree(c->d_tile_counter); cudaFree(c->d_counters); cudaFree(c->wf_state);
  if (c->wf_done) cudaEventDestroy(c->wf_done);
  cudaFree(c->o_color); cudaFree(c->o_position); cudaFree(c->o_normal); cudaFree(c->o_id);
  if (c->h_stage) cudaFreeHost(c->h_stage);
  if (c->stage_done) cudaEventDestroy(c->stage_done);
  if (c->stream) cudaStreamDestroy(c->stream);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  for (int i = 0; i < 16; ++i) if (c->band_done[i]) cudaEventDestroy(c->band_done[i]);
  delete c;
}

rt_status rt_upload_scene(rt_context* c, const void* objects, int n, const void* lights, int m,
                          void* cuda_stream, int async) {
  if (!c) return RT_ERR_INVALID_ARG;
  if (n < 0 || m < 0 || (n > 0 && !objects) || (m > 0 && !lights)) return fail(c, RT_ERR_INVALID_ARG, "rt_upload_scene: bad counts or null arrays");
  RT_CUDA(c, cudaSetDevice(c->device));
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  const rt_object* obj = (const rt_object*)objects;
  int planes = 0;
  for (int i = 0; i < n; ++i) planes += (obj[i].type == 1);
  const size_t off_plane = (size_t)3 * n, off_material = off_plane + (size_t)3 * planes;
  const size_t off_lights = off_material + (size_t)4 * n, off_pk = off_lights + (size_t)6 * m;
  const PkLayout pkl = stage_scene_packet(obj, n, (const rt_light*)lights, m, nullptr);
  // light grids of large scenes, rebuilt only when the bytes change
  if (pkl.clusters > 0 && c->grid_override != 0) {
    const size_t ob = (size_t)n * sizeof(rt_object), lb = (size_t)m * sizeof(rt_light);
    const bool same = c->grid_key.size() == ob + lb + 8 && std::memcmp(c->grid_key.data() + 8, objects, ob) == 0 &&
                      (lb == 0 || std::memcmp(c->grid_key.data() + 8 + ob, lights, lb) == 0) &&
                      std::memcmp(c->grid_key.data(), &n, 4) == 0 && std::memcmp(c->grid_key.data() + 4, &m, 4) == 0;
    if (!same) {
      GridBlob g = build_light_grids(obj, n, (const rt_light*)lights, m);
      c->grid_data.swap(g.data); c->grid_off_f4 = g.off_f4; c->grid_ent_f4 = g.ent_f4;
      c->grid_key.resize(ob + lb + 8);
      std::memcpy(c->grid_key.data(), &n, 4); std::memcpy(c->grid_key.data() + 4, &m, 4);
      if (ob) std::memcpy(c->grid_key.data() + 8, objects, ob);
      if (lb) std::memcpy(c->grid_key.data() + 8 + ob, lights, lb);
    }
  } else {
    c->grid_data.clear(); c->grid_key.clear();
  }
  const size_t off_grid = off_pk + pkl.float4s;
  const size_t total = off_grid + c->grid_data.size();
  const size_t need = (total > 0 ? total : 1) * sizeof(float4);
  if (c->stage_pending) { RT_CUDA(c, cudaEventSynchronize(c->stage_done)); c->stage_pending = false; }
  if (need > c->h_stage_cap) {
    if (c->h_stage) cudaFreeHost(c->h_stage);
    c->h_stage = nullptr; c->h_stage_cap = 0;
    RT_CUDA(c, cudaHostAlloc((void**)&c->h_stage, need * 2, cudaHostAllocDefault));
    c->h_stage_cap = need * 2;
  }
  if (need > c->d_scene_cap) {
    RT_CUDA(c, cudaStreamSynchronize(stream));
    cudaFree(c->d_scene); c->d_scene = nullptr; c->d_scene_cap = 0;
    RT_CUDA(c, cudaMalloc((void**)&c->d_scene, need * 2));
    c->d_scene_cap = need * 2;
  }
  stage_scene(obj, n, (const rt_light*)lights, m, planes, c->h_stage, off_plane, off_material, off_lights);
  if (pkl.float4s) stage_scene_packet(obj, n, (const rt_light*)lights, m, c->h_stage + off_pk);
  if (!c->grid_data.empty()) std::memcpy(c->h_stage + off_grid, c->grid_data.data(), c->grid_data.size() * sizeof(float4));
  c->off_grid = off_grid; c->grid_f4s = c->grid_data.size();
  RT_CUDA(c, cudaMemcpyAsync(c->d_scene, c->h_stage, total * sizeof(float4), cudaMemcpyHostToDevice, stream));
  c->n_obj = n; c->n_planes = planes; c->n_lights = m;
  c->off_plane = off_plane; c->off_material = off_material; c->off_lights = off_lights;
  c->off_pk = off_pk; c->pk_float4s = pkl.float4s; c->pk_spheres = pkl.spheres; c->pk_planes = pkl.planes;
  c->pk_lights_offset = pkl.lights_offset; c->pk_clusters = pkl.clusters; c->pk_material_offset = pkl.material_offset;
  bool has_pcss = false;                                 // the analytic probe path only pays where a PCSS light exists
  for (int j = 0; j < m; ++j) has_pcss = has_pcss || ((const rt_light*)lights)[j].shadowType == 2;
  c->pk_fan_ok = (pkl.fan_ok && has_pcss) ? 1 : 0;
  if (async) {
    RT_CUDA(c, cudaEventRecord(c->stage_done, stream));
    c->stage_pending = true;
  } else {
    RT_CUDA(c, cudaStreamSynchronize(stream));
  }
  return RT_OK;
}

rt_status rt_set_skybox(rt_context* c, const void* faces, int face_size, int channels, int is_fp16) {
  if (!c) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  RT_CUDA(c, cudaDeviceSynchronize());
  if (c->sky_tex) { cudaDestroyTextureObject(c->sky_tex); c->sky_tex = 0; }
  if (c->sky_array) { cudaFreeArray(c->sky_array); c->sky_array = nullptr; }
  cudaFree(c->d_sky); c->d_sky = nullptr; c->sky_size = 0;
  if (!faces) return RT_OK;
  if (face_size <= 0 || face_size > 16384 || (channels != 3 && channels != 4)) return fail(c, RT_ERR_INVALID_ARG, "rt_set_skybox: face_size/channels");
  const size_t texels = (size_t)6 * face_size * face_size;
  std::vector<__half> rgba(texels * 4);
  for (size_t t = 0; t < texels; ++t) {
    for (int ch = 0; ch < 4; ++ch) {
      __half v = __float2half_rn(1.0f);
      if (ch < channels) v = is_fp16 ? ((const __half*)faces)[t * channels + ch] : __float2half_rn(((const float*)faces)[t * channels + ch]);
      rgba[t * 4 + ch] = v;
    }
  }
  RT_CUDA(c, cudaMalloc((void**)&c->d_sky, texels * sizeof(uint2)));
  RT_CUDA(c, cudaMemcpy(c->d_sky, rgba.data(), texels * sizeof(uint2), cudaMemcpyHostToDevice));
  c->sky_size = face_size;
  // Cubemap texture object for the RT_FLAG_SKYBOX_HW miss path (pure CUDA, no graphics API).
  cudaChannelFormatDesc desc = cudaCreateChannelDescHalf4();
  cudaError_t e = cudaMalloc3DArray(&c->sky_array, &desc, make_cudaExtent(face_size, face_size, 6), cudaArrayCubemap);
  if (e == cudaSuccess) {
    cudaMemcpy3DParms cp = {};
    cp.srcPtr = make_cudaPitchedPtr(rgba.data(), (size_t)face_size * sizeof(uint2), face_size, face_size);
    cp.dstArray = c->sky_array;
    cp.extent = make_cudaExtent(face_size, face_size, 6);
    cp.kind = cudaMemcpyHostToDevice;
    e = cudaMemcpy3D(&cp);
  }
  if (e == cudaSuccess) {
    cudaResourceDesc rd = {};
    rd.resType = cudaResourceTypeArray; rd.res.array.array = c->sky_array;
    cudaTextureDesc td = {};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 1;
    e = cudaCreateTextureObject(&c->sky_tex, &rd, &td, nullptr);
  }
  if (e != cudaSuccess) {   // the exact (manual) miss path does not need the texture object
    cudaGetLastError();
    c->sky_tex = 0;
    c->err = std::string("cubemap texture object unavailable: ") + cudaGetErrorString(e);
  }
  return RT_OK;
}

rt_status rt_set_noise(rt_context* c, const unsigned char* r8, int w, int h) {
  if (!c) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  RT_CUDA(c, cudaDeviceSynchronize());
  cudaFree(c->d_noise); c->d_noise = nullptr; c->noise_w = c->noise_h = 0;
  if (!r8) return RT_OK;
  if (w <= 0 || h <= 0) return fail(c, RT_ERR_INVALID_ARG, "rt_set_noise: size");
  RT_CUDA(c, cudaMalloc((void**)&c->d_noise, (size_t)w * h));
  RT_CUDA(c, cudaMemcpy(c->d_noise, r8, (size_t)w * h, cudaMemcpyHostToDevice));
  c->noise_w = w; c->noise_h = h;
  return RT_OK;
}

// ---- variant table (rt_context::TuneEntry)
static void fill_tuning(const rt_context* c, const rt_context::TuneEntry* t, rt_context::Variant run, int state, rt_tuning* out) {
  std::memset(out, 0, sizeof(*out));
  out->kernel = run.kernel; out->registers = run.dense; out->phase_level = run.coherent; out->state = state;
  if (!t) return;
  out->num_variants = t->nv;
  for (int i = 0; i < t->nv && i < 8; ++i) {
    out->variant_ms[i] = t->ms[i]; out->variant_kernel[i] = t->v[i].kernel;
    out->variant_registers[i] = t->v[i].dense; out->variant_phase_level[i] = t->v[i].coherent;
  }
  (void)c;
}

// Everything rt_render decides before it launches: validated arguments -> KernelParams, whether the packet
// kernel can take the scene, the variants that apply, and the table entry of this launch shape.
struct LaunchPlan {
  rt::KernelParams kp;
  bool use_packet = false, par = false, skip = false;
  int nv = 0; rt_context::Variant v[rt_context::kTuneVariants];
  long long key[10];
};

static rt_status plan_launch(rt_context* c, const rt_frame_params* p, LaunchPlan& L) {
  if (p->width <= 0 || p->height <= 0 || p->row_begin < 0 || p->row_end > p->height || p->row_begin > p->row_end)
    return fail(c, RT_ERR_INVALID_ARG, "rt_render: image size / row window");
  if (p->num_objects < 0 || p->num_objects > c->n_obj || p->num_lights < 0 || p->num_lights > c->n_lights)
    return fail(c, RT_ERR_INVALID_ARG, "rt_render: num_objects/num_lights exceed the uploaded scene");
  if (p->max_ray_depth < 0 || p->num_samples < 0) return fail(c, RT_ERR_INVALID_ARG, "rt_render: depth/samples");
  if (p->use_skybox && !c->d_sky) return fail(c, RT_ERR_INVALID_ARG, "rt_render: use_skybox without rt_set_skybox");
  if (p->row_band < 0 || (p->row_band > 0 && (p->row_band % 4 != 0 || p->row_period <= 0 || p->row_phase < 0 || p->row_phase >= p->row_period)))
    return fail(c, RT_ERR_INVALID_ARG, "rt_render: row_band must be a multiple of 4, 0 <= row_phase < row_period");
  if (p->row_begin == p->row_end) { L.skip = true; return RT_OK; }
  rt::KernelParams& kp = L.kp;
  std::memset(&kp, 0, sizeof(kp));
  kp.fp = *p;
  kp.tan_fov = std::tan(p->fov_deg * 0.017453292519943295f * 0.5f);   // tan(radians(fov) * 0.5), glsl:209
  kp.aspect = (float)p->width / (float)p->height;                      // glsl:208
  rt::DeviceScene& sc = kp.sc;
  sc.isect = c->d_scene; sc.plane = c->d_scene + c->off_plane;
  sc.material = c->d_scene + c->off_material; sc.lights = c->d_scene + c->off_lights;
  sc.num_objects = p->num_objects; sc.num_planes = c->n_planes; sc.num_lights = p->num_lights;
  const size_t smem = rt::scene_smem_bytes(sc.num_objects, sc.num_planes, sc.num_lights);
  sc.scene_in_smem = smem <= (size_t)c->max_smem_optin - 2048 ? 1 : 0;
  sc.sky = c->d_sky; sc.sky_tex = (unsigned long long)c->sky_tex; sc.sky_size = c->sky_size;
  sc.noise = c->d_noise; sc.noise_w = c->noise_w; sc.noise_h = c->noise_h;
  kp.tiles_x = (unsigned)((p->width + 7) / 8);
  int local_rows = p->row_end - p->row_begin;
  if (p->row_band > 0) {                                   // rows of the bands this caller owns
    const int nbands = (local_rows + p->row_band - 1) / p->row_band;
    const int mine = nbands > p->row_phase ? (nbands - p->row_phase + p->row_period - 1) / p->row_period : 0;
    if (mine == 0) { L.skip = true; return RT_OK; }
    local_rows = mine * p->row_band;                       // the last band may be cut by row_end (checked per pixel)
  }
  kp.tiles_y = (unsigned)((local_rows + 3) / 4);
  const unsigned bty = (unsigned)rt::packet_block_tiles_y();
  kp.blocks_x = (kp.tiles_x + 3) / 4; kp.blocks_y = (kp.tiles_y + bty - 1) / bty;
  kp.counters = c->d_counters;
  // Kernel selection: the packet kernel (rt_packet.cuh) unless the caller asks for the literal
  // one (counters need its reference-literal loops), renders a prefix of the uploaded arrays
  // (the packet blob is staged for the full scene), or the scene does not fit shared memory.
  const bool full_scene = p->num_objects == c->n_obj && p->num_lights == c->n_lights;
  L.use_packet = !(p->flags & (RT_FLAG_COUNTERS | RT_FLAG_LITERAL_KERNEL)) && full_scene && c->pk_float4s > 0 &&
                 rt::packet_smem_bytes((int)c->pk_float4s) <= (size_t)c->max_smem_optin - 1024;
  sc.pk_blob = c->d_scene + c->off_pk; sc.pk_float4s = L.use_packet ? (int)c->pk_float4s : 0;
  sc.pk_num_spheres = c->pk_spheres; sc.pk_num_planes = c->pk_planes; sc.pk_lights_offset = c->pk_lights_offset;
  sc.pk_num_clusters = c->pk_clusters;
  sc.pk_material_offset = c->pk_material_offset;
  sc.pk_grid = c->grid_f4s ? c->d_scene + c->off_grid : nullptr; sc.pk_grid_off_f4 = c->grid_off_f4; sc.pk_grid_ent_f4 = c->grid_ent_f4;
  sc.pk_fan_ok = c->fan_override >= 0 ? (c->fan_override && c->pk_fan_ok) : c->pk_fan_ok;
  // CTA-coherent phases pay off while the per-ray object loop is short (instruction-fetch bound);
  // with long loops the barrier imbalance costs more than the fetch locality gains.
  sc.pk_big = c->pk_spheres + c->pk_planes > 48 ? 1 : 0;
  L.par = (p->flags & RT_FLAG_PARITY_MATH) != 0u;
  // ---- the variants that apply (identical images; see rt_context::TuneEntry). v[0] is the default.
  L.nv = 0;
  if (L.use_packet) {
    if (!sc.pk_big) {
      // small scenes: barrier per block (1), + per bounce (2), + per light (3). Measured on B200: with bounces the
      // finer phases win slightly (C2 0.4366 / 0.4341 / 0.4318 ms), a single-bounce frame has too few phases for
      // the barriers to pay (C1 0.1033 / 0.1053 / 0.1074 ms).
      const int coh = c->coherent_override >= 0 ? c->coherent_override : (p->max_ray_depth <= 1 ? 1 : 3);
      L.v[L.nv++] = {RT_KERNEL_MEGAKERNEL, 0, coh};
    } else {
      for (int coh = 3; coh >= 0; coh -= 3)
        for (int dense = 0; dense <= 1; ++dense) {
          if (c->dense_override >= 0 && dense != (c->dense_override ? 1 : 0)) continue;
          if (c->coherent_override >= 0 && coh != 3) continue;
          L.v[L.nv++] = {RT_KERNEL_MEGAKERNEL, dense, c->coherent_override >= 0 ? c->coherent_override : coh};
        }
    }
    if (sc.num_lights <= rt::regroup_max_lights() && c->regroup_override != 0 &&
        rt::regroup_smem_bytes((int)c->pk_float4s, sc.num_lights, sc.pk_big != 0) <= (size_t)c->max_smem_optin - 1024) {
      if (c->regroup_override == 1) L.nv = 0;              // RT_PK_REGROUP=1 pins it
      L.v[L.nv++] = {RT_KERNEL_REGROUP, 0, 0};
    }
    if (rt::wavefront_supported(kp) && c->wavefront_override != 0 && c->regroup_override != 1) {
      if (c->wavefront_override == 1) L.nv = 0;            // RT_PK_WAVEFRONT=1 pins it
      L.v[L.nv++] = {RT_KERNEL_WAVEFRONT, 0, 0};
    }
  }
  const long long key[10] = {c->pk_spheres, c->pk_planes, c->n_lights, p->width, p->height, local_rows, p->max_ray_depth, p->num_samples,
                             (long long)(L.par ? 1 : 0) | ((long long)L.nv << 2) | ((long long)p->row_band << 8) | ((long long)p->row_period << 24),
                             (long long)(p->use_skybox ? 1 : 0) | ((long long)(p->flags & RT_FLAG_SKYBOX_HW) << 1)};
  std::memcpy(L.key, key, sizeof(key));
  return RT_OK;
}

static rt_context::TuneEntry* find_tune(rt_context* c, const LaunchPlan& L, bool create) {
  for (auto& t : c->tune)
    if (std::memcmp(t.key, L.key, sizeof(L.key)) == 0) { t.last_use = ++c->tune_clock; return &t; }
  if (!create) return nullptr;
  if ((int)c->tune.size() >= rt_context::kTuneEntries) {   // evict the least recently used shape
    size_t lru = 0;
    for (size_t i = 1; i < c->tune.size(); ++i) if (c->tune[i].last_use < c->tune[lru].last_use) lru = i;
    release_tune_events(c->tune[lru]);
    c->tune.erase(c->tune.begin() + (long)lru);
  }
  c->tune.emplace_back();
  rt_context::TuneEntry& t = c->tune.back();
  std::memcpy(t.key, L.key, sizeof(L.key));
  t.nv = L.nv;
  for (int i = 0; i < L.nv; ++i) t.v[i] = L.v[i];
  t.last_use = ++c->tune_clock;
  return &t;
}

// Applies a pin (rt_set_tuning) to the variant list: the first variant matching every pinned field.
static int pinned_variant(const rt_context* c, const LaunchPlan& L) {
  for (int i = 0; i < L.nv; ++i) {
    if (c->pin.kernel >= 0 && L.v[i].kernel != c->pin.kernel) continue;
    if (L.v[i].kernel == RT_KERNEL_MEGAKERNEL) {
      if (c->pin.dense >= 0 && L.kp.sc.pk_big && L.v[i].dense != c->pin.dense) continue;
      if (c->pin.coherent >= 0 && L.kp.sc.pk_big && L.v[i].coherent != c->pin.coherent) continue;
    }
    return i;
  }
  return 0;
}

// The wavefront pipeline for plan L (rt_wavefront.cuh): per chunk of tile rows and per sample, 2 D + 1 launches
// over the context's queue buffers. Asynchronous on `stream`; renders of one context that use the queues are
// ordered behind each other by an event.
static rt_status wavefront_render(rt_context* c, LaunchPlan& L, cudaStream_t stream) {
  const rt::KernelParams& kp = L.kp;
  const int nl = kp.sc.num_lights, D = kp.fp.max_ray_depth, nsamp = kp.fp.num_samples > 0 ? kp.fp.num_samples : 1;
  const unsigned per_row = kp.tiles_x * 32u;
  unsigned rows_per_chunk = (unsigned)(c->wf_chunk_pixels / per_row);
  if (rows_per_chunk < 1u) rows_per_chunk = 1u;
  if (rows_per_chunk > kp.tiles_y) rows_per_chunk = kp.tiles_y;
  const size_t cap = (size_t)rows_per_chunk * per_row;
  const size_t planes = 2 * 3 + 2 * 2 + 2 * (size_t)(nl > 0 ? nl : 1);
  const size_t acc_bytes = nsamp > 1 ? (size_t)kp.fp.width * kp.fp.height * sizeof(float4) : 0;
  const size_t need = planes * cap * sizeof(float4) + (size_t)(nl > 0 ? nl : 1) * cap * sizeof(unsigned) + acc_bytes + rt::kWfCounters * sizeof(unsigned) + 1024;
  if (need > c->wf_bytes) {
    RT_CUDA(c, cudaDeviceSynchronize());
    cudaFree(c->wf_state); c->wf_state = nullptr; c->wf_bytes = 0;
    RT_CUDA(c, cudaMalloc(&c->wf_state, need));
    c->wf_bytes = need;
  }
  if (!c->wf_done) RT_CUDA(c, cudaEventCreateWithFlags(&c->wf_done, cudaEventDisableTiming));
  else RT_CUDA(c, cudaStreamWaitEvent(stream, c->wf_done, 0));
  rt::WfParams wp;
  std::memset(&wp, 0, sizeof(wp));
  float4* f = (float4*)c->wf_state;
  wp.path[0] = f; f += 3 * cap; wp.path[1] = f; f += 3 * cap;
  wp.hit[0] = f; f += 2 * cap; wp.hit[1] = f; f += 2 * cap;
  wp.brdf[0] = f; f += (size_t)(nl > 0 ? nl : 1) * cap; wp.brdf[1] = f; f += (size_t)(nl > 0 ? nl : 1) * cap;
  wp.acc = nsamp > 1 ? f : nullptr; f += acc_bytes / sizeof(float4);
  wp.jobq = (unsigned*)f;
  wp.counters = wp.jobq + (size_t)(nl > 0 ? nl : 1) * cap;
  wp.cap = (unsigned)cap;
  for (unsigned row0 = 0; row0 < kp.tiles_y; row0 += rows_per_chunk) {
    wp.tile_row0 = row0;
    wp.tile_rows = kp.tiles_y - row0 < rows_per_chunk ? kp.tiles_y - row0 : rows_per_chunk;
    for (int sidx = 0; sidx < nsamp; ++sidx) {
      wp.sample = sidx;
      RT_CUDA(c, cudaMemsetAsync(wp.counters, 0, rt::kWfCounters * sizeof(unsigned), stream));
      for (int d = 0; d <= D; ++d) {
        wp.depth = d;
        cudaError_t e = L.par ? rt::launch_wf_bounce_parity(kp, wp, c->sm_count - c->sm_reserve, stream) : rt::launch_wf_bounce_fast(kp, wp, c->sm_count - c->sm_reserve, stream);
        if (e != cudaSuccess) return cuda_fail(c, e, "wavefront bounce launch");
        c->launches += 1;
        if (d == D || nl == 0) continue;
        e = L.par ? rt::launch_wf_shadow_parity(kp, wp, c->sm_count - c->sm_reserve, stream) : rt::launch_wf_shadow_fast(kp, wp, c->sm_count - c->sm_reserve, stream);
        if (e != cudaSuccess) return cuda_fail(c, e, "wavefront shadow launch");
        c->launches += 1;
      }
    }
  }
  RT_CUDA(c, cudaEventRecord(c->wf_done, stream));
  return RT_OK;
}

// One launch of variant `v` (or the literal kernel) of plan L on `stream`.
static rt_status launch_variant(rt_context* c, LaunchPlan& L, rt_context::Variant v, float* d_color, float* d_position, void* d_normal,
                                int32_t* d_object_id, cudaStream_t stream) {
  rt::KernelParams& kp = L.kp;
  kp.color = (float4*)d_color; kp.position = (float4*)d_position; kp.normal = d_normal; kp.object_id = d_object_id;
  kp.sc.pk_dense = v.dense; kp.sc.pk_coherent = v.coherent;
  if (L.use_packet && v.kernel == RT_KERNEL_WAVEFRONT) {   // its launches take their work from the queue counters, not from a slot
    rt_status st = wavefront_render(c, L, stream);
    if (st != RT_OK) return st;
    c->last_kernel = v.kernel; c->last_dense = 0; c->last_coherent = 0;
    return RT_OK;
  }
  rt_context::CounterSlot& slot = c->slots[c->next_slot];
  if (slot.used) RT_CUDA(c, cudaStreamWaitEvent(stream, slot.done, 0));   // a no-op when the previous user ran on this stream
  kp.tile_counter = c->d_tile_counter + 32 * c->next_slot; kp.tile_base = slot.base;
  unsigned advance = 0;
  cudaError_t e;
  if (L.use_packet && v.kernel == RT_KERNEL_REGROUP) {
    e = L.par ? rt::launch_regroup_parity(kp, c->sm_count - c->sm_reserve, stream, &advance) : rt::launch_regroup_fast(kp, c->sm_count - c->sm_reserve, stream, &advance);
    c->launches += 1;
  } else if (L.use_packet) {
    e = L.par ? rt::launch_packet_parity(kp, c->sm_count - c->sm_reserve, stream, &advance) : rt::launch_packet_fast(kp, c->sm_count - c->sm_reserve, stream, &advance);
    c->launches += 1;
  } else {
    e = L.par ? rt::launch_render_parity(kp, c->sm_count - c->sm_reserve, stream, &advance) : rt::launch_render_fast(kp, c->sm_count - c->sm_reserve, stream, &advance);
    c->launches += 1;
  }
  if (e != cudaSuccess) return cuda_fail(c, e, "rt_render launch");
  slot.base += kp.tiles_x * kp.tiles_y + advance;   // tiles + the over-fetch the launcher reports; unsigned wrap-around is intended
  RT_CUDA(c, cudaEventRecord(slot.done, stream));
  slot.used = true;
  c->next_slot = (c->next_slot + 1) % rt_context::kCounterSlots;
  c->last_kernel = L.use_packet ? v.kernel : RT_KERNEL_LITERAL; c->last_dense = v.dense; c->last_coherent = v.coherent;
  return RT_OK;
}

// Reads the finished trials of an entry and fixes its choice.
static void conclude_tune(rt_context::TuneEntry& t) {
  int arg = 0; float best = 0.0f; bool ok = true;
  for (int i = 0; i < t.nv && ok; ++i) {
    float vbest = 0.0f;
    for (int k = 0; k < rt_context::kTuneTrials; ++k) {
      float ms = 0.0f;
      if (cudaEventSynchronize(t.ev[i][k][1]) != cudaSuccess || cudaEventElapsedTime(&ms, t.ev[i][k][0], t.ev[i][k][1]) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
      if (k == 0 || ms < vbest) vbest = ms;
    }
    t.ms[i] = vbest;
    if (i == 0 || vbest < best) { best = vbest; arg = i; }
  }
  t.choice = ok ? arg : 0;
  t.chosen = true;
}

rt_status rt_render(rt_context* c, const rt_frame_params* p, float* d_color, float* d_position, void* d_normal,
                    int32_t* d_object_id, void* cuda_stream) {
  if (!c || !p) return RT_ERR_INVALID_ARG;
  LaunchPlan L;
  rt_status st = plan_launch(c, p, L);
  if (st != RT_OK || L.skip) return st;
  RT_CUDA(c, cudaSetDevice(c->device));
  nvtxRangePushA("rt_render");
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  rt_context::Variant run = L.nv ? L.v[0] : rt_context::Variant{RT_KERNEL_LITERAL, 0, 0};
  int trial = -1, trial_k = 0;
  rt_context::TuneEntry* t = nullptr;
  if (L.nv > 1 && c->pinned) run = L.v[pinned_variant(c, L)];
  else if (L.nv > 1) {
    t = find_tune(c, L, c->autotune);
    if (t && t->chosen) run = t->v[t->choice];
    else if (t && c->autotune) {
      // launch 1 of a new shape runs the default untimed (cold caches, module load); the next nv * kTuneTrials
      // launches are timed trials (variant-major), the one after reads the events and fixes the choice
      const int total = t->nv * rt_context::kTuneTrials;
      if (t->state == 0) t->state = 1;
      else if (t->state <= total) {
        trial = (t->state - 1) / rt_context::kTuneTrials; trial_k = (t->state - 1) % rt_context::kTuneTrials;
        for (int j = 0; j < 2; ++j)
          if (!t->ev[trial][trial_k][j]) { cudaError_t e = cudaEventCreate(&t->ev[trial][trial_k][j]); if (e != cudaSuccess) { nvtxRangePop(); return cuda_fail(c, e, "cudaEventCreate"); } }
        run = t->v[trial];
        t->state += 1;
      } else { conclude_tune(*t); run = t->v[t->choice]; }
    }
  }
  if (trial >= 0) cudaEventRecord(t->ev[trial][trial_k][0], stream);
  st = launch_variant(c, L, run, d_color, d_position, d_normal, d_object_id, stream);
  if (trial >= 0) cudaEventRecord(t->ev[trial][trial_k][1], stream);
  nvtxRangePop();
  return st;
}

rt_status rt_tune(rt_context* c, const rt_frame_params* p, rt_tuning* out) {
  if (!c || !p) return RT_ERR_INVALID_ARG;
  LaunchPlan L;
  rt_status st = plan_launch(c, p, L);
  if (st != RT_OK) return st;
  if (L.skip || L.nv <= 1) {
    if (out) fill_tuning(c, nullptr, L.nv ? L.v[0] : rt_context::Variant{RT_KERNEL_LITERAL, 0, 0}, RT_TUNING_CHOSEN, out);
    return RT_OK;
  }
  RT_CUDA(c, cudaSetDevice(c->device));
  st = ensure_outputs(c, (size_t)p->width * p->height);
  if (st != RT_OK) return st;
  rt_context::TuneEntry* t = find_tune(c, L, true);
  cudaEvent_t e0, e1;
  RT_CUDA(c, cudaEventCreate(&e0)); RT_CUDA(c, cudaEventCreate(&e1));
  const bool acc = (p->flags & RT_FLAG_ACCUMULATE) != 0u;
  // Warm-up first: a device that has been idle runs its first milliseconds below its boost clock (B200: a 0.36 ms frame
  // reads 0.41 ms when it is timed 10 ms after the context was created), which would penalise whichever variant is timed
  // first. >= 60 ms of default-variant launches, checked every 8 launches.
  {
    cudaEventRecord(e0, c->stream);
    for (int round = 0; round < 4096 && st == RT_OK; ++round) {
      for (int k = 0; k < 8 && st == RT_OK; ++k)
        st = launch_variant(c, L, t->v[0], (float*)c->o_color, (float*)c->o_position, c->o_normal, nullptr, c->stream);
      cudaEventRecord(e1, c->stream);
      float ms = 0.0f;
      if (cudaEventSynchronize(e1) != cudaSuccess || cudaEventElapsedTime(&ms, e0, e1) != cudaSuccess || ms >= 60.0f) break;
    }
  }
  // Each timed launch is enqueued BEHIND an untimed launch of the same variant, without synchronising in between: the
  // device is still busy while the host enqueues (e0, launch, e1), so the events bracket device time only — timing a
  // launch into an idle device would add the host's launch path (tens of microseconds, several launches for the
  // wavefront pipeline) to a 0.1 - 0.4 ms frame.
  for (int i = 0; i < t->nv && st == RT_OK; ++i) {
    float best = 0.0f;
    for (int k = 0; k < rt_context::kTuneTrials && st == RT_OK; ++k) {
      if (acc) cudaMemsetAsync(c->o_color, 0, (size_t)p->width * p->height * sizeof(float4), c->stream);
      st = launch_variant(c, L, t->v[i], (float*)c->o_color, (float*)c->o_position, c->o_normal, nullptr, c->stream);
      if (st != RT_OK) break;
      if (acc) cudaMemsetAsync(c->o_color, 0, (size_t)p->width * p->height * sizeof(float4), c->stream);
      cudaEventRecord(e0, c->stream);
      st = launch_variant(c, L, t->v[i], (float*)c->o_color, (float*)c->o_position, c->o_normal, nullptr, c->stream);
      cudaEventRecord(e1, c->stream);
      if (st != RT_OK) break;
      float ms = 0.0f;
      if (cudaEventSynchronize(e1) != cudaSuccess || cudaEventElapsedTime(&ms, e0, e1) != cudaSuccess) { st = cuda_fail(c, cudaGetLastError(), "rt_tune"); break; }
      if (k == 0 || ms < best) best = ms;
    }
    t->ms[i] = best;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (st != RT_OK) return st;
  int arg = 0;
  for (int i = 1; i < t->nv; ++i) if (t->ms[i] < t->ms[arg]) arg = i;
  t->choice = arg; t->chosen = true; t->state = t->nv * rt_context::kTuneTrials + 2;
  if (out) fill_tuning(c, t, t->v[arg], RT_TUNING_CHOSEN, out);
  return RT_OK;
}

rt_status rt_get_tuning(rt_context* c, const rt_frame_params* p, rt_tuning* out) {
  if (!c || !out) return RT_ERR_INVALID_ARG;
  if (!p) {
    fill_tuning(c, nullptr, rt_context::Variant{c->last_kernel, c->last_dense, c->last_coherent}, RT_TUNING_DEFAULT, out);
    return RT_OK;
  }
  LaunchPlan L;
  rt_status st = plan_launch(c, p, L);
  if (st != RT_OK) return st;
  if (L.skip || L.nv == 0) { fill_tuning(c, nullptr, rt_context::Variant{RT_KERNEL_LITERAL, 0, 0}, RT_TUNING_DEFAULT, out); return RT_OK; }
  rt_context::TuneEntry* t = L.nv > 1 ? find_tune(c, L, false) : nullptr;
  if (L.nv > 1 && c->pinned) fill_tuning(c, t, L.v[pinned_variant(c, L)], RT_TUNING_PINNED, out);
  else if (t && t->chosen) fill_tuning(c, t, t->v[t->choice], RT_TUNING_CHOSEN, out);
  else if (t && t->state > 0) fill_tuning(c, t, L.v[0], RT_TUNING_TRIALS, out);
  else fill_tuning(c, t, L.v[0], RT_TUNING_DEFAULT, out);
  if (!t) {                                                // report the applicable variants even when nothing is measured
    out->num_variants = L.nv;
    for (int i = 0; i < L.nv && i < 8; ++i) { out->variant_kernel[i] = L.v[i].kernel; out->variant_registers[i] = L.v[i].dense; out->variant_phase_level[i] = L.v[i].coherent; }
  }
  return RT_OK;
}

rt_status rt_set_tuning(rt_context* c, const rt_tuning* tuning) {
  if (!c) return RT_ERR_INVALID_ARG;
  if (!tuning) { c->pinned = false; return RT_OK; }
  if (tuning->kernel > RT_KERNEL_REGROUP || tuning->kernel == RT_KERNEL_LITERAL || tuning->registers > 1 || tuning->phase_level > 3)
    return fail(c, RT_ERR_INVALID_ARG, "rt_set_tuning: kernel / registers / phase_level out of range");
  c->pinned = true;
  c->pin = {tuning->kernel, tuning->registers, tuning->phase_level};
  return RT_OK;
}

rt_status rt_set_sm_reserve(rt_context* c, int sms) {
  if (!c) return RT_ERR_INVALID_ARG;
  if (sms < 0 || sms >= c->sm_count) return fail(c, RT_ERR_INVALID_ARG, "rt_set_sm_reserve: sms must be in [0, SM count)");
  c->sm_reserve = sms;
  return RT_OK;
}

rt_status rt_set_autotune(rt_context* c, int on) {
  if (!c) return RT_ERR_INVALID_ARG;
  c->autotune = on != 0;
  return RT_OK;
}

// The wavefront pipeline (rt_wavefront.cuh) with its ray queues in g_static_scratch. Calls no cudaMalloc / cudaFree and no
// allocating routine: the array's address was resolved in rt_create, the launch configurations come from the per-kernel
// cache, and the frame is cut into chunks of tile rows that fit the array. The queue counters are the only part of the
// scratch that must start at zero: cleared per chunk on the caller's stream. Everything else is written before it is read.
rt_status rt_render_static(rt_context* c, const rt_frame_params* p, float* d_color, float* d_position, void* d_normal,
                           int32_t* d_object_id, void* cuda_stream) {
  if (!c || !p) return RT_ERR_INVALID_ARG;
  if (!c->static_scratch) return fail(c, RT_ERR_CUDA, "rt_render_static: cudaGetSymbolAddress of the scratch array failed at rt_create: " + c->static_scratch_err);
  LaunchPlan L;
  rt_status st = plan_launch(c, p, L);
  if (st != RT_OK || L.skip) return st;
  rt::KernelParams& kp = L.kp;
  const int nsamp = p->num_samples > 0 ? p->num_samples : 1;
  if (nsamp != 1) return fail(c, RT_ERR_INVALID_ARG, "rt_render_static: num_samples must be 1 (a multi-sample frame needs a full-frame accumulator; use rt_render)");
  if (p->flags & (RT_FLAG_COUNTERS | RT_FLAG_LITERAL_KERNEL)) return fail(c, RT_ERR_INVALID_ARG, "rt_render_static: COUNTERS / LITERAL_KERNEL are not supported");
  if (!L.use_packet || !rt::wavefront_supported(kp))
    return fail(c, RT_ERR_INVALID_ARG, "rt_render_static: the scene must be fully staged in shared memory and have at most 32 lights, depth <= 16 (use rt_render)");
  RT_CUDA(c, cudaSetDevice(c->device));
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  const int nl = kp.sc.num_lights, D = kp.fp.max_ray_depth, nlp = nl > 0 ? nl : 1;
  const size_t counter_bytes = (rt::kWfCounters * sizeof(unsigned) + 255) / 256 * 256;
  const size_t slot_bytes = (size_t)(2 * 3 + 2 * 2 + 2 * nlp) * sizeof(float4) + (size_t)nlp * sizeof(unsigned);
  const size_t per_row = (size_t)kp.tiles_x * 32u;              // queue slots of one tile row (4 image rows)
  size_t rows_per_chunk = (kStaticScratchBytes - counter_bytes) / (per_row * slot_bytes);
  if (rows_per_chunk < 1) {
    char msg[256];
    snprintf(msg, sizeof(msg), "rt_render_static: one tile row of width %d with %d lights needs %zu bytes of queues, the static scratch holds %zu "
             "(maximum width at this light count: %zu pixels)", p->width, nl, per_row * slot_bytes + counter_bytes, kStaticScratchBytes,
             (kStaticScratchBytes - counter_bytes) / (slot_bytes * 32u) * 8u);
    return fail(c, RT_ERR_INVALID_ARG, msg);
  }
  if (rows_per_chunk > kp.tiles_y) rows_per_chunk = kp.tiles_y;
  const size_t cap = rows_per_chunk * per_row;
  kp.color = (float4*)d_color; kp.position = (float4*)d_position; kp.normal = d_normal; kp.object_id = d_object_id;
  kp.sc.pk_dense = 0; kp.sc.pk_coherent = 0;
  rt::WfParams wp;
  std::memset(&wp, 0, sizeof(wp));
  unsigned char* base = (unsigned char*)c->static_scratch;
  wp.counters = (unsigned*)base;
  float4* f = (float4*)(base + counter_bytes);
  wp.path[0] = f; f += 3 * cap; wp.path[1] = f; f += 3 * cap;
  wp.hit[0] = f; f += 2 * cap; wp.hit[1] = f; f += 2 * cap;
  wp.brdf[0] = f; f += (size_t)nlp * cap; wp.brdf[1] = f; f += (size_t)nlp * cap;
  wp.jobq = (unsigned*)f;
  wp.acc = nullptr;
  wp.cap = (unsigned)cap;
  wp.sample = 0;
  nvtxRangePushA("rt_render_static");
  struct Pop { ~Pop() { nvtxRangePop(); } } pop;
  for (unsigned row0 = 0; row0 < kp.tiles_y; row0 += (unsigned)rows_per_chunk) {
    wp.tile_row0 = row0;
    wp.tile_rows = kp.tiles_y - row0 < (unsigned)rows_per_chunk ? kp.tiles_y - row0 : (unsigned)rows_per_chunk;
    RT_CUDA(c, cudaMemsetAsync(wp.counters, 0, rt::kWfCounters * sizeof(unsigned), stream));
    for (int d = 0; d <= D; ++d) {
      wp.depth = d;
      cudaError_t e = L.par ? rt::launch_wf_bounce_parity(kp, wp, c->sm_count, stream) : rt::launch_wf_bounce_fast(kp, wp, c->sm_count, stream);
      if (e != cudaSuccess) return cuda_fail(c, e, "rt_render_static: bounce launch");
      c->launches += 1;
      if (d == D || nl == 0) continue;
      e = L.par ? rt::launch_wf_shadow_parity(kp, wp, c->sm_count, stream) : rt::launch_wf_shadow_fast(kp, wp, c->sm_count, stream);
      if (e != cudaSuccess) return cuda_fail(c, e, "rt_render_static: shadow launch");
      c->launches += 1;
    }
  }
  c->last_kernel = RT_KERNEL_WAVEFRONT; c->last_dense = 0; c->last_coherent = 0;
  return RT_OK;
}

rt_status rt_render_host(rt_context* c, const rt_frame_params* p, const void* objects, int n, const void* lights, int m,
                         float* h_color, float* h_position, void* h_normal, int32_t* h_object_id) {
  if (!c || !p) return RT_ERR_INVALID_ARG;
  if (p->width <= 0 || p->height <= 0) return fail(c, RT_ERR_INVALID_ARG, "rt_render_host: image size");
  RT_CUDA(c, cudaSetDevice(c->device));
  rt_status st;
  if (objects || lights) {                               // ssbo.update() + lightSSBO.update(), every frame
    st = rt_upload_scene(c, objects, n, lights, m, c->stream, 1);
    if (st != RT_OK) return st;
  }
  // How the three images reach the host (rt_set_host_path):
  //   ZEROCOPY  pinned host buffers are mapped into the device address space (UVA), so the
  //             kernel's float4 stores go straight over PCIe while later tiles are still being
  //             traced: no device-side frame, no separate copy.
  //   BANDED    rows are rendered in bands; band k is copied D2H on a second stream while band
  //             k+1 renders.
  //   STAGED    render the frame, then copy it (the v0 behaviour).
  int path = c->host_path;
  if (p->row_band > 0 && path != RT_HOST_PATH_AUTO && path != RT_HOST_PATH_ZEROCOPY)
    return fail(c, RT_ERR_INVALID_ARG, "rt_render_host: interleaved row bands need the zero-copy path");
  auto mapped = [](const void* h) -> bool {
    if (!h) return true;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, h) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost && a.devicePointer == h;
  };
  const bool all_mapped = mapped(h_color) && mapped(h_position) && mapped(h_normal) && mapped(h_object_id);
  if (path == RT_HOST_PATH_AUTO) path = all_mapped ? RT_HOST_PATH_ZEROCOPY : RT_HOST_PATH_BANDED;   // measured: profiles/README.md
  if (path == RT_HOST_PATH_ZEROCOPY && !all_mapped) path = RT_HOST_PATH_STAGED;
  if (p->row_band > 0 && path != RT_HOST_PATH_ZEROCOPY)
    return fail(c, RT_ERR_INVALID_ARG, "rt_render_host: interleaved row bands need pinned, device-mapped host buffers");
  if (path == RT_HOST_PATH_ZEROCOPY) {
    st = rt_render(c, p, h_color, h_position, h_normal, h_object_id, c->stream);
    if (st != RT_OK) return st;
    RT_CUDA(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
  }
  const size_t pixels = (size_t)p->width * p->height;
  st = ensure_outputs(c, pixels);
  if (st != RT_OK) return st;
  const size_t nbytes = (p->flags & RT_FLAG_NORMAL_FP32) ? sizeof(float4) : sizeof(uint2);
  const int rows = p->row_end - p->row_begin;
  int bands = 1;
  if (path == RT_HOST_PATH_BANDED) {                     // ~2 MB+ of output per band, at most 16 bands
    const size_t row_bytes = (size_t)p->width * 40;
    bands = (int)(((size_t)rows * row_bytes) / (6u << 20));
    if (bands < 1) bands = 1;
    if (bands > 16) bands = 16;
  }
  int per = ((rows + bands - 1) / bands + 3) / 4 * 4;    // band height: a multiple of the tile height
  if (per < 4) per = 4;
  int b = 0;
  for (int r0 = p->row_begin; r0 < p->row_end; r0 += per, ++b) {
    rt_frame_params q = *p;
    q.row_begin = r0;
    q.row_end = r0 + per < p->row_end ? r0 + per : p->row_end;
    st = rt_render(c, &q, h_color ? (float*)c->o_color : nullptr, h_position ? (float*)c->o_position : nullptr,
                   h_normal ? c->o_normal : nullptr, h_object_id ? c->o_id : nullptr, c->stream);
    if (st != RT_OK) return st;
    cudaStream_t cs = bands > 1 ? c->copy_stream : c->stream;
    if (bands > 1) {
      RT_CUDA(c, cudaEventRecord(c->band_done[b & 15], c->stream));
      RT_CUDA(c, cudaStreamWaitEvent(cs, c->band_done[b & 15], 0));
    }
    const size_t first = (size_t)q.row_begin * p->width, count = (size_t)(q.row_end - q.row_begin) * p->width;
    if (h_color) RT_CUDA(c, cudaMemcpyAsync(h_color + first * 4, c->o_color + first, count * sizeof(float4), cudaMemcpyDeviceToHost, cs));
    if (h_position) RT_CUDA(c, cudaMemcpyAsync(h_position + first * 4, c->o_position + first, count * sizeof(float4), cudaMemcpyDeviceToHost, cs));
    if (h_normal) RT_CUDA(c, cudaMemcpyAsync((char*)h_normal + first * nbytes, (char*)c->o_normal + first * nbytes, count * nbytes, cudaMemcpyDeviceToHost, cs));
    if (h_object_id) RT_CUDA(c, cudaMemcpyAsync(h_object_id + first, c->o_id + first, count * sizeof(int), cudaMemcpyDeviceToHost, cs));
  }
  RT_CUDA(c, cudaStreamSynchronize(c->stream));
  if (bands > 1) RT_CUDA(c, cudaStreamSynchronize(c->copy_stream));
  return RT_OK;
}

rt_status rt_set_host_path(rt_context* c, int path) {
  if (!c || path < RT_HOST_PATH_AUTO || path > RT_HOST_PATH_ZEROCOPY) return RT_ERR_INVALID_ARG;
  c->host_path = path;
  return RT_OK;
}

rt_status rt_synchronize(rt_context* c) {
  if (!c) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  RT_CUDA(c, cudaDeviceSynchronize());
  return RT_OK;
}

rt_status rt_render_gl(rt_context* c, const rt_frame_params*, unsigned, unsigned, unsigned) {
  return fail(c, RT_ERR_UNAVAILABLE, "rt_render_gl: no OpenGL/EGL stack on this machine (CUDA-GL interop unavailable on this box)");
}

rt_status rt_get_counters(rt_context* c, uint64_t* out, int reset) {
  if (!c || !out) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  RT_CUDA(c, cudaDeviceSynchronize());
  RT_CUDA(c, cudaMemcpy(out, c->d_counters, sizeof(uint64_t) * RT_CNT_COUNT, cudaMemcpyDeviceToHost));
  if (reset) RT_CUDA(c, cudaMemset(c->d_counters, 0, sizeof(uint64_t) * RT_CNT_COUNT));
  return RT_OK;
}

rt_status rt_device_alloc(rt_context* c, void** out, size_t bytes) {
  if (!c || !out) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  RT_CUDA(c, cudaMalloc(out, bytes ? bytes : 1));
  return RT_OK;
}
rt_status rt_device_free(rt_context* c, void* p) {
  if (!c) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  RT_CUDA(c, cudaFree(p));
  return RT_OK;
}
rt_status rt_ipc_export(rt_context* c, void* device_ptr, unsigned char handle[64]) {
  if (!c || !device_ptr || !handle) return RT_ERR_INVALID_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
  RT_CUDA(c, cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  RT_CUDA(c, cudaIpcGetMemHandle(&h, device_ptr));
  std::memcpy(handle, &h, 64);
  return RT_OK;
}
rt_status rt_ipc_open(rt_context* c, const unsigned char handle[64], void** out) {
  if (!c || !handle || !out) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, 64);
  RT_CUDA(c, cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return RT_OK;
}
rt_status rt_ipc_close(rt_context* c, void* p) {
  if (!c) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  RT_CUDA(c, cudaIpcCloseMemHandle(p));
  return RT_OK;
}

rt_status rt_host_alloc(void** out, size_t bytes) {
  if (!out) return RT_ERR_INVALID_ARG;
  cudaError_t e = cudaHostAlloc(out, bytes, cudaHostAllocDefault);
  return e == cudaSuccess ? RT_OK : (e == cudaErrorMemoryAllocation ? RT_ERR_OOM : RT_ERR_CUDA);
}

rt_status rt_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? RT_OK : RT_ERR_CUDA; }

rt_status rt_measure_fp32_peak(rt_context* c, double* tflops, int* sm_count, double* ms) {
  if (!c || !tflops) return RT_ERR_INVALID_ARG;
  RT_CUDA(c, cudaSetDevice(c->device));
  double t = 0, m = 0;
  cudaError_t e = rt::measure_fp32_peak(c->sm_count, c->stream, &t, &m);
  if (e != cudaSuccess) return cuda_fail(c, e, "rt_measure_fp32_peak");
  c->launches += 4;
  *tflops = t;
  if (sm_count) *sm_count = c->sm_count;
  if (ms) *ms = m;
  return RT_OK;
}

uint64_t rt_launch_count(rt_context* c) { return c ? c->launches : 0; }

size_t rt_debug_light_grids(const void* objects, int n, const void* lights, int m, void* out, size_t capacity_f4,
                            int* off_f4, int* ent_f4, int* slot_to_object) {
  if (n < 0 || m < 0 || (n > 0 && !objects) || (m > 0 && !lights)) return 0;
  const rt_object* obj = (const rt_object*)objects;
  const GridBlob g = build_light_grids(obj, n, (const rt_light*)lights, m);
  if (off_f4) *off_f4 = g.off_f4;
  if (ent_f4) *ent_f4 = g.ent_f4;
  if (slot_to_object) {
    const std::vector<int> order = sphere_slot_order(obj, n, true);
    for (size_t k = 0; k < order.size(); ++k) slot_to_object[k] = order[k];
  }
  if (out && g.data.size() <= capacity_f4 && !g.data.empty()) std::memcpy(out, g.data.data(), g.data.size() * sizeof(float4));
  return g.data.size();
}

}  // extern "C"
