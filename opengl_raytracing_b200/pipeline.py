"""Host-side mirror of the reference's interface for the ray-tracing stage.

The reference drives the pass from one call site, ``ForwardShadingPipline::Render``
(/root/reference/src/ForwardShadingPipeline.cpp:155-182): ``ssbo.update()``,
``lightSSBO.update()``, ten ``raytracingShader.set*`` uniform calls, ``glDispatchCompute``,
``glMemoryBarrier``. The classes below keep those names and meanings — ``SSBO``,
``LightSSBO``, ``RaytracingShader.setInt/setFloat/setVec3/setVec2/setBool``,
``ForwardShadingPipeline.Init/Render`` — over the C ABI in include/rt_abi.h, so a test
written against the reference's host code reads the same here. GL objects become headless
device buffers (torch tensors used purely as device memory + streams).

There is no CPU path: constructing ``RtContext`` without the CUDA library or a GPU raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import abi
from . import scene as S


def _ptr(a) -> C.c_void_p:
    if a is None:
        return C.c_void_p(0)
    if isinstance(a, C.c_void_p):
        return a                                   # raw device address (e.g. a peer-mapped frame)
    if isinstance(a, int):
        return C.c_void_p(a)
    if isinstance(a, np.ndarray):
        return C.c_void_p(a.ctypes.data)
    return C.c_void_p(int(a.data_ptr()))          # torch tensor


class RtContext:
    """Owns one ``rt_context`` (one CUDA device). Thin, 1:1 with the C entry points."""

    def __init__(self, device: int = 0, lib_path: str | None = None):
        self.lib = abi.load_library(lib_path)
        self.handle = C.c_void_p(0)
        self.device = device
        st = self.lib.rt_create(C.byref(self.handle), device)
        if st != abi.RT_OK:
            raise abi.RtError(st, "rt_create failed: " + self.lib.rt_status_string(st).decode()
                              + " (a CUDA device is required; there is no CPU fallback)")
        self._keep = {}

    def _check(self, st: int) -> None:
        if st != abi.RT_OK:
            msg = self.lib.rt_last_error(self.handle)
            raise abi.RtError(st, (msg or b"").decode() or self.lib.rt_status_string(st).decode())

    def close(self) -> None:
        if self.handle:
            self.lib.rt_destroy(self.handle)
            self.handle = C.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # SSBO::update + LightSSBO::update (src/SSBO.h:16-23, src/LightSSBO.h:16-25)
    def upload_scene(self, objects: np.ndarray, lights: np.ndarray, stream: int = 0, async_: bool = False) -> None:
        objects = np.ascontiguousarray(objects, dtype=S.OBJECT_DTYPE)
        lights = np.ascontiguousarray(lights, dtype=S.LIGHT_DTYPE)
        self._keep["scene"] = (objects, lights)
        self._check(self.lib.rt_upload_scene(self.handle, _ptr(objects), len(objects), _ptr(lights), len(lights),
                                             C.c_void_p(stream), 1 if async_ else 0))

    # glBindTexture(GL_TEXTURE_CUBE_MAP, ...) (src/ForwardShadingPipeline.cpp:167-170)
    def set_skybox(self, faces: np.ndarray | None) -> None:
        if faces is None:
            self._check(self.lib.rt_set_skybox(self.handle, None, 0, 0, 0))
            return
        assert faces.ndim == 4 and faces.shape[0] == 6 and faces.shape[1] == faces.shape[2]
        assert faces.dtype in (np.float16, np.float32) and faces.shape[3] in (3, 4)
        faces = np.ascontiguousarray(faces)
        self._check(self.lib.rt_set_skybox(self.handle, _ptr(faces), faces.shape[1], faces.shape[3],
                                           1 if faces.dtype == np.float16 else 0))

    # InitBlueNoiseTex (src/ForwardShadingPipeline.cpp:39-49)
    def set_noise(self, noise: np.ndarray | None) -> None:
        if noise is None:
            self._check(self.lib.rt_set_noise(self.handle, None, 0, 0))
            return
        noise = np.ascontiguousarray(noise, dtype=np.uint8)
        self._check(self.lib.rt_set_noise(self.handle, _ptr(noise), noise.shape[1], noise.shape[0]))

    # glDispatchCompute + glMemoryBarrier (src/ForwardShadingPipeline.cpp:175-180)
    def render(self, params: abi.FrameParams, color=None, position=None, normal=None, object_id=None,
               stream: int = 0) -> None:
        self._check(self.lib.rt_render(self.handle, C.byref(params), _ptr(color), _ptr(position), _ptr(normal),
                                       _ptr(object_id), C.c_void_p(stream)))

    # rt_render_static: rt_render without any allocation (wavefront queues in a static device array, include/rt_abi.h)
    def render_static(self, params: abi.FrameParams, color=None, position=None, normal=None, object_id=None,
                      stream: int = 0) -> None:
        self._check(self.lib.rt_render_static(self.handle, C.byref(params), _ptr(color), _ptr(position), _ptr(normal),
                                              _ptr(object_id), C.c_void_p(stream)))

    def render_host(self, params: abi.FrameParams, objects, lights, color, position, normal, object_id=None) -> None:
        n = 0 if objects is None else len(objects)
        m = 0 if lights is None else len(lights)
        self._check(self.lib.rt_render_host(self.handle, C.byref(params), _ptr(objects), n, _ptr(lights), m,
                                            _ptr(color), _ptr(position), _ptr(normal), _ptr(object_id)))

    def synchronize(self) -> None:
        self._check(self.lib.rt_synchronize(self.handle))

    def counters(self, reset: bool = True) -> dict:
        buf = (C.c_uint64 * abi.RT_CNT_COUNT)()
        self._check(self.lib.rt_get_counters(self.handle, buf, 1 if reset else 0))
        return {k: int(buf[i]) for i, k in enumerate(abi.COUNTER_NAMES)}

    def measure_fp32_peak(self) -> dict:
        tf, sm, ms = C.c_double(0), C.c_int(0), C.c_double(0)
        self._check(self.lib.rt_measure_fp32_peak(self.handle, C.byref(tf), C.byref(sm), C.byref(ms)))
        return {"tflops": tf.value, "sm_count": sm.value, "ms": ms.value}

    def launch_count(self) -> int:
        return int(self.lib.rt_launch_count(self.handle))

    # kernel variants (identical images): rt_tune / rt_get_tuning / rt_set_tuning / rt_set_autotune
    def tune(self, params: abi.FrameParams) -> dict:
        t = abi.Tuning()
        self._check(self.lib.rt_tune(self.handle, C.byref(params), C.byref(t)))
        return t.as_dict()

    def get_tuning(self, params: abi.FrameParams | None = None) -> dict:
        t = abi.Tuning()
        self._check(self.lib.rt_get_tuning(self.handle, C.byref(params) if params is not None else None, C.byref(t)))
        return t.as_dict()

    def set_tuning(self, kernel: int | None = None, registers: int = -1, phase_level: int = -1) -> None:
        """kernel None removes the pin; fields < 0 are left to the library."""
        if kernel is None:
            self._check(self.lib.rt_set_tuning(self.handle, None))
            return
        t = abi.Tuning()
        t.kernel, t.registers, t.phase_level = kernel, registers, phase_level
        self._check(self.lib.rt_set_tuning(self.handle, C.byref(t)))

    def set_autotune(self, on: bool) -> None:
        self._check(self.lib.rt_set_autotune(self.handle, 1 if on else 0))

    def set_sm_reserve(self, sms: int) -> None:
        """Render grids are sized for (SM count - sms) SMs: room for a kernel of another stream (rt_set_sm_reserve)."""
        self._check(self.lib.rt_set_sm_reserve(self.handle, int(sms)))


def frame_params(width: int, height: int, num_objects: int, num_lights: int, camera: S.Camera | None = None,
                 max_ray_depth: int = 3, frame_count: int = 0, num_samples: int = 1, use_skybox: bool = False,
                 row_begin: int = 0, row_end: int | None = None, flags: int = 0,
                 noise_scale: float = 1.0 / 1024.0, row_band: int = 0, row_period: int = 1,
                 row_phase: int = 0) -> abi.FrameParams:
    """The uniform block of one dispatch with the reference's values: ``fov`` = Camera::FOV
    (45), ``focalLength`` 1.0 and ``maxRayDistance`` 114514.0 (shader defaults the host never
    overrides, glsl:80,85), ``noiseScale`` 1/1024 (src/ForwardShadingPipeline.cpp:164)."""
    cam = camera or S.Camera()
    p = abi.FrameParams()
    p.width, p.height = width, height
    p.row_begin, p.row_end = row_begin, height if row_end is None else row_end
    p.num_objects, p.num_lights = num_objects, num_lights
    p.camera_pos[:] = cam.position
    p.camera_dir[:] = cam.front
    p.camera_up[:] = cam.up
    p.camera_right[:] = cam.right
    p.fov_deg, p.focal_length, p.max_ray_distance = cam.fov, cam.focal_length, 114514.0
    p.use_skybox = 1 if use_skybox else 0
    p.noise_scale[:] = (noise_scale, noise_scale)
    p.frame_count, p.num_samples, p.max_ray_depth = frame_count, num_samples, max_ray_depth
    p.flags = flags
    p.row_band, p.row_period, p.row_phase = row_band, row_period, row_phase
    return p


def workload_params(w, flags: int = 0, **over) -> abi.FrameParams:
    """FrameParams for a ``synthetic.Workload``."""
    kw = dict(width=w.width, height=w.height, num_objects=len(w.objects), num_lights=len(w.lights),
              camera=w.camera, max_ray_depth=w.max_ray_depth, frame_count=w.frame_count,
              num_samples=w.num_samples, use_skybox=w.use_skybox, flags=flags)
    kw.update(over)
    return frame_params(**kw)


class MultiGpu:
    """``rt_multi`` (include/rt_multi.h): one process driving several GPUs of the box — row bands stored straight
    into the root GPU's images over NVLink, or sample ranges summed by the peer pull-reduce kernel."""

    def __init__(self, devices):
        self.lib = abi.load_library()
        self.handle = C.c_void_p(0)
        self.devices = list(devices)
        arr = (C.c_int * len(self.devices))(*self.devices)
        st = self.lib.rt_multi_create(C.byref(self.handle), arr, len(self.devices))
        if st != abi.RT_OK:
            raise abi.RtError(st, "rt_multi_create failed: " + self.lib.rt_status_string(st).decode())
        self._keep = {}

    def _check(self, st: int) -> None:
        if st != abi.RT_OK:
            raise abi.RtError(st, (self.lib.rt_multi_last_error(self.handle) or b"").decode() or self.lib.rt_status_string(st).decode())

    def bind_workload(self, w) -> None:
        objects = np.ascontiguousarray(w.objects, dtype=S.OBJECT_DTYPE)
        lights = np.ascontiguousarray(w.lights, dtype=S.LIGHT_DTYPE)
        self._keep["scene"] = (objects, lights)
        self._check(self.lib.rt_multi_upload_scene(self.handle, _ptr(objects), len(objects), _ptr(lights), len(lights)))
        if w.use_skybox and w.skybox is not None:
            faces = np.ascontiguousarray(w.skybox)
            self._check(self.lib.rt_multi_set_skybox(self.handle, _ptr(faces), faces.shape[1], faces.shape[3], 1 if faces.dtype == np.float16 else 0))
        else:
            self._check(self.lib.rt_multi_set_skybox(self.handle, None, 0, 0, 0))
        if w.noise is None:
            self._check(self.lib.rt_multi_set_noise(self.handle, None, 0, 0))
        else:
            noise = np.ascontiguousarray(w.noise, dtype=np.uint8)
            self._check(self.lib.rt_multi_set_noise(self.handle, _ptr(noise), noise.shape[1], noise.shape[0]))

    def render(self, params: abi.FrameParams, sharding: int, color=None, position=None, normal=None, object_id=None) -> None:
        self._check(self.lib.rt_multi_render(self.handle, C.byref(params), sharding, _ptr(color), _ptr(position), _ptr(normal), _ptr(object_id)))

    def synchronize(self) -> None:
        self._check(self.lib.rt_multi_synchronize(self.handle))

    def render_host(self, params: abi.FrameParams, sharding: int, objects, lights, color, position, normal, object_id=None) -> None:
        n = 0 if objects is None else len(objects)
        m = 0 if lights is None else len(lights)
        self._check(self.lib.rt_multi_render_host(self.handle, C.byref(params), sharding, _ptr(objects), n, _ptr(lights), m,
                                                  _ptr(color), _ptr(position), _ptr(normal), _ptr(object_id)))

    def launch_count(self) -> int:
        return int(self.lib.rt_multi_launch_count(self.handle))

    def close(self) -> None:
        if self.handle:
            self.lib.rt_multi_destroy(self.handle)
            self.handle = C.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# --------------------------------------------------------------------------- reference-named mirror
class SSBO:
    """src/SSBO.h: ``objects`` vector + ``update()`` (re-uploaded whole, every frame)."""

    def __init__(self, pipeline: "ForwardShadingPipeline"):
        self._p = pipeline
        self.objects = S.new_objects(0)

    def update(self) -> None:
        self._p._scene_dirty = True


class LightSSBO:
    """src/LightSSBO.h: ``lights`` vector + ``update()``."""

    def __init__(self, pipeline: "ForwardShadingPipeline"):
        self._p = pipeline
        self.lights = S.new_lights(0)

    def update(self) -> None:
        self._p._scene_dirty = True


class RaytracingShader:
    """src/Shader.h:106-128 ``use`` / ``set*`` by uniform NAME, for raytracingCs.glsl's
    uniforms (glsl:72-89). Unknown names are ignored, as glGetUniformLocation == -1 is."""

    _VEC3 = {"cameraPos": "camera_pos", "cameraDir": "camera_dir", "cameraUp": "camera_up", "cameraRight": "camera_right"}
    _FLOAT = {"fov": "fov_deg", "focalLength": "focal_length", "maxRayDistance": "max_ray_distance"}
    _INT = {"numObjects": "num_objects", "numLights": "num_lights", "frameCount": "frame_count"}

    def __init__(self):
        # shader-side defaults: fov 60, focalLength 1, maxRayDistance 114514 (glsl:79-85)
        self.params = frame_params(0, 0, 0, 0)
        self.params.fov_deg = 60.0

    def use(self) -> None:
        pass

    def setInt(self, name: str, v: int) -> None:
        if name in self._INT:
            setattr(self.params, self._INT[name], int(v))

    def setBool(self, name: str, v: bool) -> None:
        if name == "useSkybox":
            self.params.use_skybox = 1 if v else 0

    def setFloat(self, name: str, v: float) -> None:
        if name in self._FLOAT:
            setattr(self.params, self._FLOAT[name], float(v))

    def setVec3(self, name: str, v) -> None:
        if name in self._VEC3:
            getattr(self.params, self._VEC3[name])[:] = tuple(float(x) for x in v)

    def setVec2(self, name: str, v) -> None:
        if name == "noiseScale":
            self.params.noise_scale[:] = tuple(float(x) for x in v)


class ForwardShadingPipeline:
    """The ray-tracing stage of ``ForwardShadingPipline`` (src/ForwardShadingPipeline.{h,cpp}):
    ``Init`` creates the three output images (outputTex RGBA32F :57-65, gPositionTex RGBA32F
    :115-120, gNormalTex RGBA16F :121-126) as device buffers; ``Render`` performs the
    :155-182 slice of one frame. The window, UI and post chain are out of scope."""

    def __init__(self, width: int = 800, height: int = 800, device: int = 0, max_ray_depth: int = 3):
        import torch
        self._torch = torch
        self.WIDTH, self.HEIGHT = width, height           # src/global.cpp:4-5
        self.ctx = RtContext(device)
        self.camera = S.Camera()
        self.ssbo = SSBO(self)
        self.lightSSBO = LightSSBO(self)
        self.raytracingShader = RaytracingShader()
        self.max_ray_depth = max_ray_depth                # #define MAX_RAY_DEPTH (glsl:4)
        self.frameCount = 0
        self.skyboxEnabled = False
        self.taaEnabled = True                            # frameCount advances only under TAA (:232-254)
        self._scene_dirty = True
        self.Init()

    def Init(self) -> None:
        t, dev = self._torch, f"cuda:{self.ctx.device}"
        self.outputTex = t.zeros(self.HEIGHT, self.WIDTH, 4, dtype=t.float32, device=dev)
        self.gPositionTex = t.zeros(self.HEIGHT, self.WIDTH, 4, dtype=t.float32, device=dev)
        self.gNormalTex = t.zeros(self.HEIGHT, self.WIDTH, 4, dtype=t.float16, device=dev)

    def SetSkybox(self, faces) -> None:
        self.ctx.set_skybox(faces)
        self.skyboxEnabled = faces is not None

    def SetNoise(self, noise) -> None:
        self.ctx.set_noise(noise)

    def Render(self, flags: int = 0):
        """One frame's ray-tracing stage; returns (outputTex, gPositionTex, gNormalTex)."""
        if self._scene_dirty:                             # ssbo.update() / lightSSBO.update()
            self.ctx.upload_scene(self.ssbo.objects, self.lightSSBO.lights)
            self._scene_dirty = False
        sh, cam = self.raytracingShader, self.camera
        sh.use()
        sh.setInt("numObjects", len(self.ssbo.objects))
        sh.setInt("numLights", len(self.lightSSBO.lights))
        sh.setVec3("cameraPos", cam.position)
        sh.setVec3("cameraDir", cam.front)
        sh.setVec3("cameraUp", cam.up)
        sh.setVec3("cameraRight", cam.right)
        sh.setFloat("fov", cam.fov)
        sh.setInt("frameCount", self.frameCount)
        sh.setVec2("noiseScale", (1.0 / 1024.0, 1.0 / 1024.0))
        sh.setBool("useSkybox", self.skyboxEnabled)
        p = sh.params
        p.width, p.height, p.row_begin, p.row_end = self.WIDTH, self.HEIGHT, 0, self.HEIGHT
        p.num_samples, p.max_ray_depth, p.flags = 1, self.max_ray_depth, flags
        stream = self._torch.cuda.current_stream(self.ctx.device).cuda_stream
        self.ctx.render(p, self.outputTex, self.gPositionTex, self.gNormalTex, None, stream)   # dispatch + barrier
        if self.taaEnabled:
            self.frameCount += 1
        return self.outputTex, self.gPositionTex, self.gNormalTex


# --------------------------------------------------------------------------- convenience
def bind_workload(ctx: RtContext, w) -> None:
    """Upload a ``synthetic.Workload``'s scene, skybox and noise (what Init + the first
    ssbo.update() do in the reference)."""
    ctx.upload_scene(w.objects, w.lights)
    ctx.set_skybox(w.skybox if w.use_skybox else None)
    ctx.set_noise(w.noise)


def render_workload(ctx: RtContext, w, flags: int = 0, params: abi.FrameParams | None = None, bind: bool = True):
    """Render a workload into fresh device buffers and return host copies:
    dict(color (H,W,4) f32, position (H,W,4) f32, normal (H,W,4) f32 — from the RGBA16F image
    unless FLAG_NORMAL_FP32 —, object_id (H,W) i32). Row 0 = gl y 0."""
    import torch
    if bind:
        bind_workload(ctx, w)
    p = params or workload_params(w, flags=flags)
    dev = f"cuda:{ctx.device}"
    H, W = p.height, p.width
    color = torch.zeros(H, W, 4, dtype=torch.float32, device=dev)
    position = torch.zeros(H, W, 4, dtype=torch.float32, device=dev)
    ndt = torch.float32 if (p.flags & abi.FLAG_NORMAL_FP32) else torch.float16
    normal = torch.zeros(H, W, 4, dtype=ndt, device=dev)
    oid = torch.full((H, W), -3, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(ctx.device).cuda_stream
    ctx.render(p, color, position, normal, oid, stream)
    torch.cuda.synchronize(ctx.device)
    return {"color": color.cpu().numpy(), "position": position.cpu().numpy(),
            "normal": normal.float().cpu().numpy(), "object_id": oid.cpu().numpy()}
