"""oracle — TEST INFRASTRUCTURE. CPU checkers for the ray-tracing pass.

Two shared libraries export the same C API (oracle_api.h):
  * ``port``      oracle/librt_oracle.so      — rt_oracle.cpp, this repo's CPU restatement
  * ``reference`` oracle/_ref/libref_shader.so — the reference's own GLSL compiled as C++
                                                 (make_ref.py; needs /root/reference to BUILD,
                                                 not to run)
Only tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import this
package. The product (opengl_raytracing_b200) never does and has no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from opengl_raytracing_b200 import abi
from opengl_raytracing_b200.pipeline import workload_params

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "librt_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libref_shader.so")
REFERENCE_ROOT = "/root/reference"

_libs: dict = {}


def build_port(force: bool = False) -> str:
    src = os.path.join(HERE, "rt_oracle.cpp")
    deps = [src, os.path.join(HERE, "oracle_api.h"), os.path.join(HERE, "..", "include", "rt_abi.h")]
    if not force and os.path.isfile(PORT_SO) and all(os.path.getmtime(PORT_SO) >= os.path.getmtime(d) for d in deps):
        return PORT_SO
    subprocess.run(["g++", "-O2", "-std=c++17", "-fopenmp", "-ffp-contract=off", "-fno-fast-math", "-fPIC",
                    "-shared", "-Wall", "-o", PORT_SO, src], check=True)
    _libs.pop("port", None)
    return PORT_SO


def build_ref(force: bool = False) -> str | None:
    """Builds oracle/_ref from /root/reference when that tree is present (this container);
    on the GPU box the prebuilt .so that travelled with the snapshot is used as is."""
    from . import make_ref
    if not os.path.isfile(make_ref.REF_SHADER):
        return REF_SO if os.path.isfile(REF_SO) else None
    deps = [make_ref.REF_SHADER, os.path.join(HERE, "ref_harness.cpp"), os.path.join(HERE, "glsl_shim.h"),
            os.path.join(HERE, "make_ref.py")]
    if not force and os.path.isfile(REF_SO) and all(os.path.getmtime(REF_SO) >= os.path.getmtime(d) for d in deps):
        return REF_SO
    make_ref.build()
    _libs.pop("reference", None)
    return REF_SO


def have(kind: str) -> bool:
    return os.path.isfile(PORT_SO if kind == "port" else REF_SO)


def load(kind: str = "port") -> C.CDLL:
    if kind in _libs:
        return _libs[kind]
    path = PORT_SO if kind == "port" else REF_SO
    if kind == "port" and not os.path.isfile(path):
        build_port()
    lib = C.CDLL(path)
    lib.oracle_kind.restype = C.c_char_p
    vp, ip = C.c_void_p, C.c_int
    lib.oracle_render.argtypes = [C.POINTER(abi.FrameParams), vp, vp, vp, ip, vp, ip, ip,
                                  ip, ip, ip, ip, ip, ip, vp, vp, vp, vp, vp, ip]
    lib.oracle_render.restype = ip
    if kind == "port":
        lib.oracle_render_margins.argtypes = lib.oracle_render.argtypes + [vp]
        lib.oracle_render_margins.restype = ip
    assert lib.oracle_kind().decode() == kind
    _libs[kind] = lib
    return lib


_sky_cache: dict = {}


def _sky_f32(sky: np.ndarray | None):
    """(6,S,S,3|4) fp16/fp32 faces -> RGBA32F holding FP16-rounded values (what the GL_RGB16F
    cube map of src/TextureLoader.cpp:143 stores)."""
    if sky is None:
        return None
    key = id(sky)
    if key not in _sky_cache:
        s = sky.astype(np.float16).astype(np.float32)
        if s.shape[-1] == 3:
            s = np.concatenate([s, np.ones(s.shape[:-1] + (1,), np.float32)], axis=-1)
        _sky_cache.clear()
        _sky_cache[key] = np.ascontiguousarray(s)
    return _sky_cache[key]


def render(w, kind: str = "port", x0: int = 0, y0: int = 0, sx: int = 1, sy: int = 1,
           nx: int | None = None, ny: int | None = None, threads: int = 0, params: abi.FrameParams | None = None,
           margins: bool = False):
    """Runs one of the CPU oracles on a ``synthetic.Workload``. Renders pixels
    (x0+i*sx, y0+j*sy); default = the full frame. Returns a dict of (ny,nx,4) float32 arrays
    ``color``/``position``/``normal`` (row 0 = gl y 0) and ``counters``. ``margins`` (port only) adds
    ``margin`` (ny,nx): the smallest relative margin of any geometric decision on the pixel's path
    (rt_oracle.cpp margin_of_object) — what tests/parity.py classifies grazing pixels with."""
    lib = load(kind)
    p = params or workload_params(w)
    if nx is None:
        nx = (p.width - x0 + sx - 1) // sx
    if ny is None:
        ny = (p.height - y0 + sy - 1) // sy
    color = np.zeros((ny, nx, 4), np.float32)
    position = np.zeros((ny, nx, 4), np.float32)
    normal = np.zeros((ny, nx, 4), np.float32)
    object_id = np.zeros((ny, nx), np.int32)
    counters = (C.c_uint64 * abi.RT_CNT_COUNT)()
    objects = np.ascontiguousarray(w.objects)
    lights = np.ascontiguousarray(w.lights)
    sky = _sky_f32(w.skybox) if w.use_skybox else None
    noise = None if w.noise is None else np.ascontiguousarray(w.noise, dtype=np.uint8)
    args = [C.byref(p), objects.ctypes.data, lights.ctypes.data,
            None if sky is None else sky.ctypes.data, 0 if sky is None else sky.shape[1],
            None if noise is None else noise.ctypes.data,
            0 if noise is None else noise.shape[1], 0 if noise is None else noise.shape[0],
            x0, y0, sx, sy, nx, ny, color.ctypes.data, position.ctypes.data, normal.ctypes.data, object_id.ctypes.data,
            C.cast(counters, C.c_void_p), threads]
    margin = None
    if margins:
        assert kind == "port", "grazing margins come from the instrumented port (bit-identical to the reference build)"
        margin = np.full((ny, nx), np.inf, np.float32)
        rc = lib.oracle_render_margins(*args, margin.ctypes.data)
    else:
        rc = lib.oracle_render(*args)
    if rc != 0:
        raise RuntimeError(f"oracle_render({kind}) failed: {rc}")
    out = {"color": color, "position": position, "normal": normal, "object_id": object_id,
           "counters": {k: int(counters[i]) for i, k in enumerate(abi.COUNTER_NAMES)}}
    if margin is not None:
        out["margin"] = margin
    return out


# --------------------------------------------------------------------------- f1-f3 stages (post_api.h)
POST_PORT_SO = os.path.join(HERE, "librt_post_oracle.so")
POST_REF_SO = os.path.join(HERE, "_ref", "libref_post.so")


def build_post_port(force: bool = False) -> str:
    src = os.path.join(HERE, "post_oracle.cpp")
    deps = [src, os.path.join(HERE, "post_api.h"), os.path.join(HERE, "..", "include", "rt_abi.h")]
    if not force and os.path.isfile(POST_PORT_SO) and all(os.path.getmtime(POST_PORT_SO) >= os.path.getmtime(d) for d in deps):
        return POST_PORT_SO
    subprocess.run(["g++", "-O2", "-std=c++17", "-fopenmp", "-ffp-contract=off", "-fno-fast-math", "-fPIC",
                    "-shared", "-Wall", "-o", POST_PORT_SO, src], check=True)
    _libs.pop("post_port", None)
    return POST_PORT_SO


def build_post_ref(force: bool = False) -> str | None:
    """oracle/_ref/libref_post.so from the reference's fragment shaders, when the tree is present."""
    from . import make_ref
    srcs = make_ref.post_sources()
    if not all(os.path.isfile(f) for f in srcs):
        return POST_REF_SO if os.path.isfile(POST_REF_SO) else None
    deps = srcs + [os.path.join(HERE, f) for f in ("post_harness.cpp", "post_shim.h", "glsl_shim.h", "post_api.h", "make_ref.py")]
    if not force and os.path.isfile(POST_REF_SO) and all(os.path.getmtime(POST_REF_SO) >= os.path.getmtime(d) for d in deps):
        return POST_REF_SO
    make_ref.build_post()
    _libs.pop("post_reference", None)
    return POST_REF_SO


def have_post(kind: str) -> bool:
    return os.path.isfile(POST_PORT_SO if kind == "port" else POST_REF_SO)


def load_post(kind: str = "port") -> C.CDLL:
    key = "post_" + kind
    if key in _libs:
        return _libs[key]
    if kind == "port" and not os.path.isfile(POST_PORT_SO):
        build_post_port()
    lib = C.CDLL(POST_PORT_SO if kind == "port" else POST_REF_SO)
    lib.post_oracle_kind.restype = C.c_char_p
    vp, ip, fl = C.c_void_p, C.c_int, C.c_float
    lib.post_bloom_extract.argtypes = [vp, ip, ip, fl, vp]
    lib.post_gaussian_blur.argtypes = [vp, ip, ip, ip, vp]
    lib.post_bloom_combine.argtypes = [vp, vp, ip, ip, fl, vp]
    lib.post_taa.argtypes = [vp, vp, vp, ip, ip, fl, fl, fl, vp]
    lib.post_equirect_to_cubemap.argtypes = [vp, ip, ip, ip, vp]
    lib.post_generate_aabbs.argtypes = [vp, ip]
    lib.post_decode_rgbe.argtypes = [vp, C.c_size_t, vp]
    assert lib.post_oracle_kind().decode() == kind
    _libs[key] = lib
    return lib


def display_rgba8(display: np.ndarray) -> np.ndarray:
    """The window's default framebuffer (8-bit UNORM; no sRGB enable anywhere in the reference's
    src/) as it stores the combine pass's colour (src/ForwardShadingPipeline.cpp:220-230):
    OpenGL 4.3 section 2.3.4.2 / 17.3.9 — clamp to [0, 1], scale by 255 in FP32, round to nearest (GL leaves
    the tie rule to the implementation: ties to even here), NaN -> 0. Checker of
    rt_frame_read_display(RT_DISPLAY_RGBA8)."""
    x = np.ascontiguousarray(display, dtype=np.float32)
    x = np.where(np.isnan(x), np.float32(0), x)
    x = np.clip(x, np.float32(0), np.float32(1)).astype(np.float32)
    return np.rint(x * np.float32(255)).astype(np.uint8)


def _img(a: np.ndarray, ch: int = 4) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.float32)
    assert a.ndim == 3 and a.shape[2] == ch, a.shape
    return a


class Post:
    """numpy front end of one post oracle ("port" or "reference"). RGBA16F images are float32
    arrays holding half-rounded values."""

    def __init__(self, kind: str = "port"):
        self.kind, self.lib = kind, load_post(kind)

    def bloom_extract(self, color, threshold=1.0):
        color = _img(color); out = np.empty_like(color)
        assert self.lib.post_bloom_extract(color.ctypes.data, color.shape[1], color.shape[0], threshold, out.ctypes.data) == 0
        return out

    def gaussian_blur(self, img, horizontal: bool):
        img = _img(img); out = np.empty_like(img)
        assert self.lib.post_gaussian_blur(img.ctypes.data, img.shape[1], img.shape[0], 1 if horizontal else 0, out.ctypes.data) == 0
        return out

    def bloom_combine(self, color, bloom, strength=0.5):
        color, bloom = _img(color), _img(bloom); out = np.empty_like(color)
        assert self.lib.post_bloom_combine(color.ctypes.data, bloom.ctypes.data, color.shape[1], color.shape[0], strength, out.ctypes.data) == 0
        return out

    def taa(self, color, history, normal, blend, jx, jy):
        color, history, normal = _img(color), _img(history), _img(normal); out = np.empty_like(color)
        assert self.lib.post_taa(color.ctypes.data, history.ctypes.data, normal.ctypes.data, color.shape[1], color.shape[0],
                                 blend, jx, jy, out.ctypes.data) == 0
        return out

    def post_chain(self, color, normal, history, threshold=1.0, strength=0.5, passes=10, taa=True, blend=0.1, jx=0.0, jy=0.0):
        """src/ForwardShadingPipeline.cpp:185-254 -> (display, new history or None)."""
        a = self.bloom_extract(color, threshold)
        horizontal = True
        for _ in range(passes):
            a = self.gaussian_blur(a, horizontal)
            horizontal = not horizontal
        display = self.bloom_combine(color, a, strength)
        return display, (self.taa(color, history, normal, blend, jx, jy) if taa else None)

    def equirect_to_cubemap(self, equirect_rgb, face_size):
        e = _img(equirect_rgb, 3); out = np.empty((6, face_size, face_size, 4), np.float32)
        assert self.lib.post_equirect_to_cubemap(e.ctypes.data, e.shape[1], e.shape[0], face_size, out.ctypes.data) == 0
        return out

    def generate_aabbs(self, objects: np.ndarray) -> np.ndarray:
        o = np.frombuffer(bytearray(np.ascontiguousarray(objects).tobytes()), dtype=objects.dtype)   # keeps the padding bytes
        assert self.lib.post_generate_aabbs(o.ctypes.data, len(o)) == 0, "port only"
        return o

    def decode_rgbe(self, rgbe: np.ndarray) -> np.ndarray:
        q = np.ascontiguousarray(rgbe, dtype=np.uint8).reshape(-1, 4)
        out = np.empty((len(q), 3), np.float32)
        assert self.lib.post_decode_rgbe(q.ctypes.data, len(q), out.ctypes.data) == 0, "port only"
        return out
