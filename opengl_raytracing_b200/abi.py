"""ctypes binding of librt_b200.so (include/rt_abi.h) — the reference-side stub.

This is the Python spelling of the binding INTEGRATION.md shows for the C++ host: every
function below is one ``extern "C"`` entry point. There is no fallback: if the CUDA
library is missing or fails to load, importing callers get ``RtLibraryError``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "csrc", "librt_b200.so")   # env: dev A/B builds only

RT_OK, RT_ERR_INVALID_ARG, RT_ERR_CUDA, RT_ERR_NCCL, RT_ERR_OOM, RT_ERR_UNAVAILABLE = range(6)

FLAG_NORMAL_FP32 = 0x01
FLAG_PARITY_MATH = 0x02
FLAG_SKYBOX_HW = 0x04
FLAG_COUNTERS = 0x08
FLAG_ACCUMULATE = 0x10
FLAG_LITERAL_KERNEL = 0x20
FLAG_SUM_SAMPLES = 0x40
HOST_PATH_AUTO, HOST_PATH_STAGED, HOST_PATH_BANDED, HOST_PATH_ZEROCOPY = range(4)

COUNTER_NAMES = [
    "rays", "aabb", "sphere", "plane_early", "plane_full", "hit_update_sphere", "hit_update_plane",
    "light_point", "light_dir", "light_area", "pbr", "pcf_setup", "pcf_sample", "pcss_probe",
    "sss_sample", "bounce_mirror", "bounce_refract", "bounce_diffuse", "rr", "camera", "miss",
]
RT_CNT_COUNT = len(COUNTER_NAMES)

EXPORTED_SYMBOLS = [
    "rt_abi_version", "rt_status_string", "rt_create", "rt_destroy", "rt_last_error", "rt_upload_scene",
    "rt_set_skybox", "rt_set_noise", "rt_render", "rt_render_host", "rt_synchronize", "rt_render_gl",
    "rt_get_counters", "rt_set_host_path", "rt_device_alloc", "rt_device_free", "rt_ipc_export",
    "rt_ipc_open", "rt_ipc_close", "rt_host_alloc", "rt_host_free", "rt_measure_fp32_peak", "rt_launch_count",
    "rt_debug_light_grids", "rt_render_static", "rt_tune", "rt_get_tuning", "rt_set_tuning", "rt_set_autotune", "rt_set_sm_reserve",
]
KERNEL_AUTO, KERNEL_MEGAKERNEL, KERNEL_WAVEFRONT, KERNEL_LITERAL, KERNEL_REGROUP = -1, 0, 1, 2, 3
TUNING_DEFAULT, TUNING_TRIALS, TUNING_CHOSEN, TUNING_PINNED = range(4)
KERNEL_NAMES = {-1: "auto", 0: "megakernel", 1: "wavefront", 2: "literal", 3: "regroup"}
KERNEL_IDS = {v: k for k, v in KERNEL_NAMES.items()}
POST_EXPORTED_SYMBOLS = [                   # include/rt_post.h
    "rt_generate_aabbs", "rt_generate_aabbs_device", "rt_decode_rgbe", "rt_equirect_to_cubemap",
    "rt_bloom_extract", "rt_gaussian_blur", "rt_bloom_combine", "rt_taa_resolve", "rt_post_chain", "rt_post_launch_count",
]


FRAME_EXPORTED_SYMBOLS = [                  # include/rt_frame.h
    "rt_frame_create", "rt_frame_destroy", "rt_frame_render", "rt_frame_images", "rt_frame_read_display",
    "rt_frame_render_host", "rt_frame_launch_count",
]
DISPLAY_RGBA32F, DISPLAY_RGBA8 = 0, 1
MULTI_EXPORTED_SYMBOLS = [                  # include/rt_multi.h
    "rt_multi_create", "rt_multi_destroy", "rt_multi_last_error", "rt_multi_size", "rt_multi_context", "rt_multi_upload_scene",
    "rt_multi_set_skybox", "rt_multi_set_noise", "rt_multi_render", "rt_multi_synchronize", "rt_multi_render_host",
    "rt_multi_launch_count", "rt_peer_reduce", "rt_peer_reduce_status",
]
MULTI_ROWS, MULTI_SAMPLES = 0, 1
RT_MAX_PEERS, RT_PEER_FLAG_WORDS = 16, 64


class RtLibraryError(RuntimeError):
    pass


class RtError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"rt_status {status}: {msg}")
        self.status = status


class FrameParams(C.Structure):
    """rt_frame_params (include/rt_abi.h) == the shader's uniforms, glsl:72-89."""
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32),
        ("row_begin", C.c_int32), ("row_end", C.c_int32),
        ("num_objects", C.c_int32), ("num_lights", C.c_int32),
        ("camera_pos", C.c_float * 3), ("camera_dir", C.c_float * 3),
        ("camera_up", C.c_float * 3), ("camera_right", C.c_float * 3),
        ("fov_deg", C.c_float), ("focal_length", C.c_float), ("max_ray_distance", C.c_float),
        ("use_skybox", C.c_int32),
        ("noise_scale", C.c_float * 2),
        ("frame_count", C.c_int32), ("num_samples", C.c_int32), ("max_ray_depth", C.c_int32),
        ("flags", C.c_uint32),
        ("row_band", C.c_int32), ("row_period", C.c_int32), ("row_phase", C.c_int32),
    ]


class Tuning(C.Structure):
    """rt_tuning (include/rt_abi.h): which build of the pass runs, and what was measured."""
    _fields_ = [
        ("kernel", C.c_int32), ("registers", C.c_int32), ("phase_level", C.c_int32), ("state", C.c_int32),
        ("num_variants", C.c_int32), ("variant_ms", C.c_float * 8),
        ("variant_kernel", C.c_int32 * 8), ("variant_registers", C.c_int32 * 8), ("variant_phase_level", C.c_int32 * 8),
    ]

    def as_dict(self) -> dict:
        return {"kernel": KERNEL_NAMES.get(self.kernel, str(self.kernel)), "registers": 80 if self.registers else 128,
                "phase_level": self.phase_level, "state": ("default", "trials", "chosen", "pinned")[self.state],
                "variants": [{"kernel": KERNEL_NAMES.get(self.variant_kernel[i], "?"), "registers": 80 if self.variant_registers[i] else 128,
                              "phase_level": self.variant_phase_level[i], "ms": round(float(self.variant_ms[i]), 4)}
                             for i in range(self.num_variants)]}


class PeerReduceDesc(C.Structure):
    """rt_peer_reduce_desc (include/rt_multi.h)."""
    _fields_ = [
        ("rank", C.c_int32), ("world", C.c_int32),
        ("partial", C.c_void_p * 16), ("result", C.c_void_p), ("flags", C.c_void_p * 16),
        ("pixels", C.c_uint64), ("scale", C.c_float), ("epoch", C.c_uint32),
    ]


class PostParams(C.Structure):
    """rt_post_params (include/rt_post.h): the post chain's uniforms, src/ForwardShadingPipeline.cpp:185-254."""
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32),
        ("bloom_threshold", C.c_float), ("bloom_strength", C.c_float),
        ("blur_passes", C.c_int32), ("taa_enabled", C.c_int32),
        ("taa_blend", C.c_float), ("jitter_x", C.c_float), ("jitter_y", C.c_float),
        ("flags", C.c_uint32),
    ]


POST_LITERAL_PASSES = 0x01

_lib = None


def load_library(path: str | None = None) -> C.CDLL:
    """dlopen librt_b200.so and declare every prototype. Fails loudly."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.isfile(p):
        raise RtLibraryError(
            f"{p} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback.")
    try:
        lib = C.CDLL(p, mode=C.RTLD_GLOBAL)
    except OSError as e:  # missing libcudart etc.
        raise RtLibraryError(f"cannot load {p}: {e}") from e
    vp, ip = C.c_void_p, C.c_int
    lib.rt_abi_version.restype = ip
    lib.rt_status_string.restype = C.c_char_p
    lib.rt_status_string.argtypes = [ip]
    lib.rt_create.argtypes = [C.POINTER(vp), ip]
    lib.rt_destroy.argtypes = [vp]
    lib.rt_destroy.restype = None
    lib.rt_last_error.argtypes = [vp]
    lib.rt_last_error.restype = C.c_char_p
    lib.rt_upload_scene.argtypes = [vp, vp, ip, vp, ip, vp, ip]
    lib.rt_set_skybox.argtypes = [vp, vp, ip, ip, ip]
    lib.rt_set_noise.argtypes = [vp, vp, ip, ip]
    lib.rt_render.argtypes = [vp, C.POINTER(FrameParams), vp, vp, vp, vp, vp]
    lib.rt_render_static.argtypes = [vp, C.POINTER(FrameParams), vp, vp, vp, vp, vp]
    lib.rt_render_host.argtypes = [vp, C.POINTER(FrameParams), vp, ip, vp, ip, vp, vp, vp, vp]
    lib.rt_synchronize.argtypes = [vp]
    lib.rt_set_host_path.argtypes = [vp, ip]
    lib.rt_render_gl.argtypes = [vp, C.POINTER(FrameParams), C.c_uint, C.c_uint, C.c_uint]
    lib.rt_get_counters.argtypes = [vp, C.POINTER(C.c_uint64), ip]
    lib.rt_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
    lib.rt_device_alloc.argtypes = [vp, C.POINTER(vp), C.c_size_t]
    lib.rt_device_free.argtypes = [vp, vp]
    lib.rt_ipc_export.argtypes = [vp, vp, C.c_char_p]
    lib.rt_ipc_open.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
    lib.rt_ipc_close.argtypes = [vp, vp]
    lib.rt_host_free.argtypes = [vp]
    lib.rt_measure_fp32_peak.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(ip), C.POINTER(C.c_double)]
    lib.rt_launch_count.argtypes = [vp]
    lib.rt_launch_count.restype = C.c_uint64
    lib.rt_debug_light_grids.argtypes = [vp, ip, vp, ip, vp, C.c_size_t, C.POINTER(ip), C.POINTER(ip), vp]
    lib.rt_debug_light_grids.restype = C.c_size_t
    lib.rt_tune.argtypes = [vp, C.POINTER(FrameParams), C.POINTER(Tuning)]
    lib.rt_get_tuning.argtypes = [vp, C.POINTER(FrameParams), C.POINTER(Tuning)]
    lib.rt_set_tuning.argtypes = [vp, C.POINTER(Tuning)]
    lib.rt_set_autotune.argtypes = [vp, ip]
    lib.rt_set_sm_reserve.argtypes = [vp, ip]
    # include/rt_post.h
    pp = C.POINTER(PostParams)
    lib.rt_generate_aabbs.argtypes = [vp, vp, ip]
    lib.rt_generate_aabbs_device.argtypes = [vp, vp, ip, vp]
    lib.rt_decode_rgbe.argtypes = [vp, vp, vp, C.c_size_t, vp]
    lib.rt_equirect_to_cubemap.argtypes = [vp, vp, ip, ip, vp, ip, vp]
    lib.rt_bloom_extract.argtypes = [vp, pp, vp, vp, vp]
    lib.rt_gaussian_blur.argtypes = [vp, pp, vp, vp, ip, vp]
    lib.rt_bloom_combine.argtypes = [vp, pp, vp, vp, vp, vp]
    lib.rt_taa_resolve.argtypes = [vp, pp, vp, vp, vp, vp, vp]
    lib.rt_post_chain.argtypes = [vp, pp, vp, vp, vp, vp, vp, vp]
    lib.rt_post_launch_count.argtypes = [vp]
    lib.rt_post_launch_count.restype = C.c_uint64
    # include/rt_frame.h
    fp = C.POINTER(FrameParams)
    lib.rt_frame_create.argtypes = [vp, ip, ip, C.POINTER(vp)]
    lib.rt_frame_destroy.argtypes = [vp]
    lib.rt_frame_destroy.restype = None
    lib.rt_frame_render.argtypes = [vp, fp, pp, vp]
    lib.rt_frame_images.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    lib.rt_frame_read_display.argtypes = [vp, vp, ip, vp]
    lib.rt_frame_render_host.argtypes = [vp, fp, pp, vp, ip, vp, ip, vp, ip]
    lib.rt_frame_launch_count.argtypes = [vp]
    lib.rt_frame_launch_count.restype = C.c_uint64
    # include/rt_multi.h
    lib.rt_multi_create.argtypes = [C.POINTER(vp), C.POINTER(ip), ip]
    lib.rt_multi_destroy.argtypes = [vp]
    lib.rt_multi_destroy.restype = None
    lib.rt_multi_last_error.argtypes = [vp]
    lib.rt_multi_last_error.restype = C.c_char_p
    lib.rt_multi_size.argtypes = [vp]
    lib.rt_multi_context.argtypes = [vp, ip]
    lib.rt_multi_context.restype = vp
    lib.rt_multi_upload_scene.argtypes = [vp, vp, ip, vp, ip]
    lib.rt_multi_set_skybox.argtypes = [vp, vp, ip, ip, ip]
    lib.rt_multi_set_noise.argtypes = [vp, vp, ip, ip]
    lib.rt_multi_render.argtypes = [vp, fp, ip, vp, vp, vp, vp]
    lib.rt_multi_synchronize.argtypes = [vp]
    lib.rt_multi_render_host.argtypes = [vp, fp, ip, vp, ip, vp, ip, vp, vp, vp, vp]
    lib.rt_multi_launch_count.argtypes = [vp]
    lib.rt_multi_launch_count.restype = C.c_uint64
    lib.rt_peer_reduce.argtypes = [vp, C.POINTER(PeerReduceDesc), vp]
    lib.rt_peer_reduce_status.argtypes = [vp, vp]
    if path is None:
        _lib = lib
    return lib
