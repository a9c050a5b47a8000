"""Host-side mirror of include/rt_frame.h: one whole iteration of ``ForwardShadingPipline::Render``
(src/ForwardShadingPipeline.cpp:139-272) with every intermediate image resident on the device —
ray-tracing dispatch, bloom, combine into the display image, TAA resolve into the history
ping-pong — and the presentation step (read-back of the display image as RGBA32F or as the 8-bit
UNORM the reference's window holds). 1:1 with the C entry points; no CPU fallback."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import abi
from .pipeline import RtContext, _ptr


class _Dev:
    """A device image of a frame, exposed through ``__cuda_array_interface__`` (zero-copy view)."""

    def __init__(self, ptr: int, shape: tuple, typestr: str):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 3, "strides": None}


class Frame:
    """Owns one ``rt_frame``: the images ``ForwardShadingPipline::Init`` creates (:51-137)."""

    def __init__(self, ctx: RtContext, width: int, height: int):
        self.ctx, self.lib, self.width, self.height = ctx, ctx.lib, width, height
        self.handle = C.c_void_p(0)
        ctx._check(self.lib.rt_frame_create(ctx.handle, width, height, C.byref(self.handle)))

    def close(self) -> None:
        if self.handle and self.ctx.handle:            # a frame must go before its context
            self.lib.rt_frame_destroy(self.handle)
        self.handle = C.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def render(self, rp: abi.FrameParams, pp: abi.PostParams, stream: int = 0) -> None:
        self.ctx._check(self.lib.rt_frame_render(self.handle, C.byref(rp), C.byref(pp), C.c_void_p(stream)))

    def images(self) -> dict:
        """Zero-copy torch views of the device images: color / position / display / history
        (H, W, 4) float32, normal (H, W, 4) float16."""
        import torch
        p = [C.c_void_p(0) for _ in range(5)]
        self.ctx._check(self.lib.rt_frame_images(self.handle, *[C.byref(x) for x in p]))
        shape = (self.height, self.width, 4)
        dev = f"cuda:{self.ctx.device}"
        names = ("color", "position", "normal", "display", "history")
        return {n: torch.as_tensor(_Dev(int(x.value), shape, "<f2" if n == "normal" else "<f4"), device=dev)
                for n, x in zip(names, p)}

    def read_display(self, dst=None, fmt: int = abi.DISPLAY_RGBA32F, stream: int = 0):
        """-> (H, W, 4) float32 or uint8 host array (``dst``: a pinned torch tensor / ndarray to fill)."""
        if dst is None:
            dst = np.empty((self.height, self.width, 4), dtype=np.float32 if fmt == abi.DISPLAY_RGBA32F else np.uint8)
        self.ctx._check(self.lib.rt_frame_read_display(self.handle, _ptr(dst), fmt, C.c_void_p(stream)))
        return dst

    def render_host(self, rp: abi.FrameParams, pp: abi.PostParams, objects, lights, dst, fmt: int = abi.DISPLAY_RGBA8):
        n = 0 if objects is None else len(objects)
        m = 0 if lights is None else len(lights)
        self.ctx._check(self.lib.rt_frame_render_host(self.handle, C.byref(rp), C.byref(pp), _ptr(objects), n, _ptr(lights), m,
                                                      _ptr(dst), fmt))
        return dst

    def launch_count(self) -> int:
        return int(self.lib.rt_frame_launch_count(self.handle))
