"""Host-side mirror of the stages either side of the ray-tracing pass (include/rt_post.h):
the post chain of ``ForwardShadingPipline::Render`` (src/ForwardShadingPipeline.cpp:185-254),
``ConvertHDRToCubemap`` (src/TextureLoader.cpp:118-194) and the per-object
``GenerateAABBForObject`` loop (src/SceneIO.h:75-104). Images are torch CUDA tensors
(H, W, 4): float32 for RGBA32F, float16 for RGBA16F; row 0 = GL y 0. No CPU fallback."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import abi, scene as S
from .pipeline import RtContext, _ptr


def halton_sequence(index: int, base: int) -> float:
    """``haltonSequence`` of the host (src/global.cpp), float arithmetic."""
    f, r = np.float32(1.0), np.float32(0.0)
    while index > 0:
        f = np.float32(f / np.float32(base))
        r = np.float32(r + f * np.float32(index % base))
        index //= base
    return float(r)


def taa_jitter(frame_count: int, width: int, height: int) -> tuple:
    """uJitterX / uJitterY as the dispatch site computes them (:242-243)."""
    return (np.float32(halton_sequence(frame_count % 8, 2) * 0.5 / width).item(),
            np.float32(halton_sequence(frame_count % 8, 3) * 0.5 / height).item())


def post_params(width: int, height: int, threshold: float = 1.0, strength: float = 0.5, passes: int = 10,
                taa: bool = True, blend: float = 0.1, jitter=(0.0, 0.0), flags: int = 0) -> abi.PostParams:
    p = abi.PostParams()
    p.width, p.height = width, height
    p.bloom_threshold, p.bloom_strength, p.blur_passes = threshold, strength, passes
    p.taa_enabled, p.taa_blend = 1 if taa else 0, blend
    p.jitter_x, p.jitter_y = jitter
    p.flags = flags
    return p


class PostChain:
    """Bloom + TAA behind an ``RtContext``. Method names follow the shaders."""

    def __init__(self, ctx: RtContext):
        import torch
        self._t, self.ctx, self.lib = torch, ctx, ctx.lib
        self.dev = f"cuda:{ctx.device}"

    def _stream(self):
        return C.c_void_p(self._t.cuda.current_stream(self.ctx.device).cuda_stream)

    def _new(self, p, dtype):
        return self._t.empty(p.height, p.width, 4, dtype=dtype, device=self.dev)

    def bloom_extract(self, p, color):
        out = self._new(p, self._t.float16)
        self.ctx._check(self.lib.rt_bloom_extract(self.ctx.handle, C.byref(p), _ptr(color), _ptr(out), self._stream()))
        return out

    def gaussian_blur(self, p, img, horizontal: bool):
        out = self._new(p, self._t.float16)
        self.ctx._check(self.lib.rt_gaussian_blur(self.ctx.handle, C.byref(p), _ptr(img), _ptr(out), 1 if horizontal else 0, self._stream()))
        return out

    def bloom_combine(self, p, color, bloom):
        out = self._new(p, self._t.float32)
        self.ctx._check(self.lib.rt_bloom_combine(self.ctx.handle, C.byref(p), _ptr(color), _ptr(bloom), _ptr(out), self._stream()))
        return out

    def taa_resolve(self, p, color, history_prev, normal):
        out = self._new(p, self._t.float32)
        self.ctx._check(self.lib.rt_taa_resolve(self.ctx.handle, C.byref(p), _ptr(color), _ptr(history_prev), _ptr(normal), _ptr(out), self._stream()))
        return out

    def post_chain(self, p, color, normal, history_prev=None, display=None, history_out=None):
        """-> (display RGBA32F, new history RGBA32F or None)."""
        display = self._new(p, self._t.float32) if display is None else display
        if p.taa_enabled and history_out is None:
            history_out = self._new(p, self._t.float32)
        self.ctx._check(self.lib.rt_post_chain(self.ctx.handle, C.byref(p), _ptr(color), _ptr(normal), _ptr(history_prev),
                                               _ptr(history_out), _ptr(display), self._stream()))
        return display, (history_out if p.taa_enabled else None)

    def launch_count(self) -> int:
        return int(self.lib.rt_post_launch_count(self.ctx.handle))


def generate_aabbs(ctx: RtContext, objects: np.ndarray) -> np.ndarray:
    """``GenerateAABBForObject`` for every record, on the GPU; returns a copy with ``bounds`` filled."""
    o = np.frombuffer(bytearray(np.ascontiguousarray(objects, dtype=S.OBJECT_DTYPE).tobytes()), dtype=S.OBJECT_DTYPE)   # keeps padding
    ctx._check(ctx.lib.rt_generate_aabbs(ctx.handle, o.ctypes.data, len(o)))
    return o


def equirect_to_cubemap(ctx: RtContext, equirect_rgb: np.ndarray, face_size: int = 512) -> np.ndarray:
    """``ConvertHDRToCubemap`` after the file decode: (h, w, 3) float32 -> (6, S, S, 4) float16,
    ready for ``RtContext.set_skybox``."""
    import torch
    e = torch.from_numpy(np.ascontiguousarray(equirect_rgb, dtype=np.float32)).to(f"cuda:{ctx.device}")
    faces = torch.empty(6, face_size, face_size, 4, dtype=torch.float16, device=e.device)
    stream = C.c_void_p(torch.cuda.current_stream(ctx.device).cuda_stream)
    ctx._check(ctx.lib.rt_equirect_to_cubemap(ctx.handle, _ptr(e), e.shape[1], e.shape[0], _ptr(faces), face_size, stream))
    torch.cuda.synchronize(ctx.device)
    return faces.cpu().numpy()


def decode_rgbe(ctx: RtContext, rgbe: np.ndarray) -> np.ndarray:
    """(n, 4) uint8 RGBE texels -> (n, 3) float32 on the GPU."""
    import torch
    q = torch.from_numpy(np.ascontiguousarray(rgbe, dtype=np.uint8).reshape(-1, 4)).to(f"cuda:{ctx.device}")
    out = torch.empty(q.shape[0], 3, dtype=torch.float32, device=q.device)
    stream = C.c_void_p(torch.cuda.current_stream(ctx.device).cuda_stream)
    ctx._check(ctx.lib.rt_decode_rgbe(ctx.handle, _ptr(q), _ptr(out), q.shape[0], stream))
    torch.cuda.synchronize(ctx.device)
    return out.cpu().numpy()


# --------------------------------------------------------------------------- Radiance .hdr container
def read_hdr(data: bytes) -> tuple:
    """Radiance RGBE file -> ((h, w, 4) uint8 RGBE texels in FILE row order, w, h). Restates the
    container parse stb_image performs for ``stbi_loadf`` (header lines, ``-Y h +X w``, new-style
    per-channel RLE scanlines and flat data); the RGBE -> float step runs on the GPU
    (``decode_rgbe``). The reference then flips rows (``stbi_set_flip_vertically_on_load``):
    use ``[::-1]`` on the decoded image."""
    pos = data.find(b"\n")
    head = data[:pos]
    if head not in (b"#?RADIANCE", b"#?RGBE"):
        raise ValueError("not a Radiance HDR file")
    fmt_ok = False
    while True:
        nxt = data.find(b"\n", pos + 1)
        if nxt < 0:
            raise ValueError("truncated HDR header")
        line = data[pos + 1:nxt]
        pos = nxt
        if line == b"":
            break
        if line == b"FORMAT=32-bit_rle_rgbe":
            fmt_ok = True
    if not fmt_ok:
        raise ValueError("unsupported HDR format")
    nxt = data.find(b"\n", pos + 1)
    tok = data[pos + 1:nxt].split()
    if len(tok) != 4 or tok[0] != b"-Y" or tok[2] != b"+X":
        raise ValueError("unsupported HDR data layout")
    h, w = int(tok[1]), int(tok[3])
    buf = np.frombuffer(data, dtype=np.uint8, offset=nxt + 1)
    out = np.empty((h, w, 4), np.uint8)
    if w < 8 or w >= 32768:                                   # flat data
        out[:] = buf[:h * w * 4].reshape(h, w, 4)
        return out, w, h
    i = 0
    for y in range(h):
        if buf[i] != 2 or buf[i + 1] != 2 or (buf[i + 2] & 0x80):   # not RLE: the whole image is flat
            if y != 0:
                raise ValueError("corrupt HDR: mixed scanline encodings")
            out[:] = buf[:h * w * 4].reshape(h, w, 4)
            return out, w, h
        if ((int(buf[i + 2]) << 8) | int(buf[i + 3])) != w:
            raise ValueError("corrupt HDR: scanline width")
        i += 4
        for ch in range(4):
            x = 0
            while x < w:
                count = int(buf[i]); i += 1
                if count > 128:                               # run
                    count -= 128
                    out[y, x:x + count, ch] = buf[i]; i += 1
                else:                                         # literal
                    out[y, x:x + count, ch] = buf[i:i + count]; i += count
                x += count
    return out, w, h


def write_hdr(rgbe: np.ndarray) -> bytes:
    """(h, w, 4) uint8 RGBE -> a flat (non-RLE) Radiance file; test helper and inverse of read_hdr."""
    h, w, _ = rgbe.shape
    return b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n" % (h, w) + np.ascontiguousarray(rgbe, np.uint8).tobytes()
