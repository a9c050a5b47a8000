"""GPU parity of the f1-f3 stages (include/rt_post.h) against the CPU oracle port, through the C
ABI. Bar: bit-exact for bloom extract / blur / combine / TAA / AABBs / RGBE (pure IEEE FP32
multiply-add-divide-sqrt, unfused on both sides); the equirect -> cube pass goes through
atan2f/asinf (CUDA: 2 ulp, glibc: < 1 ulp), so a few half-precision texels may land on the
neighbouring half or filter-weight step: >= 99 % identical, the rest within 2^-8 relative."""
import os

import numpy as np
import pytest

import oracle
from opengl_raytracing_b200 import abi, post as P, scene as S, synthetic

pytestmark = pytest.mark.gpu

from test_post_oracle import post_inputs, same  # noqa: E402


@pytest.fixture(scope="module")
def ctx():
    from opengl_raytracing_b200.pipeline import RtContext
    c = RtContext(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def port():
    oracle.build_post_port()
    return oracle.Post("port")


def dev(a, dtype=None):
    import torch
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return t if dtype is None else t.to(dtype)


def host(t):
    return t.float().cpu().numpy()


@pytest.mark.parametrize("shape", [(29, 45), (1, 1), (2, 33), (40, 7), (135, 240), (270, 481)])
def test_each_pass_is_bit_identical_to_the_oracle(ctx, port, shape):
    import torch
    h, w = shape
    color, normal, history = post_inputs(13, h, w)
    pc = P.PostChain(ctx)
    p = P.post_params(w, h)
    d_color = dev(color)
    a = pc.bloom_extract(p, d_color)
    want_a = port.bloom_extract(color, 1.0)
    assert same(host(a), want_a)
    for horizontal in (True, False):
        assert same(host(pc.gaussian_blur(p, a, horizontal)), port.gaussian_blur(want_a, horizontal))
    assert same(host(pc.bloom_combine(p, d_color, a)), port.bloom_combine(color, want_a, 0.5))
    d_hist, d_normal = dev(history), dev(normal, torch.float16)
    for frame in (0, 3, 7):
        jx, jy = P.taa_jitter(frame, w, h)
        p2 = P.post_params(w, h, blend=0.1, jitter=(jx, jy))
        assert same(host(pc.taa_resolve(p2, d_color, d_hist, d_normal)), port.taa(color, history, normal, 0.1, jx, jy))
    p3 = P.post_params(w, h, blend=0.5, jitter=(-0.7 / w, 1.3 / h))
    assert same(host(pc.taa_resolve(p3, d_color, d_hist, d_normal)), port.taa(color, history, normal, 0.5, -0.7 / w, 1.3 / h))


@pytest.mark.parametrize("flags", [0, abi.POST_LITERAL_PASSES])
@pytest.mark.parametrize("shape,passes", [((29, 45), 10), ((135, 240), 10), ((64, 64), 3), ((37, 19), 0), ((270, 481), 10)])
def test_post_chain_is_bit_identical_to_the_oracle(ctx, port, shape, passes, flags):
    import torch
    h, w = shape
    color, normal, history = post_inputs(17, h, w)
    pc = P.PostChain(ctx)
    jx, jy = P.taa_jitter(5, w, h)
    p = P.post_params(w, h, passes=passes, blend=0.1, jitter=(jx, jy), flags=flags)
    before = pc.launch_count()
    display, hist = pc.post_chain(p, dev(color), dev(normal, torch.float16), dev(history))
    assert pc.launch_count() > before
    want_d, want_h = port.post_chain(color, normal, history, passes=passes, blend=0.1, jx=jx, jy=jy)
    assert same(host(display), want_d)
    assert same(host(hist), want_h)
    # TAA off: display only, history pointers may be null
    p.taa_enabled = 0
    display2, none = pc.post_chain(p, dev(color), None, None)
    assert none is None and same(host(display2), want_d)


def test_post_chain_on_a_rendered_frame_matches_golden_free_properties(ctx, port):
    """The chain consumes the ray tracer's three images directly (device to device)."""
    import torch
    from opengl_raytracing_b200.pipeline import render_workload
    w = synthetic.config(2, 320, 180)
    k = render_workload(ctx, w)
    pc = P.PostChain(ctx)
    jx, jy = P.taa_jitter(0, 320, 180)
    p = P.post_params(320, 180, jitter=(jx, jy))
    hist0 = np.zeros_like(k["color"])
    display, hist = pc.post_chain(p, dev(k["color"]), dev(k["normal"], torch.float16), dev(hist0))
    want_d, want_h = port.post_chain(k["color"], k["normal"], hist0, jx=jx, jy=jy)
    assert same(host(display), want_d) and same(host(hist), want_h)
    assert (host(display)[..., :3] >= k["color"][..., :3] - 1e-6).all()      # bloom only adds light


def test_invalid_arguments_fail_loudly(ctx):
    import torch
    pc = P.PostChain(ctx)
    p = P.post_params(0, 4)
    with pytest.raises(abi.RtError):
        pc.post_chain(p, torch.zeros(4, 4, 4, device="cuda"), None, None, display=torch.zeros(4, 4, 4, device="cuda"))
    p = P.post_params(4, 4)
    img = torch.zeros(4, 4, 4, dtype=torch.float16, device="cuda")
    with pytest.raises(abi.RtError):                      # aliased blur
        ctx._check(ctx.lib.rt_gaussian_blur(ctx.handle, __import__("ctypes").byref(p), img.data_ptr(), img.data_ptr(), 1, None))
    with pytest.raises(abi.RtError):                      # TAA on, no history
        pc.post_chain(p, torch.zeros(4, 4, 4, device="cuda"), img, None)


def test_aabbs_on_device_are_bit_identical(ctx, port):
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scenes.npz"))
    for name in ("default", "SIMPLE", "performance_test"):
        objs = g[name + "_objects"].copy().view(S.OBJECT_DTYPE)
        want = objs.tobytes()
        objs["bounds"]["min"] = 7
        objs["bounds"]["max"] = -7
        assert P.generate_aabbs(ctx, objs).tobytes() == want
    for cfg in (3, 5):
        w = synthetic.config(cfg)
        blank = np.frombuffer(bytearray(w.objects.tobytes()), dtype=S.OBJECT_DTYPE)
        blank["bounds"]["min"] = 0
        blank["bounds"]["max"] = 0
        assert P.generate_aabbs(ctx, blank).tobytes() == w.objects.tobytes()
    # tilted planes, degenerate normals, unknown types keep their bounds
    rng = np.random.default_rng(2)
    o = S.new_objects(4096)
    o["type"] = rng.integers(0, 3, size=4096)
    o["position"] = rng.standard_normal((4096, 3)) * 10
    o["radius"] = rng.random(4096) * 3
    n = rng.standard_normal((4096, 3)); n /= np.linalg.norm(n, axis=1, keepdims=True)
    n[:7] = [(0, 1, 0), (0, -1, 0), (0, 0, 0), (0, 0.9, 0.43589), (0, 0.91, 0.4146), (1, 0, 0), (0, 0, 1)]
    o["normal"] = n
    o["size"] = rng.random((4096, 2)) * 20
    o["bounds"]["min"] = 3
    o["bounds"]["max"] = 5
    got, want = P.generate_aabbs(ctx, o), port.generate_aabbs(o)
    assert np.array_equal(got["bounds"]["min"], want["bounds"]["min"], equal_nan=True)
    assert np.array_equal(got["bounds"]["max"], want["bounds"]["max"], equal_nan=True)
    other = o["type"] == 2
    assert (got["bounds"]["min"][other] == 3).all() and (got["bounds"]["max"][other] == 5).all()


def float_to_rgbe(rgb):
    """Radiance encoding (shared exponent of the largest channel, mantissas truncated)."""
    m = rgb.max(axis=-1)
    e = np.where(m > 1e-32, np.floor(np.log2(np.maximum(m, 1e-38))) + 1, -128).astype(np.int32)
    mant = np.floor(rgb / np.exp2(e)[..., None] * 256.0).clip(0, 255)
    out = np.concatenate([mant, (e + 128)[..., None]], axis=-1).astype(np.uint8)
    out[m <= 1e-32] = 0
    return out


def test_rgbe_decode_and_equirect_to_cubemap(ctx, port):
    rng = np.random.default_rng(4)
    q = rng.integers(0, 256, size=(64, 128, 4), dtype=np.uint8)          # arbitrary bytes: exact decode
    q[0, :9, 3] = 0
    assert np.array_equal(P.decode_rgbe(ctx, q), port.decode_rgbe(q))
    # a smooth sky (gradient + sun) through the whole f2 path: .hdr bytes -> RGBE -> float -> cube
    h, w, S_ = 128, 256, 64
    v, u = np.meshgrid((np.arange(h) + 0.5) / h, (np.arange(w) + 0.5) / w, indexing="ij")
    sky = np.stack([0.3 + 0.5 * v, 0.4 + 0.4 * v, 0.9 - 0.3 * v], axis=-1)
    sky += 40.0 * np.exp(-((u - 0.3) ** 2 + (v - 0.7) ** 2) / 0.002)[..., None]
    texels, ww, hh = P.read_hdr(P.write_hdr(float_to_rgbe(sky)))
    assert (ww, hh) == (w, h)
    rgb = P.decode_rgbe(ctx, texels).reshape(h, w, 3)
    assert np.array_equal(rgb, port.decode_rgbe(texels).reshape(h, w, 3))
    assert np.allclose(rgb, sky, rtol=2.0 ** -6, atol=1e-6)                # 8-bit shared-exponent mantissas
    equirect = np.ascontiguousarray(rgb[::-1])                             # stbi_set_flip_vertically_on_load(true)
    got = P.equirect_to_cubemap(ctx, equirect, S_).astype(np.float32)
    want = port.equirect_to_cubemap(equirect, S_)
    assert got.shape == want.shape == (6, S_, S_, 4)
    assert (got == want).mean() >= 0.99
    assert np.allclose(got, want, rtol=2.0 ** -8, atol=2.0 ** -14)
    assert (got[..., 3] == 1).all()
    # the faces feed rt_set_skybox unchanged
    ctx.set_skybox(got.astype(np.float16))
    ctx.set_skybox(None)
