"""CPU: the post / f1 / f2 oracle port against (a) oracle/_ref/libref_post.so — the reference's own
fragment shaders compiled as C++ — when it is built, and (b) the committed golden vectors that
library produced (tests/golden/post.npz, generator: tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

import oracle
from opengl_raytracing_b200 import post as P, scene as S, synthetic

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "post.npz")


def post_inputs(seed=7, h=29, w=45):
    """Seeded HDR-ish frame: colours up to ~4 with some exact zeros, a NaN and an inf texel,
    half-precision normals (some zero, as on a primary miss) and a history image."""
    rng = np.random.default_rng(seed)
    color = (rng.random((h, w, 4), dtype=np.float32) ** 3 * 4).astype(np.float32)
    color[rng.random((h, w)) < 0.2] = 0
    color[..., 3] = 1
    color[3 % h, 5 % w, 0] = np.nan
    color[(h - 2) % h, (w - 3) % w, 1] = np.inf
    n = rng.standard_normal((h, w, 3)).astype(np.float32)
    n /= np.linalg.norm(n, axis=2, keepdims=True)
    n[rng.random((h, w)) < 0.3] = 0
    normal = np.concatenate([n, np.ones((h, w, 1), np.float32)], axis=2).astype(np.float16).astype(np.float32)
    history = (rng.random((h, w, 4), dtype=np.float32) * 4).astype(np.float32)
    history[..., 3] = 1
    return color, normal, history


def same(a, b):
    return np.array_equal(a, b, equal_nan=True)


@pytest.fixture(scope="module")
def port():
    oracle.build_post_port()
    return oracle.Post("port")


@pytest.fixture(scope="module")
def ref():
    oracle.build_post_ref()
    if not oracle.have_post("reference"):
        pytest.skip("oracle/_ref/libref_post.so not built (reference tree absent)")
    return oracle.Post("reference")


@pytest.mark.parametrize("shape", [(29, 45), (1, 1), (2, 33), (40, 7)])
def test_port_matches_reference_shaders(port, ref, shape):
    color, normal, history = post_inputs(11, *shape)
    h, w = shape
    a = port.bloom_extract(color, 1.0)
    assert same(a, ref.bloom_extract(color, 1.0))
    assert same(port.gaussian_blur(a, True), ref.gaussian_blur(a, True))
    assert same(port.gaussian_blur(a, False), ref.gaussian_blur(a, False))
    assert same(port.bloom_combine(color, a, 0.5), ref.bloom_combine(color, a, 0.5))
    for frame in range(8):
        jx, jy = P.taa_jitter(frame, w, h)
        assert same(port.taa(color, history, normal, 0.1, jx, jy), ref.taa(color, history, normal, 0.1, jx, jy))
    assert same(port.taa(color, history, normal, 0.5, -0.7 / w, 1.3 / h), ref.taa(color, history, normal, 0.5, -0.7 / w, 1.3 / h))
    d0, h0 = port.post_chain(color, normal, history, blend=0.1, jx=0.25 / w, jy=1 / 6 / h)
    d1, h1 = ref.post_chain(color, normal, history, blend=0.1, jx=0.25 / w, jy=1 / 6 / h)
    assert same(d0, d1) and same(h0, h1)


def test_port_equirect_matches_reference_shader(port, ref):
    rng = np.random.default_rng(3)
    e = (rng.random((48, 96, 3), dtype=np.float32) ** 4 * 50).astype(np.float32)
    assert same(port.equirect_to_cubemap(e, 24), ref.equirect_to_cubemap(e, 24))


def test_port_matches_golden_vectors(port):
    g = np.load(GOLDEN)
    color, normal, history = post_inputs()
    assert same(port.bloom_extract(color, 1.0), g["extract"])
    assert same(port.gaussian_blur(g["extract"], True), g["blur_h"])
    assert same(port.gaussian_blur(g["extract"], False), g["blur_v"])
    d, hn = port.post_chain(color, normal, history, blend=0.1, jx=float(g["jitter"][0]), jy=float(g["jitter"][1]))
    assert same(d, g["display"]) and same(hn, g["history"])
    assert same(port.equirect_to_cubemap(g["equirect"], 16), g["cube"])


def test_sampling_at_texel_centres_is_the_texel(port):
    """bloom_combine with strength 0 returns the colour image; a blur of a constant is ~constant."""
    color, _, _ = post_inputs(5, 9, 17)
    color = np.nan_to_num(color, nan=0.5, posinf=2.0)
    assert same(port.bloom_combine(color, np.zeros_like(color), 0.5)[..., :3], color[..., :3])
    flat = np.full((9, 17, 4), 0.75, np.float32)
    out = port.gaussian_blur(flat, True)
    assert np.allclose(out[..., :3], 0.75, atol=1e-3) and (out[..., 3] == 1).all()


def test_aabb_port_matches_host_mirror_and_golden_scenes(port):
    g = np.load(os.path.join(os.path.dirname(GOLDEN), "scenes.npz"))
    for name in ("default", "SIMPLE", "performance_test"):
        objs = g[name + "_objects"].copy().view(S.OBJECT_DTYPE)      # (a structured .copy() drops padding bytes)
        want = objs.tobytes()
        objs["bounds"]["min"] = 0
        objs["bounds"]["max"] = 0
        assert port.generate_aabbs(objs).tobytes() == want
    w = synthetic.config(3)
    blank = np.frombuffer(bytearray(w.objects.tobytes()), dtype=S.OBJECT_DTYPE)
    blank["bounds"]["min"] = 0
    blank["bounds"]["max"] = 0
    assert port.generate_aabbs(blank).tobytes() == w.objects.tobytes()
    assert S.generate_aabbs(blank).tobytes() == w.objects.tobytes()


def test_rgbe_port_and_hdr_container(port):
    rng = np.random.default_rng(9)
    q = rng.integers(0, 256, size=(6, 40, 4), dtype=np.uint8)
    q[0, :5, 3] = 0
    rgb = port.decode_rgbe(q)
    m = q.reshape(-1, 4).astype(np.float64)
    want = np.where(m[:, 3:4] == 0, 0.0, m[:, :3] * np.exp2(m[:, 3:4] - 136.0))
    assert np.array_equal(rgb.astype(np.float64), want)          # exact: 8-bit mantissa times a power of two
    back, w, h = P.read_hdr(P.write_hdr(q))
    assert (w, h) == (40, 6) and np.array_equal(back, q)
    # new-style RLE scanlines: a run, a literal, per channel
    w = 16
    scan = bytearray([2, 2, 0, w])
    for ch in range(4):
        scan += bytes([128 + 10, 7 + ch]) + bytes([6]) + bytes(range(ch, ch + 6))
    data = b"#?RADIANCE\n# comment\nFORMAT=32-bit_rle_rgbe\nEXPOSURE=1.0\n\n-Y 2 +X 16\n" + bytes(scan) * 2
    img, ww, hh = P.read_hdr(data)
    assert (ww, hh) == (16, 2)
    for ch in range(4):
        assert list(img[1, :, ch]) == [7 + ch] * 10 + list(range(ch, ch + 6))
