"""GPU parity of include/rt_frame.h (one whole iteration of ForwardShadingPipline::Render with the
images resident on the device) through the C ABI. Bar: the frame's images are those of rt_render,
and its display / history images those of the post-chain oracle fed with them, bit for bit; the
8-bit presentation equals the oracle's framebuffer conversion exactly."""
import numpy as np
import pytest

import oracle
from opengl_raytracing_b200 import abi, post as P, synthetic
from opengl_raytracing_b200.frame import Frame

pytestmark = pytest.mark.gpu

from test_post_oracle import same  # noqa: E402


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


def _host(t):
    return t.float().cpu().numpy()


@pytest.mark.parametrize("cfg,size", [(2, (320, 180)), (1, (97, 61)), (3, (128, 72))])
def test_frame_sequence_is_rt_render_then_post_chain_bit_for_bit(ctx, port, cfg, size):
    """Four consecutive frames with TAA on: frame_count advances, the history ping-pongs as
    :141/:236/:249 do, every image equals the separate calls' / the oracle's."""
    import torch
    from opengl_raytracing_b200.pipeline import bind_workload, render_workload, workload_params
    W, H = size
    w = synthetic.config(cfg, W, H)
    bind_workload(ctx, w)
    fr = Frame(ctx, W, H)
    hist_prev = np.zeros((H, W, 4), dtype=np.float32)           # rt_frame_create zero-fills the history images
    for fc in range(4):
        rp = workload_params(w, frame_count=fc)
        jx, jy = P.taa_jitter(fc, W, H)
        pp = P.post_params(W, H, jitter=(jx, jy))
        fr.render(rp, pp)
        torch.cuda.synchronize()
        img = {k: _host(v) for k, v in fr.images().items()}
        k = render_workload(ctx, w, params=rp, bind=False)       # the ray-tracing pass alone, same inputs
        assert img["color"].tobytes() == k["color"].tobytes()
        assert img["position"].tobytes() == k["position"].tobytes()
        assert img["normal"].tobytes() == np.ascontiguousarray(k["normal"], dtype=np.float32).tobytes()
        want_d, want_h = port.post_chain(k["color"], k["normal"], hist_prev, jx=jx, jy=jy)
        assert same(img["display"], want_d) and same(img["history"], want_h)
        d32 = fr.read_display(fmt=abi.DISPLAY_RGBA32F)
        assert same(d32, want_d)
        d8 = fr.read_display(fmt=abi.DISPLAY_RGBA8)
        assert d8.dtype == np.uint8 and d8.tobytes() == oracle.display_rgba8(want_d).tobytes()
        hist_prev = want_h
    assert fr.launch_count() == 4
    fr.close()


def test_frame_render_host_end_to_end_and_taa_off(ctx, port):
    import torch
    from opengl_raytracing_b200.pipeline import render_workload, workload_params
    W, H = 160, 96
    w = synthetic.config(2, W, H)
    ctx.set_skybox(w.skybox)
    ctx.set_noise(w.noise)
    fr = Frame(ctx, W, H)
    rp = workload_params(w)
    pp = P.post_params(W, H, taa=False, strength=0.75, threshold=0.5)
    pinned = torch.empty(H, W, 4, dtype=torch.uint8).pin_memory()
    fr.render_host(rp, pp, w.objects, w.lights, pinned, abi.DISPLAY_RGBA8)       # uploads the scene itself
    k = render_workload(ctx, w, params=rp, bind=False)
    want_d, none = port.post_chain(k["color"], k["normal"], None, threshold=0.5, strength=0.75, taa=False)
    assert none is None
    assert pinned.numpy().tobytes() == oracle.display_rgba8(want_d).tobytes()
    pageable = np.empty((H, W, 4), dtype=np.float32)
    fr.render_host(rp, pp, None, None, pageable, abi.DISPLAY_RGBA32F)            # keeps the uploaded scene
    assert same(pageable, want_d)
    fr.close()


def test_display_rgba8_edge_values(ctx):
    """NaN -> 0, negatives and values above 1 clamp, ties round to even."""
    import torch
    W, H = 8, 2
    fr = Frame(ctx, W, H)
    vals = np.array([np.nan, -1.0, 0.0, 0.5 / 255, 1.5 / 255, 2.5 / 255, 0.999, 1.0, 7.0, np.inf, -np.inf, 0.5,
                     1e-30, 254.5 / 255, 0.25, 0.75], dtype=np.float32)
    d = np.resize(vals, (H, W, 4)).astype(np.float32)
    fr.images()["display"].copy_(torch.from_numpy(d).cuda())
    got = fr.read_display(fmt=abi.DISPLAY_RGBA8)
    assert got.tobytes() == oracle.display_rgba8(d).tobytes()
    fr.close()


def test_frame_invalid_arguments_fail_loudly(ctx):
    from opengl_raytracing_b200.pipeline import bind_workload, workload_params
    w = synthetic.config(1, 64, 32)
    bind_workload(ctx, w)
    with pytest.raises(abi.RtError):
        Frame(ctx, 0, 32)
    fr = Frame(ctx, 64, 32)
    pp = P.post_params(64, 32)
    with pytest.raises(abi.RtError):                       # size mismatch
        fr.render(workload_params(synthetic.config(1, 32, 32)), pp)
    with pytest.raises(abi.RtError):                       # row-sharded frame
        fr.render(workload_params(w, row_end=16), pp)
    with pytest.raises(abi.RtError):                       # gNormalTex is RGBA16F
        fr.render(workload_params(w, flags=abi.FLAG_NORMAL_FP32), pp)
    with pytest.raises(abi.RtError):
        fr.read_display(fmt=7)
    fr.close()
