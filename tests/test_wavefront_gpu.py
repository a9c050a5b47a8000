"""The two ray-regrouping variants — the wavefront pipeline (csrc/rt_wavefront.cuh: bounce / shadow-job launches
over ballot-compacted ray queues in HBM) and the regrouping megakernel (csrc/rt_regroup.cuh: the same rounds
inside a persistent CTA, queues in shared memory) — against the plain persistent megakernel, through the C ABI
(rt_set_tuning pins the variant). All run the same per-pixel arithmetic — only which lane runs it differs — so
every image must agree BIT FOR BIT, in both
math builds: the five BASELINE configs, multi-sample sums, row windows and interleaved bands, ragged sizes,
the edge-case scenes, several chunkings of the frame, and the explicit tuning entry points."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from opengl_raytracing_b200 import abi, synthetic  # noqa: E402
from opengl_raytracing_b200.pipeline import RtContext, render_workload, workload_params, bind_workload  # noqa: E402
from edge_cases import edge_workloads  # noqa: E402

pytestmark = pytest.mark.gpu

ZERO = {"color": 0, "position": 0, "normal": 0, "object_id": 0}


def _bits_equal(a: dict, b: dict) -> dict:
    d = {k: int((a[k].view(np.uint32) != b[k].view(np.uint32)).sum()) if a[k].dtype == np.float32
         else int((a[k] != b[k]).sum()) for k in ("color", "position", "normal", "object_id")}
    if hasattr(b, "other"):                    # the regrouping megakernel's image rides along with the wavefront's
        for k, v in _bits_equal(a, b.other).items():
            d[k] += v
    return d


class _Many(dict):
    pass


def _both(ctx, w, flags, params=None, bind=True):
    """Renders with the megakernel, the wavefront pipeline and the regrouping megakernel; returns (megakernel image,
    images of the other two merged for _bits_equal: a differing element in EITHER counts)."""
    ctx.set_tuning(abi.KERNEL_MEGAKERNEL)
    a = render_workload(ctx, w, flags=flags, params=params, bind=bind)
    assert ctx.get_tuning()["kernel"] == "megakernel"
    others = []
    for kern, name in ((abi.KERNEL_WAVEFRONT, "wavefront"), (abi.KERNEL_REGROUP, "regroup")):
        ctx.set_tuning(kern)
        others.append(render_workload(ctx, w, flags=flags, params=params, bind=False))
        assert ctx.get_tuning()["kernel"] == name
    ctx.set_tuning(None)
    b = _Many(others[0])
    b.other = others[1]
    return a, b


SIZES = {1: (1280, 720), 2: (1920, 1080), 3: (960, 540), 4: (480, 270), 5: (1920, 1080)}


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5])
@pytest.mark.parametrize("math", ["fast", "parity"])
def test_wavefront_is_bit_identical_to_megakernel(rt_ctx, n, math):
    w = synthetic.config(n, *SIZES[n])
    if n == 4:
        w.num_samples = 5
    fl = abi.FLAG_NORMAL_FP32 | (abi.FLAG_PARITY_MATH if math == "parity" else 0)
    a, b = _both(rt_ctx, w, fl)
    assert _bits_equal(a, b) == ZERO, (n, math)


def test_wavefront_half_normals_and_null_outputs(rt_ctx):
    import torch
    w = synthetic.config(2, 640, 360)
    a, b = _both(rt_ctx, w, 0)
    assert _bits_equal(a, b) == ZERO
    # colour only (the sample-range partial-sum path passes NULL for the other images)
    bind_workload(rt_ctx, w)
    rt_ctx.set_tuning(abi.KERNEL_WAVEFRONT)
    color = torch.zeros(w.height, w.width, 4, device="cuda")
    rt_ctx.render(workload_params(w), color, None, None, None, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    rt_ctx.set_tuning(None)
    assert np.array_equal(color.cpu().numpy().view(np.uint32), a["color"].view(np.uint32))


@pytest.mark.parametrize("name", sorted(edge_workloads().keys()))
def test_wavefront_on_edge_cases(rt_ctx, name):
    w = edge_workloads(size=(96, 64))[name]
    for fl in (abi.FLAG_NORMAL_FP32, abi.FLAG_NORMAL_FP32 | abi.FLAG_PARITY_MATH):
        a, b = _both(rt_ctx, w, fl)
        assert _bits_equal(a, b) == ZERO, (name, fl)


@pytest.mark.parametrize("size", [(1, 1), (7, 3), (33, 5), (9, 37)])
def test_wavefront_ragged_sizes(rt_ctx, size):
    for n in (2, 3):
        w = synthetic.config(n, *size)
        a, b = _both(rt_ctx, w, abi.FLAG_NORMAL_FP32)
        assert _bits_equal(a, b) == ZERO, (n, size)


def test_wavefront_depth_zero_and_deep(rt_ctx):
    for depth in (0, 1, 7):
        w = synthetic.config(3, 320, 180)
        w.max_ray_depth = depth
        a, b = _both(rt_ctx, w, abi.FLAG_NORMAL_FP32)
        assert _bits_equal(a, b) == ZERO, depth


@pytest.mark.parametrize("chunk", ["256", "5000", "40000"])
def test_wavefront_chunking_does_not_change_the_image(monkeypatch, chunk):
    """The frame is processed in chunks of tile rows (RT_WF_CHUNK pixels each) that reuse the queue buffers;
    a chunk of one tile row, of a few rows and of the whole frame must give the same bits."""
    w = synthetic.config(3, 400, 220)
    w.num_samples = 3
    monkeypatch.setenv("RT_WF_CHUNK", chunk)
    c = RtContext(0)
    a, b = _both(c, w, abi.FLAG_NORMAL_FP32)
    c.close()
    assert _bits_equal(a, b) == ZERO


def test_wavefront_row_windows_bands_and_sample_modes(rt_ctx):
    import torch
    w = synthetic.config(3, 320, 200)
    w.num_samples = 3
    bind_workload(rt_ctx, w)
    H, W = w.height, w.width
    st = torch.cuda.current_stream().cuda_stream
    outs = {}
    for kern in (abi.KERNEL_MEGAKERNEL, abi.KERNEL_WAVEFRONT, abi.KERNEL_REGROUP):
        rt_ctx.set_tuning(kern)
        color = torch.full((H, W, 4), 0.25, device="cuda"); pos = torch.zeros(H, W, 4, device="cuda")
        nrm = torch.zeros(H, W, 4, dtype=torch.float16, device="cuda"); oid = torch.full((H, W), -9, dtype=torch.int32, device="cuda")
        # a row window, then interleaved bands of three "ranks", then partial sums on top (ACCUMULATE / SUM_SAMPLES)
        rt_ctx.render(workload_params(w, row_begin=0, row_end=52), color, pos, nrm, oid, st)
        for g in range(3):
            rt_ctx.render(workload_params(w, row_begin=52, row_band=8, row_period=3, row_phase=g), color, pos, nrm, oid, st)
        acc = torch.zeros(H, W, 4, device="cuda")
        rt_ctx.render(workload_params(w, flags=abi.FLAG_ACCUMULATE, frame_count=3), acc, None, None, None, st)
        rt_ctx.render(workload_params(w, flags=abi.FLAG_ACCUMULATE, frame_count=6), acc, None, None, None, st)
        sums = torch.zeros(H, W, 4, device="cuda")
        rt_ctx.render(workload_params(w, flags=abi.FLAG_SUM_SAMPLES), sums, None, None, None, st)
        torch.cuda.synchronize()
        outs[kern] = [t.cpu().numpy() for t in (color, pos, nrm.float(), oid, acc, sums)]
    rt_ctx.set_tuning(None)
    for other in (abi.KERNEL_WAVEFRONT, abi.KERNEL_REGROUP):
        for x, y in zip(outs[abi.KERNEL_MEGAKERNEL], outs[other]):
            assert np.array_equal(x, y, equal_nan=True), other
    assert (outs[abi.KERNEL_WAVEFRONT][3] >= -1).all()          # every pixel of the frame was written exactly by its window


def test_shipped_scene_eight_lights(rt_ctx):
    from edge_cases import shipped_scene
    w = shipped_scene("performance_test", (960, 540), 3)
    w.use_skybox, w.skybox, w.noise = True, synthetic.skybox_s1(), synthetic.noise_n1()
    a, b = _both(rt_ctx, w, abi.FLAG_NORMAL_FP32)
    assert _bits_equal(a, b) == ZERO


def test_overlapping_launches_on_two_streams(rt_ctx):
    """rt_render launches of one context may overlap on different streams: each takes its own work counter
    (megakernel) / is ordered behind the previous user of the queue buffers (wavefront)."""
    import torch
    w = synthetic.config(3, 512, 288)
    bind_workload(rt_ctx, w)
    ref = render_workload(rt_ctx, w, bind=False)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for kern in (abi.KERNEL_MEGAKERNEL, abi.KERNEL_WAVEFRONT, abi.KERNEL_REGROUP):
        rt_ctx.set_tuning(kern)
        bufs = [torch.zeros(w.height, w.width, 4, device="cuda") for _ in range(6)]
        torch.cuda.synchronize()
        for i, b in enumerate(bufs):
            rt_ctx.render(workload_params(w), b, None, None, None, (s1 if i % 2 == 0 else s2).cuda_stream)
        torch.cuda.synchronize()
        for b in bufs:
            assert np.array_equal(b.cpu().numpy().view(np.uint32), ref["color"].view(np.uint32)), kern
    rt_ctx.set_tuning(None)


def test_rt_tune_picks_a_variant_and_never_changes_the_image():
    c = RtContext(0)
    w = synthetic.config(3, 640, 360)
    bind_workload(c, w)
    p = workload_params(w, flags=abi.FLAG_NORMAL_FP32)
    before = c.get_tuning(p)
    assert before["state"] == "default" and len(before["variants"]) >= 2
    c.set_autotune(False)
    base = render_workload(c, w, flags=abi.FLAG_NORMAL_FP32, bind=False)
    assert c.get_tuning(p)["state"] == "default"                 # autotune off: no trial frames
    t = c.tune(p)
    assert t["state"] == "chosen" and all(v["ms"] > 0 for v in t["variants"])
    best = min(t["variants"], key=lambda v: v["ms"])
    assert (t["kernel"], t["registers"], t["phase_level"]) == (best["kernel"], best["registers"], best["phase_level"])
    after = render_workload(c, w, flags=abi.FLAG_NORMAL_FP32, bind=False)
    ran = c.get_tuning()
    assert (ran["kernel"], ran["registers"], ran["phase_level"]) == (t["kernel"], t["registers"], t["phase_level"])
    assert _bits_equal(base, after) == ZERO
    # autotune on a fresh context: trials run on the caller's launches, then the choice sticks
    c.close()
    c = RtContext(0)
    frames = [render_workload(c, w, flags=abi.FLAG_NORMAL_FP32, bind=(i == 0)) for i in range(2 + 2 * len(t["variants"]))]
    assert c.get_tuning(p)["state"] == "chosen"
    for f in frames:
        assert _bits_equal(f, base) == ZERO
    c.close()


def test_sm_reserve_shrinks_the_grid_not_the_image(rt_ctx):
    """rt_set_sm_reserve: render grids sized for fewer SMs (room for a kernel of another stream); same bits, bad values rejected."""
    w = synthetic.config(3, 480, 270)
    base = render_workload(rt_ctx, w, flags=abi.FLAG_NORMAL_FP32)
    try:
        for sms in (8, 100, 147):
            rt_ctx.set_sm_reserve(sms)
            a, b = _both(rt_ctx, w, abi.FLAG_NORMAL_FP32, bind=False)
            assert _bits_equal(base, a) == ZERO and _bits_equal(base, b) == ZERO, sms
        for bad in (-1, 148, 10 ** 6):
            with pytest.raises(Exception):
                rt_ctx.set_sm_reserve(bad)
    finally:
        rt_ctx.set_sm_reserve(0)
