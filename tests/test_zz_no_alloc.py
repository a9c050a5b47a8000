"""rt_render_static (include/rt_abi.h): the entry point that allocates nothing — the wavefront pipeline with its ray queues
in a file-scope `static __device__` array — against rt_render, through the C ABI. Same per-pixel arithmetic, so the images
must be the existing entry point's BIT FOR BIT (stricter than the parity tolerances). Two calls in a row on different scenes:
the array is shared and keeps the previous call's contents. Collected last on purpose."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from opengl_raytracing_b200 import abi, synthetic  # noqa: E402
from opengl_raytracing_b200.pipeline import bind_workload, render_workload, workload_params  # noqa: E402

pytestmark = pytest.mark.gpu


def _static(ctx, w, flags):
    import torch
    H, W = w.height, w.width
    color = torch.full((H, W, 4), -5.0, device="cuda"); pos = torch.full((H, W, 4), -5.0, device="cuda")
    nrm = torch.full((H, W, 4), -5.0, device="cuda"); oid = torch.full((H, W), -9, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    ctx.render_static(workload_params(w, flags=flags), color, pos, nrm, oid, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return {"color": color.cpu().numpy(), "position": pos.cpu().numpy(), "normal": nrm.cpu().numpy(), "object_id": oid.cpu().numpy()}, free0 - free1


def _same(a, b):
    return all(np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32)) for k in ("color", "position", "normal")) and \
        np.array_equal(a["object_id"], b["object_id"])


def test_render_static_matches_rt_render_on_two_calls_in_a_row(rt_ctx):
    fl = abi.FLAG_NORMAL_FP32
    scenes = [synthetic.config(3, 640, 360), synthetic.config(2, 1920, 1080)]      # 260 objects / 4 lights, then 4 objects / 3 lights
    refs = []
    for w in scenes:
        refs.append(render_workload(rt_ctx, w, flags=fl))                          # the existing entry point (whatever variant it picks)
    # a first call loads the kernels' module (lazy loading may take device memory once); the checked calls follow
    bind_workload(rt_ctx, scenes[0])
    _static(rt_ctx, scenes[0], fl)
    used = []
    for w, ref in list(zip(scenes, refs)) + [(scenes[1], refs[1])]:                # consecutive calls: C3, C2, C2 again
        bind_workload(rt_ctx, w)
        got, delta = _static(rt_ctx, w, fl)
        assert _same(got, ref), w.name
        used.append(delta)
    assert rt_ctx.get_tuning()["kernel"] == "wavefront"
    assert all(d == 0 for d in used[1:]), used                                     # no device memory was taken by the calls
    # parity-math build too
    w = scenes[0]
    bind_workload(rt_ctx, w)
    ref = render_workload(rt_ctx, w, flags=fl | abi.FLAG_PARITY_MATH, bind=False)
    got, _ = _static(rt_ctx, w, fl | abi.FLAG_PARITY_MATH)
    assert _same(got, ref)


def test_render_static_chunks_a_frame_larger_than_its_scratch(rt_ctx):
    """C5 at 2560x1440 with 2 lights needs 3.7 M queue slots x 232 B = 855 MB of queues in one piece: the call walks it in
    chunks of tile rows that fit the 64 MiB array."""
    w = synthetic.config(5, 2560, 1440)
    ref = render_workload(rt_ctx, w, flags=abi.FLAG_NORMAL_FP32)
    got, _ = _static(rt_ctx, w, abi.FLAG_NORMAL_FP32)
    assert _same(got, ref)


def test_render_static_rejects_what_it_cannot_hold(rt_ctx):
    w = synthetic.config(3, 64, 36)
    bind_workload(rt_ctx, w)
    with pytest.raises(abi.RtError) as ei:                                         # multi-sample frames need a full-frame accumulator
        rt_ctx.render_static(workload_params(w, num_samples=2), None, None, None, None)
    assert ei.value.status == abi.RT_ERR_INVALID_ARG and "num_samples" in str(ei.value)
    with pytest.raises(abi.RtError) as ei:                                         # one tile row of queues beyond 64 MiB (4 lights: 304 B per slot)
        rt_ctx.render_static(workload_params(w, width=60000, height=8), None, None, None, None)
    assert ei.value.status == abi.RT_ERR_INVALID_ARG and "static scratch" in str(ei.value)
    with pytest.raises(abi.RtError):
        rt_ctx.render_static(workload_params(w, flags=abi.FLAG_COUNTERS), None, None, None, None)
