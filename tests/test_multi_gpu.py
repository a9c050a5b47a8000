"""include/rt_multi.h through the C ABI: one process driving several contexts (rt_multi: row bands stored straight
into the root's images; sample ranges summed by the peer pull-reduce kernel) and the one-process-per-GPU reduce
(rt_peer_reduce over CUDA-IPC mapped frames, flag words in peer memory, no NCCL). On a one-GPU box the devices are
listed twice / both ranks share cuda:0 — the protocol is the same, the transfers stay on one GPU."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from opengl_raytracing_b200 import abi, build as B, scene as S, synthetic  # noqa: E402
from opengl_raytracing_b200.pipeline import MultiGpu, RtContext, render_workload, workload_params  # noqa: E402

pytestmark = pytest.mark.gpu


def _devices(n):
    import torch
    have = torch.cuda.device_count()
    return [i % have for i in range(n)]


def _multi_render(m, w, sharding, flags=0, **over):
    import torch
    H, W = w.height, w.width
    dev = f"cuda:{m.devices[0]}"
    color = torch.full((H, W, 4), -5.0, device=dev); pos = torch.zeros(H, W, 4, device=dev)
    ndt = torch.float32 if flags & abi.FLAG_NORMAL_FP32 else torch.float16
    nrm = torch.zeros(H, W, 4, dtype=ndt, device=dev); oid = torch.full((H, W), -9, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    m.render(workload_params(w, flags=flags, **over), sharding, color, pos, nrm, oid)
    m.synchronize()
    return {"color": color.cpu().numpy(), "position": pos.cpu().numpy(), "normal": nrm.float().cpu().numpy(), "object_id": oid.cpu().numpy()}


@pytest.mark.parametrize("n,cfg,size", [(2, 2, (640, 360)), (3, 3, (480, 270)), (4, 5, (512, 300))])
def test_rt_multi_rows_is_bit_identical_to_one_gpu(rt_ctx, n, cfg, size):
    w = synthetic.config(cfg, *size)
    ref = render_workload(rt_ctx, w, flags=abi.FLAG_NORMAL_FP32)
    m = MultiGpu(_devices(n))
    m.bind_workload(w)
    for _ in range(2):                                     # twice: the second frame reuses every buffer and event
        got = _multi_render(m, w, abi.MULTI_ROWS, flags=abi.FLAG_NORMAL_FP32)
        for k in ("color", "position", "normal"):
            assert np.array_equal(got[k].view(np.uint32), ref[k].view(np.uint32)), (n, cfg, k)
        assert np.array_equal(got["object_id"], ref["object_id"])
    assert m.launch_count() >= 2 * n
    m.close()


@pytest.mark.parametrize("n,samples", [(2, 8), (3, 7), (4, 2)])
def test_rt_multi_samples_matches_the_single_gpu_mean(rt_ctx, n, samples):
    """Sample ranges: device g renders its range into a partial sum, the reduce kernels add the partials in rank order.
    Same samples, different association of the FP32 sum -> within summation error of the one-GPU mean; geometry from sample 0
    is bit-identical. 4 devices x 2 samples: two devices have no sample at all (their partial is zero)."""
    w = synthetic.config(4, 384, 216)
    w.num_samples = samples
    ref = render_workload(rt_ctx, w, flags=abi.FLAG_NORMAL_FP32)
    m = MultiGpu(_devices(n))
    m.bind_workload(w)
    for _ in range(2):
        got = _multi_render(m, w, abi.MULTI_SAMPLES, flags=abi.FLAG_NORMAL_FP32)
        assert np.array_equal(got["position"].view(np.uint32), ref["position"].view(np.uint32))
        assert np.array_equal(got["object_id"], ref["object_id"])
        assert (got["color"][..., 3] == 1.0).all()
        ok = np.isfinite(ref["color"]).all(axis=-1)
        scale = max(1.0, float(np.abs(ref["color"][ok]).max()))
        assert np.abs(got["color"][ok] - ref["color"][ok]).max() <= 2e-6 * samples * scale
    # deterministic: the same frame again gives the same bits
    again = _multi_render(m, w, abi.MULTI_SAMPLES, flags=abi.FLAG_NORMAL_FP32)
    assert np.array_equal(again["color"].view(np.uint32), got["color"].view(np.uint32))
    m.close()


def test_rt_multi_render_host_and_errors(rt_ctx):
    w = synthetic.config(2, 200, 120)
    ref = render_workload(rt_ctx, w)
    m = MultiGpu(_devices(2))
    H, W = w.height, w.width
    color = np.zeros((H, W, 4), np.float32); pos = np.zeros((H, W, 4), np.float32)
    nrm = np.zeros((H, W, 4), np.float16); oid = np.zeros((H, W), np.int32)
    m.bind_workload(w)                                     # textures; the scene comes with the call below
    m.render_host(workload_params(w), abi.MULTI_ROWS, w.objects, w.lights, color, pos, nrm, oid)
    assert np.array_equal(color.view(np.uint32), ref["color"].view(np.uint32))
    assert np.array_equal(oid, ref["object_id"]) and np.array_equal(nrm.astype(np.float32), ref["normal"])
    with pytest.raises(abi.RtError) as ei:
        m.render(workload_params(w, row_begin=4), abi.MULTI_ROWS, None, None, None, None)
    assert ei.value.status == abi.RT_ERR_INVALID_ARG
    with pytest.raises(abi.RtError):
        m.render(workload_params(w, flags=abi.FLAG_ACCUMULATE), abi.MULTI_SAMPLES, None, None, None, None)
    m.close()
    with pytest.raises(abi.RtError):
        MultiGpu([99])


def test_headless_cpp_host_drives_several_gpus(tmp_path):
    """The reference-shaped C++ host (host/rt_headless, ForwardShadingPipeline::Render) over rt_multi: the frame of two
    'GPUs' == the frame of one, bit for bit (row bands), and within summation error (sample ranges)."""
    import torch
    headless = B.build_host()
    w = synthetic.config(2, 320, 200)
    text = S.save_scene(S.Scene(objects=w.objects, lights=w.lights, object_names=[], light_names=[]))
    p = tmp_path / "c2.scene"
    p.write_text(text)
    gpus = ",".join(str(d) for d in _devices(2))
    outs = {}
    for name, extra in (("one", []), ("rows", ["--gpus", gpus, "--shard", "rows"]), ("samples", ["--gpus", gpus, "--shard", "samples", "--samples", "4"]),
                        ("one4", ["--samples", "4"])):
        out = str(tmp_path / name)
        r = subprocess.run([headless, "--scene", str(p), "--size", "320", "200", "--frames", "2", "--raw", "--out", out] + extra,
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        info = json.loads(r.stdout.strip().splitlines()[-1])
        assert info["launches"] >= 2
        outs[name] = {k: np.fromfile(f"{out}_{k}.{ext}", dtype=dt) for k, ext, dt in
                      (("color", "f32", np.float32), ("position", "f32", np.float32), ("normal", "f16", np.float16))}
    for k in ("color", "position", "normal"):
        assert outs["one"][k].tobytes() == outs["rows"][k].tobytes(), k
    assert outs["one4"]["position"].tobytes() == outs["samples"]["position"].tobytes()
    a, b = outs["one4"]["color"], outs["samples"]["color"]
    ok = np.isfinite(a)
    assert np.abs(a[ok] - b[ok]).max() <= 1e-5 * max(1.0, float(np.abs(a[ok]).max()))
    assert torch.cuda.is_available()


def test_peer_reduce_two_processes():
    """rt_peer_reduce between two rank processes (gloo for the handle exchange only): every rank renders its sample
    range into its IPC-exported partial, the reduce kernels synchronise through flag words in peer memory and leave
    the mean in rank 0's frame; three consecutive frames (the partial buffers alternate)."""
    import torch
    here = os.path.dirname(os.path.abspath(__file__))
    port = 29300 + os.getpid() % 300
    env = dict(os.environ, PEER_GPUS=str(min(2, torch.cuda.device_count())))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(here, "peer_reduce_worker.py")],
                       capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0 and "PEER_REDUCE_OK" in r.stdout, (r.stdout[-2000:], r.stderr[-3000:])
