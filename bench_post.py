#!/usr/bin/env python3
"""bench_post.py — `python bench.py --stage post`: the post chain (bloom extract -> 10 x 9-tap blur
-> combine -> TAA resolve, src/ForwardShadingPipeline.cpp:185-254) on one ray-traced frame.

Same JSON contract as bench.py. A step = one rt_post_chain call over one 1920x1080 frame (C2's
frame, rendered once by the ray tracer, untimed). ``value`` = Mpixels/s with the three input images
resident in HBM (CUDA events, L2 flushed between iterations); ``e2e`` = the same through host
buffers (pinned H2D of colour/normal/history + chain + D2H of display/history inside the timed
region). Roofline: HBM. Algorithmic bytes per pixel = 72: read colour 16 + normal 8 + history 16,
write display 16 + history 16 (the bloom ping-pong images are internal: a fused chain never
writes them). N > 1: every rank processes its own frame (independent frames, no collective):
"weak". ``--impl reference`` times the CPU oracle (the reference's shaders compiled as C++ when
oracle/_ref/libref_post.so is present, else the port) with all host threads.
"""
from __future__ import annotations

import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mpixels/s"
BYTES_PER_PIXEL = 72


def _cpu(kind: str, color, normal, history, jx, jy, repeats: int = 1):
    import oracle
    o = oracle.Post(kind)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        o.post_chain(color, normal, history, blend=0.1, jx=jx, jy=jy)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return best


def _oracle_kind():
    import oracle
    oracle.build_post_port()
    return "reference" if oracle.have_post("reference") else "port"


def run_reference(args, emit) -> None:
    import numpy as np
    if int(os.environ.get("RANK", "0")) != 0:
        return
    kind = _oracle_kind()
    W, H = args.width or 1920, args.height or 1080
    rng = np.random.default_rng(5)
    color = (rng.random((H, W, 4), dtype=np.float32) ** 3 * 4).astype(np.float32)
    normal = rng.standard_normal((H, W, 4)).astype(np.float16).astype(np.float32)
    history = (rng.random((H, W, 4), dtype=np.float32) * 4).astype(np.float32)
    from opengl_raytracing_b200 import post as P
    jx, jy = P.taa_jitter(0, W, H)
    for _ in range(max(1, min(args.warmup, 1))):
        _cpu(kind, color, normal, history, jx, jy)
    steps = max(1, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(steps):
        _cpu(kind, color, normal, history, jx, jy)
    dt = (time.perf_counter() - t0) / steps
    v = W * H / dt / 1e6
    cores = os.cpu_count()
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": METRIC, "n_gpus": args.gpus, "steps": steps,
            "warmup": 1, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"post chain on a {W}x{H} frame (bloom threshold 1.0, 10 blur passes, strength 0.5, TAA blend 0.1)",
                       "width": W, "height": H},
            "cpu_baseline": {"value": v, "unit": METRIC, "cores": cores, "kind": kind,
                             "sample": f"full {W}x{H} frame, OpenMP static over rows, {steps} steps"},
            "e2e": {"value": v, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def run_ours(args, emit) -> None:
    import numpy as np
    import torch
    import torch.distributed as dist
    from bench import ClockSampler
    from opengl_raytracing_b200 import post as P, synthetic
    from opengl_raytracing_b200.pipeline import RtContext, render_workload

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    ctx = RtContext(local)
    w = synthetic.config(2, args.width, args.height)
    w.frame_count = rank                                  # every rank post-processes its own frame
    k = render_workload(ctx, w)
    H, W = w.height, w.width
    px = H * W
    color = torch.from_numpy(k["color"]).to(dev)
    normal = torch.from_numpy(k["normal"]).to(dev).to(torch.float16)
    hist_np = np.roll(k["color"], 3, axis=1).copy()       # a plausible previous frame
    history = torch.from_numpy(hist_np).to(dev)
    display = torch.empty_like(color)
    hist_out = torch.empty_like(color)
    pc = P.PostChain(ctx)
    jx, jy = P.taa_jitter(w.frame_count, W, H)
    p = P.post_params(W, H, blend=0.1, jitter=(jx, jy), flags=args.flags)
    stream = torch.cuda.current_stream(dev)
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)

    def step():
        pc.post_chain(p, color, normal, history, display=display, history_out=hist_out)

    for _ in range(max(3, args.warmup)):
        step()
    torch.cuda.synchronize(dev)
    launches0 = pc.launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    sampler = ClockSampler(local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    sampler.start()
    for i in range(args.steps):
        flush.zero_()
        ev[i][0].record(stream)
        step()
        ev[i][1].record(stream)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    launches = pc.launch_count() - launches0
    ms = sum(a.elapsed_time(b) for a, b in ev) / args.steps
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())

    # ---- e2e: host buffers in, host buffers out
    h_color = torch.from_numpy(k["color"]).pin_memory()
    h_normal = torch.from_numpy(k["normal"]).to(torch.float16).pin_memory()
    h_hist = torch.from_numpy(hist_np).pin_memory()
    h_display = torch.empty(H, W, 4, dtype=torch.float32).pin_memory()
    h_hist_out = torch.empty(H, W, 4, dtype=torch.float32).pin_memory()

    def e2e_step():
        color.copy_(h_color, non_blocking=True)
        normal.copy_(h_normal, non_blocking=True)
        history.copy_(h_hist, non_blocking=True)
        step()
        h_display.copy_(display, non_blocking=True)
        h_hist_out.copy_(hist_out, non_blocking=True)
        torch.cuda.synchronize(dev)

    for _ in range(3):
        e2e_step()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item())

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 0) or 0)
        peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
        if not peak:
            peak, peak_src = 7700.0, "fallback: nominal HBM3e bandwidth (MEASURED_PEAKS.json absent)"
        achieved = BYTES_PER_PIXEL * px / (ms * 1e-3) / 1e9
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(f"post@{W}x{H}", {}).get("bytes")
        except Exception:
            pass
        line = {"metric": METRIC, "value": world * px / (ms * 1e-3) / 1e6, "unit": METRIC, "n_gpus": world, "steps": args.steps,
                "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"post chain on a {W}x{H} frame (bloom threshold 1.0, 10 blur passes, strength 0.5, TAA blend 0.1)",
                           "width": W, "height": H, "parallelism": "single GPU" if world == 1 else f"independent frames x{world}",
                           "l2": "flushed between timed iterations (256 MB fill, untimed)",
                           "chain": "literal passes" if (args.flags & 1) else "fused"},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic, "peak_source": peak_src,
                             "algorithmic_bytes_per_pixel": BYTES_PER_PIXEL, "launches_per_step": launches / args.steps},
                "e2e": {"value": world * px / (e2e_ms * 1e-3) / 1e6, "unit": METRIC, "ms_per_step": e2e_ms,
                        "h2d_bytes_per_step": px * 40, "d2h_bytes_per_step": px * 32,
                        "call": "pinned H2D of colour/normal/history + rt_post_chain + D2H of display/history"},
                "gpu_launches": launches, "clocks": clocks}
        if not args.no_cpu_baseline:
            kind = _oracle_kind()
            jx0, jy0 = P.taa_jitter(0, W, H)
            dt = _cpu(kind, k["color"], k["normal"], hist_np, jx0, jy0, repeats=2)
            line["cpu_baseline"] = {"value": px / dt / 1e6, "unit": METRIC, "cores": os.cpu_count(), "kind": kind,
                                    "sample": f"full {W}x{H} frame, best of 2, OpenMP static over rows"}
        emit(line)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
