#!/usr/bin/env python3
"""bench_frame.py — `python bench.py --stage frame`: one whole iteration of
ForwardShadingPipline::Render (src/ForwardShadingPipeline.cpp:139-272) per step — the ray-tracing
pass, bloom (extract, 10 blurs, combine) and the TAA resolve — through include/rt_frame.h, with the
three images the pass writes staying in HBM as they do in the reference (GL textures there).

Same JSON contract as bench.py. Workload: the same frame as the default stage (C2 unless
--config / --scene says otherwise), TAA on, frame_count advancing every step (so the noise lookup,
the jitter and the history ping-pong move as they do in the app). ``value`` = reference-literal
rays of the frame / CUDA-event time of rt_frame_render (ray tracing + post chain), L2 flushed
between iterations. ``e2e`` = rt_frame_render_host: scene H2D + the frame + presentation, i.e. the
display image converted to the 8-bit UNORM the reference's window holds and copied into pinned
host memory (4 B per pixel instead of the 40 B per pixel of intermediates the default stage's e2e
moves). ``roofline`` describes the dominant kernel, the ray-tracing kernel, timed alone on the same
frame images. N > 1: independent frames per rank (a frame is not row-sharded: the post chain
reads the whole image), "weak". ``--impl reference`` = the CPU oracles of both stages back to back.
"""
from __future__ import annotations

import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mrays/s"


def _cpu_frame(w, args, budget_s: float) -> dict:
    """CPU seconds of one frame: the ray-tracing oracle on a bounded pixel subset (scaled to the
    frame) + the post-chain oracle on the whole frame."""
    import numpy as np
    import bench
    import bench_post
    import oracle
    from opengl_raytracing_b200 import post as P
    kind = "reference" if oracle.have("reference") else "port"
    cb = bench.cpu_baseline(w, kind, budget_s=budget_s, step=args.cpu_sample_step)
    step = cb["step"]
    t_rt = cb["seconds"] * step * step
    rays = cb["rays"] * step * step
    pkind = bench_post._oracle_kind()
    rng = np.random.default_rng(5)
    H, W = w.height, w.width
    color = (rng.random((H, W, 4), dtype=np.float32) ** 3 * 4).astype(np.float32)
    normal = rng.standard_normal((H, W, 4)).astype(np.float16).astype(np.float32)
    history = (rng.random((H, W, 4), dtype=np.float32) * 4).astype(np.float32)
    jx, jy = P.taa_jitter(0, W, H)
    t_post = bench_post._cpu(pkind, color, normal, history, jx, jy, repeats=1)
    return {"seconds": t_rt + t_post, "rays": rays, "cores": cb["cores"], "kind": kind if kind == pkind else "port",
            "sample": f"ray tracing: {cb['sample']} (x{step * step} to the frame: {t_rt:.3f} s); post chain: full {W}x{H} frame, {t_post:.3f} s"}


def run_reference(args, emit) -> None:
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import bench
    w = bench.workload_for(args)
    steps = max(1, min(args.steps, 3))
    best = None
    for _ in range(steps):
        c = _cpu_frame(w, args, budget_s=6.0)
        best = c if best is None or c["seconds"] < best["seconds"] else best
    v = best["rays"] / best["seconds"] / 1e6
    emit({"impl": "reference", "metric": METRIC, "value": v, "unit": METRIC, "n_gpus": args.gpus, "steps": steps, "warmup": 0,
          "ms_per_step": best["seconds"] * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
          "dtype": "f32", "data": "synthetic",
          "config": {"workload": f"whole frame (ray tracing + bloom + TAA) of {w.name}: {w.description}", "width": w.width, "height": w.height},
          "cpu_baseline": {"value": v, "unit": METRIC, "cores": best["cores"], "kind": best["kind"], "sample": best["sample"]},
          "e2e": {"value": v, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0})


def run_ours(args, emit) -> None:
    import numpy as np
    import torch
    import torch.distributed as dist
    import bench
    from opengl_raytracing_b200 import abi, post as P
    from opengl_raytracing_b200.frame import Frame
    from opengl_raytracing_b200.pipeline import RtContext, bind_workload, workload_params

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
    w = bench.workload_for(args)
    ctx = RtContext(local)
    bind_workload(ctx, w)
    H, W = w.height, w.width
    px = H * W
    fr = Frame(ctx, W, H)
    stream = torch.cuda.current_stream(dev)
    base_fc = w.frame_count + rank * 1000               # every rank renders its own frames

    def params(i):
        fc = base_fc + i
        return (workload_params(w, flags=args.flags, frame_count=fc),
                P.post_params(W, H, blend=0.1, jitter=P.taa_jitter(fc, W, H)))

    # reference-literal op counts of one frame (counter variant of the kernel, untimed)
    rp0, pp0 = params(0)
    pc = abi.FrameParams.from_buffer_copy(rp0)
    pc.flags = rp0.flags | abi.FLAG_COUNTERS
    scratch = torch.zeros(H, W, 4, dtype=torch.float32, device=dev)
    ctx.counters(reset=True)
    ctx.render(pc, scratch, None, None, None, stream.cuda_stream)
    counters = ctx.counters(reset=True)
    del scratch
    rays = counters["rays"]
    flops, sfu = bench.algorithmic_flops(counters)

    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
    nwarm = max(6, args.warmup)                          # lets the per-scene variant choice for large scenes settle (rt_abi.cu)
    plist = [params(i) for i in range(nwarm + args.steps)]
    for i in range(nwarm):
        fr.render(*plist[i], stream.cuda_stream)
    torch.cuda.synchronize(dev)
    fp32_peak = ctx.measure_fp32_peak()
    l0 = (ctx.launch_count(), P.PostChain(ctx).launch_count())
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    sampler = bench.ClockSampler(local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    sampler.start()
    for i in range(args.steps):
        flush.zero_()                                    # L2 flush between timed iterations (untimed)
        ev[i][0].record(stream)
        fr.render(*plist[nwarm + i], stream.cuda_stream)
        ev[i][1].record(stream)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    launches = (ctx.launch_count() - l0[0]) + (P.PostChain(ctx).launch_count() - l0[1])
    ms = sum(a.elapsed_time(b) for a, b in ev) / args.steps

    # the dominant kernel alone (the ray-tracing kernel into the frame's own images), for the roofline
    img = fr.images()
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for i in range(args.steps):
        flush.zero_()
        kev[i][0].record(stream)
        ctx.render(plist[nwarm + i][0], img["color"], img["position"], img["normal"], None, stream.cuda_stream)
        kev[i][1].record(stream)
    torch.cuda.synchronize(dev)
    kern_ms = sum(a.elapsed_time(b) for a, b in kev) / args.steps

    # ---- e2e: host scene in, 8-bit display image out
    objs = np.ascontiguousarray(w.objects); lights = np.ascontiguousarray(w.lights)
    h_display = torch.empty(H, W, 4, dtype=torch.uint8).pin_memory()
    e2e_steps = max(5, min(args.steps, 20))
    for i in range(3):
        fr.render_host(*plist[i], objs, lights, h_display, abi.DISPLAY_RGBA8)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        fr.render_host(*plist[nwarm + i % args.steps], objs, lights, h_display, abi.DISPLAY_RGBA8)
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    t = torch.tensor([ms, e2e_ms, kern_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms, kern_ms = (float(x) for x in t)

    if rank == 0:
        ach_tf = flops / (kern_ms * 1e-3) / 1e12
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(f"{w.name}@{W}x{H}", {}).get("bytes")
        except Exception:
            pass
        line = {"metric": METRIC, "value": world * rays / (ms * 1e-3) / 1e6, "unit": METRIC, "n_gpus": world, "steps": args.steps,
                "warmup": nwarm, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"whole frame (ray tracing + bloom + TAA) of {w.name}: {w.description}", "width": W, "height": H,
                           "objects": len(w.objects), "lights": len(w.lights), "max_ray_depth": w.max_ray_depth,
                           "parallelism": "single GPU" if world == 1 else f"independent frames x{world}",
                           "l2": "flushed between timed iterations (256 MB fill, untimed)", "rays_per_frame": rays,
                           "post": "bloom threshold 1.0, 10 blur passes, strength 0.5, TAA blend 0.1, frame_count advancing"},
                "ms_per_frame": ms, "ms_ray_tracing_kernel": kern_ms,
                "roofline": {"bound": "fp32", "achieved": ach_tf, "peak": fp32_peak["tflops"], "unit": "TFLOP/s",
                             "frac": ach_tf / fp32_peak["tflops"], "traffic": traffic,
                             "kernel": "the ray-tracing kernel, timed alone on the frame's images (its share of the step: "
                                       f"{kern_ms / ms:.2f})",
                             "peak_source": "measured in this run: FFMA-only micro-kernel (rt_measure_fp32_peak)",
                             "algorithmic_flops_per_launch": flops, "algorithmic_sfu_ops_per_launch": sfu},
                "e2e": {"value": world * rays / (e2e_ms * 1e-3) / 1e6, "unit": METRIC, "ms_per_step": e2e_ms,
                        "h2d_bytes_per_step": objs.nbytes + lights.nbytes, "d2h_bytes_per_step": px * 4,
                        "call": "rt_frame_render_host: scene H2D + rt_render + rt_post_chain on device-resident images + "
                                "RGBA8 display image into pinned host memory"},
                "gpu_launches": launches, "clocks": clocks}
        if not args.no_cpu_baseline and world == 1:
            c = _cpu_frame(w, args, budget_s=12.0)
            line["cpu_baseline"] = {"value": c["rays"] / c["seconds"] / 1e6, "unit": METRIC, "cores": c["cores"], "kind": c["kind"],
                                    "sample": c["sample"]}
        emit(line)
    fr.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
