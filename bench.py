#!/usr/bin/env python3
"""bench.py — Mrays/s of the ray-tracing pass on BASELINE.json's metric config.

  python bench.py --gpus N --steps K --warmup W            (this repo's CUDA path)
  python bench.py --impl reference --gpus N --steps K ...  (the reference's CPU path)

A "step" is one pass of the hot path over one frame of synthetic input. At N=1 the workload
is BASELINE.json configs[1] (C2: default scene, 1920x1080, PBR + PCSS area light + skybox,
depth 3, 1 spp). ``value`` is whole-job Mrays/s with inputs resident in HBM (CUDA events on
the launching stream, max over ranks); ``e2e`` is the same metric through the reference-facing
C-ABI call rt_render_host with HOST buffers (scene H2D + render + D2H of the three images
inside the timed region). 1 ray = 1 intersectObjects call of the reference-literal
algorithm, counted by the kernel's counter variant (validated against the oracle in
tests/test_parity_gpu.py::test_counters_match_oracle).

N > 1 (torchrun, one rank per GPU): sample-range sharding — every rank renders one
full frame at its own frameCount, the partial colours are summed into rank 0's frame inside the
timed region: per-GPU work is fixed ("weak"). The sum is the library's peer pull-reduce kernel
(rt_peer_reduce over CUDA-IPC mapped partials), pipelined one frame deep: the reduce of frame
i-1 runs on a second stream, in CTA slots the render grid leaves free (rt_set_sm_reserve),
inside the timed region of frame i — K timed regions hold K renders and K reduces
(``--pipeline 0`` / ``--reduce nccl``: the exchange behind the render on one stream).
``--shard rows`` gives the strong-scaling row-band variant (no collective in the data path:
every kernel stores straight into rank 0's frame over NVLink).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mrays/s"

# flop model v0: reference-literal flops per counted operation, (flops, SFU ops)
FLOP_MODEL = {
    "rays": (3, 3), "aabb": (25, 0), "sphere": (32, 2), "plane_early": (17, 1), "plane_full": (80, 3),
    "hit_update_sphere": (18, 1), "hit_update_plane": (0, 0),
    "light_point": (57, 3), "light_dir": (45, 2), "light_area": (69, 4), "pbr": (101, 4),
    "pcf_setup": (32, 1), "pcf_sample": (56, 3), "pcss_probe": (36, 2), "sss_sample": (82, 7),
    "bounce_mirror": (66, 2), "bounce_refract": (129, 5), "bounce_diffuse": (152, 9),
    "rr": (25, 3), "camera": (40, 2), "miss": (40, 1),
}


def algorithmic_flops(counters: dict) -> tuple:
    f = sum(counters[k] * v[0] for k, v in FLOP_MODEL.items())
    s = sum(counters[k] * v[1] for k, v in FLOP_MODEL.items())
    return f, s


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--stage", choices=["rt", "post", "frame"], default="rt",
                    help="rt = the ray-tracing pass (the headline metric); post = the post chain (bench_post.py)")
    ap.add_argument("--config", type=int, default=2, help="BASELINE.json config number (1..5); the metric is quoted on 2")
    ap.add_argument("--scene", choices=["default", "SIMPLE", "performance_test"], default=None,
                    help="a shipped scene file of the reference instead of a BASELINE config (1920x1080, depth 3)")
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--height", type=int, default=None)
    ap.add_argument("--samples", type=int, default=None, help="override num_samples (config 4 defaults to 64)")
    ap.add_argument("--shard", choices=["samples", "rows"], default="samples")
    ap.add_argument("--gather", choices=["peer", "nccl"], default="peer",
                    help="--shard rows: peer = interleaved bands stored straight into rank 0's frame over NVLink (CUDA IPC); "
                         "nccl = contiguous tiles gathered point to point")
    ap.add_argument("--reduce", choices=["peer", "nccl"], default="peer",
                    help="--shard samples: peer (default) = the library's deterministic pull-reduce kernel over CUDA-IPC-mapped partial "
                         "frames (rt_peer_reduce), pipelined one frame deep: it runs on its own stream in CTA slots the render grid leaves "
                         "free (rt_set_sm_reserve) while the NEXT frame renders; nccl = dist.reduce after each render on the same stream "
                         "(alone it is the faster reduce, N=8 71 us vs 110 us for a 1080p frame, but an NCCL CTA needs a whole free SM)")
    ap.add_argument("--pipeline", type=int, default=1, help="--reduce peer: 1 = overlap the reduce of frame i-1 with the render of frame i")
    ap.add_argument("--sm-reserve", type=int, default=8, help="pipelined peer reduce: SMs' worth of CTA slots the render grid leaves free")
    ap.add_argument("--no-subrecords", action="store_true", help="N > 1: skip the c4_samples / c5_rows sub-records")
    ap.add_argument("--flags", type=lambda s: int(s, 0), default=0, help="extra RT_FLAG_* bits (e.g. 0x2 parity math)")
    ap.add_argument("--host-path", type=int, default=0, help="rt_set_host_path: 0 auto, 1 staged, 2 banded, 3 zero-copy")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample-step", type=int, default=0, help="oracle pixel stride for the CPU baseline (0 = auto)")
    return ap.parse_args()


def workload_for(args):
    from opengl_raytracing_b200 import synthetic
    if args.scene:
        # one of the reference's shipped scene files (res/Scene/*.scene), from the decoded bytes
        # committed under tests/golden (the reference tree does not exist on the GPU box), as the
        # reference would show it at start-up: Camera defaults, light defaults (PCF x4), depth 3
        import numpy as np
        from opengl_raytracing_b200 import scene as S
        g = np.load(os.path.join(ROOT, "tests", "golden", "scenes.npz"))
        objs = g[args.scene + "_objects"].copy().view(S.OBJECT_DTYPE)
        lights = g[args.scene + "_lights"].copy().view(S.LIGHT_DTYPE)
        w = synthetic.Workload(args.scene, objs, lights, args.width or 1920, args.height or 1080, 3, use_skybox=True,
                               skybox=synthetic.skybox_s1(), noise=synthetic.noise_n1(),
                               description=f"res/Scene/{args.scene}.scene ({len(objs)} objects, {len(lights)} lights), Camera defaults, skybox S1, noise N1, depth 3, 1 spp")
        if args.samples is not None:
            w.num_samples = args.samples
        return w
    w = synthetic.config(args.config, args.width, args.height)
    if args.samples is not None:
        w.num_samples = args.samples
    return w


class ClockSampler:
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int, period_s: float = 0.002):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self.period_s, self.power = period_s, []
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
            "hw_power_brake": getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(self.period_s)            # the timed region of a small config lasts ~10 ms

    def start(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._t:
            self._t.join()
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), "sm_mhz_min": s[0] if s else None, "power_w_max": max(self.power) if self.power else None}


def host_threads() -> int:
    """The CPUs this process may run on. torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the CPU
    arm passes this count to the oracle explicitly (an OpenMP num_threads clause), so that setting is ignored."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return os.cpu_count() or 1


def config_for(w, args, world: int, parallelism: str, **extra) -> dict:
    """The ``config`` object — the SAME key set for both arms of the bench."""
    cfg = {"workload": f"{w.name}: {w.description}", "width": w.width, "height": w.height, "objects": len(w.objects),
           "lights": len(w.lights), "max_ray_depth": w.max_ray_depth, "num_samples": max(1, w.num_samples),
           "parallelism": parallelism, "l2": None, "rays_per_frame": None, "rays_per_pixel": None, "math": None,
           "rays": "reference-literal: 1 ray = 1 intersectObjects call of the shader on this input (the kernels prove most "
                   "of them unnecessary and skip them)",
           "variant": None, "note": None}
    cfg.update(extra)
    return cfg


def cpu_baseline(w, kind: str, budget_s: float = 15.0, step: int = 0) -> dict:
    """Times the CPU implementation of the path (oracle/_ref = the reference's own shader source
    compiled as C++, else the port) on a bounded pixel subset with all host threads."""
    import oracle
    if kind == "reference" and not oracle.have("reference"):
        kind = "port"
    if kind == "port":
        oracle.build_port()
    cores = host_threads()
    # calibrate on a sparse subset, then size the timed sample to ~budget_s of CPU work
    oracle.render(w, kind, x0=0, y0=0, sx=1, sy=1, nx=4, ny=4, threads=cores)      # warm: library load, skybox conversion
    t0 = time.perf_counter()
    cal = oracle.render(w, "port", x0=3, y0=5, sx=32, sy=32, threads=cores)
    t_cal = time.perf_counter() - t0
    px_cal = cal["color"].shape[0] * cal["color"].shape[1] * max(1, w.num_samples)
    per_px = t_cal / max(px_cal, 1)
    total_px = w.width * w.height * max(1, w.num_samples)
    if step <= 0:
        step = 1
        while per_px * total_px / (step * step) > budget_s and step < 64:
            step *= 2
    x0, y0 = (3 % step, 5 % step)
    best, rays = None, 0
    for _ in range(2):
        t0 = time.perf_counter()
        oracle.render(w, kind, x0=x0, y0=y0, sx=step, sy=step, threads=cores)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    rays = oracle.render(w, "port", x0=x0, y0=y0, sx=step, sy=step, threads=cores)["counters"]["rays"]
    return {"value": rays / best / 1e6, "unit": METRIC, "cores": cores, "kind": kind,
            "sample": f"{w.name} {w.width}x{w.height}, every {step}-th pixel in x and y (offset {x0},{y0}), "
                      f"{rays} rays in {best:.3f} s, best of 2, OpenMP dynamic over rows on {cores} threads",
            "seconds": best, "rays": rays, "step": step}


def run_reference(args) -> None:
    """The reference's own CPU implementation of the path on the box's host cores: one step =
    one bounded sample of the same frame. Only rank 0 works."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    w = workload_for(args)
    kind = "reference" if oracle.have("reference") else "port"
    base = cpu_baseline(w, kind, budget_s=6.0, step=args.cpu_sample_step)
    step, x0, y0 = base["step"], 3 % base["step"], 5 % base["step"]
    cores = base["cores"]
    for _ in range(max(0, args.warmup - 1)):
        oracle.render(w, kind, x0=x0, y0=y0, sx=step, sy=step, threads=cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.render(w, kind, x0=x0, y0=y0, sx=step, sy=step, threads=cores)
    dt = (time.perf_counter() - t0) / max(1, args.steps)
    val = base["rays"] / dt / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": METRIC, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_for(w, args, 1, f"{cores} host threads (OpenMP), no GPU", math="IEEE FP32, -ffp-contract=off",
                             note="reference shader source compiled as C++ on host cores; GL/llvmpipe unavailable on this box; "
                                  "each step = " + base["sample"]),
        "cpu_baseline": {"value": val, "unit": METRIC, "cores": cores, "kind": base["kind"], "sample": base["sample"]},
        "e2e": {"value": val, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def executed_work(name: str, W: int, H: int) -> dict | None:
    """FP32 thread-instructions one launch of this config EXECUTES, from the committed ncu capture of this build
    (profiles/executed.json, written by tools/ncu_executed.py from a `ncu --set full` report): the algorithmic
    flop count is the reference-literal algorithm's, which the kernels mostly prove unnecessary at C3-C5."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "executed.json"))).get(f"{name}@{W}x{H}")
    except Exception:
        return None


def sustained_leg(ctx, params, out, stream, dev, kern_ms: float, local: int, seconds: float = 2.0) -> dict:
    """>= `seconds` of back-to-back frames (no L2 flush, no host sync in between) with NVML sampled throughout:
    shows whether the clock of the short timed region holds under load."""
    import torch
    n = max(10, int(seconds * 1e3 / max(kern_ms, 1e-3)) + 1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(local, period_s=0.01)
    torch.cuda.synchronize(dev)
    sampler.start()
    e0.record(stream)
    for _ in range(n):
        ctx.render(params, out[0], out[1], out[2], None, stream.cuda_stream)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    c = sampler.stop()
    ms = e0.elapsed_time(e1)
    return {"frames": n, "seconds": ms * 1e-3, "ms_per_frame": ms / n, "sm_mhz_median": c["sm_mhz"], "sm_mhz_min": c.get("sm_mhz_min"),
            "power_w_max": c.get("power_w_max"), "reasons": c["reasons"], "samples": c["samples"]}


def run_ours(args) -> None:
    import numpy as np
    import torch
    import torch.distributed as dist
    from opengl_raytracing_b200 import abi, sharding, synthetic
    from opengl_raytracing_b200.pipeline import RtContext, bind_workload, workload_params

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        try:                                           # pinned host buffers first-touch on the GPU's NUMA node
            import pynvml
            pynvml.nvmlInit()
            hnd = pynvml.nvmlDeviceGetHandleByIndex(local)
            words = pynvml.nvmlDeviceGetCpuAffinity(hnd, (os.cpu_count() + 63) // 64)
            cpus = {64 * i + b for i, wd in enumerate(words) for b in range(64) if (wd >> b) & 1}
            if cpus:
                os.sched_setaffinity(0, cpus)
        except Exception:
            pass
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # (stdout is guarded as well: guard_stdout)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    w = workload_for(args)
    ctx = RtContext(local)
    bind_workload(ctx, w)
    H, W = w.height, w.width
    ns = max(1, w.num_samples)
    flags = args.flags

    # ---- sharding plan
    peer = None
    if world > 1 and args.shard == "rows" and args.gather == "peer":
        params = workload_params(w, flags=flags, **sharding.band_plan(rank, world))
        peer = sharding.PeerFrame(ctx, H, W)
        scaling, par = "strong", (f"interleaved {sharding.BAND_ROWS}-row bands x{world}, kernel stores go straight into "
                                  "rank 0's frame over NVLink (CUDA IPC peer mapping), no gather")
    elif world > 1 and args.shard == "rows":
        ra, rb = sharding.row_range(H, rank, world)
        params = workload_params(w, flags=flags, row_begin=ra, row_end=rb)
        scaling, par = "strong", f"row tiles x{world}, gathered to rank 0"
    elif world > 1:
        params = workload_params(w, flags=flags | abi.FLAG_SUM_SAMPLES, frame_count=w.frame_count + rank * ns)
        scaling, par = "weak", (f"sample ranges x{world} ({ns} spp per rank), partial sums reduced to rank 0 "
                                f"({'peer pull-reduce kernel over CUDA IPC' if args.reduce == 'peer' else 'NCCL reduce'})")
        if args.reduce == "peer" and args.pipeline:
            par += (f"; pipelined one frame deep: the reduce of frame i-1 runs on a second stream, in the CTA slots of {args.sm_reserve} SMs the "
                    "render grid leaves free, inside the timed region of frame i (K timed regions = K renders + K reduces; three rotating partials)")
    else:
        params = workload_params(w, flags=flags)
        scaling, par = "weak", "single GPU"

    color = torch.zeros(H, W, 4, dtype=torch.float32, device=dev)
    position = torch.zeros(H, W, 4, dtype=torch.float32, device=dev)
    normal = torch.zeros(H, W, 4, dtype=torch.float16, device=dev)
    write_geom = (world == 1) or args.shard == "rows" or rank == 0
    stream = torch.cuda.current_stream(dev)
    reducer = None
    if world > 1 and args.shard == "samples" and args.reduce == "peer":
        try:
            reducer = sharding.PeerReducer(ctx, H, W)  # partial frames of every rank mapped into every rank (CUDA IPC)
            color = reducer.partial                     # (the counter pass; every frame renders into reducer.partial of ITS epoch)
        except sharding.PeerUnavailable as e:           # raised on every rank together: same fallback everywhere
            if rank == 0:
                print(f"bench: peer reduce unavailable ({e}); falling back to --reduce nccl", file=sys.stderr)
            reducer, args.reduce = None, "nccl"
            args.no_subrecords = True                   # c5_rows maps rank 0's frame the same way
            par = f"sample ranges x{world} ({ns} spp per rank), partial sums reduced to rank 0 (NCCL reduce; peer mapping unavailable on this box)"
    pipelined = bool(reducer and args.pipeline)
    if pipelined:
        ctx.set_sm_reserve(args.sm_reserve)

    # ---- reference-literal op counts for THIS rank's work (untimed, counter variant)
    pc = abi.FrameParams.from_buffer_copy(params)
    pc.flags = (params.flags | abi.FLAG_COUNTERS) & ~(abi.FLAG_ACCUMULATE | abi.FLAG_SUM_SAMPLES)
    ctx.counters(reset=True)
    ctx.render(pc, color, None, None, None, stream.cuda_stream)
    counters = ctx.counters(reset=True)
    rays_rank = counters["rays"]
    flops_rank, sfu_rank = algorithmic_flops(counters)

    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)   # > 126 MB L2

    out = (peer.color, peer.position, peer.normal) if peer else (color, position if write_geom else None, normal if write_geom else None)

    def target():                                      # this frame's images: the peer reducer rotates its partial buffers by epoch
        return (reducer.partial, out[1], out[2]) if reducer else out

    def collective():
        if peer:
            dist.barrier()                             # the frame is complete on rank 0 when every rank's kernel is
        elif world > 1 and args.shard == "rows":
            for buf in (color, position, normal):
                sharding.gather_rows(buf, H, world)
        elif reducer:
            reducer.reduce(stream.cuda_stream, ns * world)
        elif world > 1:
            sharding.reduce_samples(color, ns * world)

    def step():
        if params.flags & abi.FLAG_ACCUMULATE:
            color.zero_()
        o = target()
        ctx.render(params, o[0], o[1], o[2], None, stream.cuda_stream)
        collective()

    class ReducePipeline:
        """Frame i renders on `stream` while the peer reduce of frame i-1 runs on `side`. The reduce is enqueued behind
        the START event of frame i's timed region (never inside the untimed L2 flush), and the region ends only when
        both the render and that reduce are done: region i = max(render i, reduce i-1) + what they cost each other.
        The render of frame i is ordered behind this rank's reduce of frame i-2 (sharding.PeerReducer: three partials)."""

        def __init__(self):
            self.side = torch.cuda.Stream(dev)
            self.pending = None                        # render-done event of the frame whose reduce has not been enqueued
            self.red_done = []

        def _reduce_pending(self, after):
            self.side.wait_event(after)
            self.side.wait_event(self.pending)
            reducer.reduce(self.side.cuda_stream, ns * world)
            e = torch.cuda.Event(); e.record(self.side)
            self.red_done = self.red_done[-1:] + [e]
            self.pending = None

        def frame(self, events=None, do_flush=False):
            if do_flush:
                flush.zero_()
            a, b, c = events if events else (torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event())
            a.record(stream)
            had = self.pending is not None
            if had:
                self._reduce_pending(a)
            if len(self.red_done) >= 2 or (self.red_done and not had):
                stream.wait_event(self.red_done[0] if len(self.red_done) >= 2 else self.red_done[-1])
            o = target()
            ctx.render(params, o[0], o[1], o[2], None, stream.cuda_stream)
            b.record(stream)
            self.pending = b
            if had:
                stream.wait_event(self.red_done[-1])
            c.record(stream)

        def drain(self):
            if self.pending is not None:
                a = torch.cuda.Event(); a.record(stream)
                self._reduce_pending(a)
            if self.red_done:
                stream.wait_event(self.red_done[-1])

    pipe = ReducePipeline() if pipelined else None

    # ---- the variant is chosen ONCE, explicitly (rt_tune: every applicable variant timed on this scene and
    # frame shape, identical images), so that no warm-up or timed frame is a trial frame of the autotuner
    ctx.set_autotune(False)
    tuning = ctx.tune(params)
    nwarm = max(3, args.warmup)
    for _ in range(nwarm):
        pipe.frame() if pipe else step()
    torch.cuda.synchronize(dev)
    fp32_peak = ctx.measure_fp32_peak()
    launches0 = ctx.launch_count()

    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True),
           torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    sampler = ClockSampler(local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    sampler.start()
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        if pipe:
            pipe.frame(ev[i], do_flush=True)
            continue
        flush.zero_()                                  # L2 flush between timed iterations (untimed)
        ev[i][0].record(stream)
        if params.flags & abi.FLAG_ACCUMULATE:
            color.zero_()
        o = target()
        ctx.render(params, o[0], o[1], o[2], None, stream.cuda_stream)
        ev[i][1].record(stream)
        collective()
        ev[i][2].record(stream)
    if pipe:
        pipe.drain()                                   # the last frame's reduce (its twin, the last warm-up frame's, ran in region 0)
    torch.cuda.synchronize(dev)
    if reducer:
        reducer.check()
    if world > 1:
        dist.barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop()
    launches = ctx.launch_count() - launches0
    step_ms = sum(a.elapsed_time(c) for a, _, c in ev) / args.steps
    kern_ms = sum(a.elapsed_time(b) for a, b, _ in ev) / args.steps    # includes the zero_ fill when accumulating
    t = torch.tensor([step_ms, kern_ms, float(rays_rank), float(flops_rank)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        step_ms, kern_ms = float(tmax[0]), float(tmax[1])
        rays_total, flops_total = float(tsum[2]), float(tsum[3])
    else:
        rays_total, flops_total = float(rays_rank), float(flops_rank)
    value = rays_total / (step_ms * 1e-3) / 1e6

    sustained = sustained_leg(ctx, params, target(), stream, dev, kern_ms, local) if rank == 0 or world > 1 else None
    if world > 1:
        dist.barrier()

    # ---- e2e: the SAME job through host buffers. Every step: scene H2D on every rank (ssbo.update()), this rank's
    # render, the exchange step of the sharding, and the finished frame's three images landing in pinned host
    # memory on rank 0. N = 1: the single C-ABI call rt_render_host does all of it.
    if params.row_band > 0:
        nb = (H + params.row_band - 1) // params.row_band
        px = sum(min(params.row_band, H - k * params.row_band) for k in range(params.row_phase, nb, params.row_period)) * W
    else:
        px = (params.row_end - params.row_begin) * W
    objs = np.ascontiguousarray(w.objects); lights = np.ascontiguousarray(w.lights)
    h2d = objs.nbytes + lights.nbytes
    if world == 1:
        h = {}
        for name in ("color", "position", "normal"):
            p = abi.C.c_void_p(0)
            assert ctx.lib.rt_host_alloc(abi.C.byref(p), H * W * (16 if name != "normal" else 8)) == 0
            h[name] = p
        pe = abi.FrameParams.from_buffer_copy(params)
        assert ctx.lib.rt_set_host_path(ctx.handle, args.host_path) == 0

        def e2e_step():
            st = ctx.lib.rt_render_host(ctx.handle, abi.C.byref(pe), objs.ctypes.data, len(objs), lights.ctypes.data, len(lights),
                                        h["color"], h["position"], h["normal"], None)
            assert st == 0, ctx.lib.rt_last_error(ctx.handle)
        e2e_call = "rt_render_host: scene H2D (pinned staging) + rt_render + D2H of colour/position/normal into pinned host buffers"
        d2h = px * 40
    else:
        hbuf = None
        if rank == 0:
            hbuf = (torch.empty(H, W, 4, dtype=torch.float32).pin_memory(), torch.empty(H, W, 4, dtype=torch.float32).pin_memory(),
                    torch.empty(H, W, 4, dtype=torch.float16).pin_memory())
        frame = peer.tensors() if (peer and rank == 0) else (color, position, normal)

        def e2e_step():
            ctx.upload_scene(objs, lights, stream.cuda_stream, async_=True)
            step()
            if rank == 0:
                final = (reducer.result if reducer else frame[0], frame[1], frame[2])
                for dst, src in zip(hbuf, final):
                    dst.copy_(src, non_blocking=True)
            torch.cuda.synchronize(dev)
        e2e_call = ("every rank: rt_upload_scene (H2D) + rt_render of its shard; the sharding's exchange step; rank 0: D2H of the "
                    "finished colour/position/normal into pinned host memory")
        d2h = H * W * 40
    for _ in range(3):
        e2e_step()
    if world > 1:
        dist.barrier()
    e2e_steps = max(5, min(args.steps, 20))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    if world > 1:
        dist.barrier()
    e2e_ms = (time.perf_counter() - t0) / e2e_steps * 1e3
    te = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_ms = float(te[0])
    if world == 1:
        for p in h.values():
            ctx.lib.rt_host_free(p)

    sub = None
    if world > 1 and not args.no_subrecords:
        try:                                           # the sub-records must never cost the headline line
            sub = multi_gpu_subrecords(args, rank, world, local)
        except Exception as e:                         # noqa: BLE001 (every rank fails alike: same code, same inputs)
            import traceback
            traceback.print_exc()
            sub = {"subrecords_error": f"{type(e).__name__}: {e}"}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        ach_tf = flops_rank / (kern_ms * 1e-3) / 1e12
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(f"{w.name}@{W}x{H}", {}).get("bytes")
        except Exception:
            pass
        roof = {"bound": "fp32", "achieved": ach_tf, "peak": fp32_peak["tflops"], "unit": "TFLOP/s",
                "frac": ach_tf / fp32_peak["tflops"], "traffic": traffic,
                "traffic_source": "profiles/traffic.json: dram__bytes_read.sum + dram__bytes_write.sum of one launch from the committed "
                                  "ncu --set full capture of this config (a constant of the build, not measured in this run)",
                "peak_source": "measured in this run: FFMA-only micro-kernel (rt_measure_fp32_peak); "
                               "MEASURED_PEAKS.json has no FP32 entry",
                "algorithmic_flops_per_launch": flops_rank, "algorithmic_sfu_ops_per_launch": sfu_rank,
                "flop_model": "v0 (SURVEY.md 8d), reference-literal op counts x fixed constants",
                "hbm_write": {"bytes_per_launch": px * 40, "achieved_gbs": px * 40 / (kern_ms * 1e-3) / 1e9,
                              "peak_gbs": peaks.get("hbm_gbs"), "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)"}}
        ex = executed_work(w.name, W, H)
        if ex:
            fl = ex["fadd"] + ex["fmul"] + 2 * ex["ffma"]
            etf = fl / (kern_ms * 1e-3) / 1e12
            roof["executed"] = {"fp32_flops_per_launch": fl, "achieved": etf, "frac": etf / fp32_peak["tflops"], "unit": "TFLOP/s",
                                "fp32_lane_utilisation": (ex["fadd"] + ex["fmul"] + ex["ffma"]) / (kern_ms * 1e-3) / (fp32_peak["tflops"] * 1e12 / 2),
                                "source": ex.get("source"), "kernels": ex.get("kernels"),
                                "note": "FADD + FMUL + 2 FFMA thread-instructions of the committed ncu capture of this config and build / this run's time"}
        line = {
            "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "warmup_effective": nwarm, "ms_per_step": step_ms, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": config_for(w, args, world, par, l2="flushed between timed iterations (256 MB fill, untimed)",
                                 rays_per_frame=rays_rank, rays_per_pixel=rays_rank / (px * 1.0),
                                 math="parity" if (flags & abi.FLAG_PARITY_MATH) else "fast", variant=tuning),
            "ms_per_frame": kern_ms,
            "roofline": roof,
            "e2e": {"value": rays_total / (e2e_ms * 1e-3) / 1e6, "unit": METRIC, "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h, "call": e2e_call},
            "gpu_launches": launches,
            "clocks": clocks,
            "sustained": sustained,
            "wall_s_timed_region": t_wall,
        }
        if sub:
            line.update(sub)
        if not args.no_cpu_baseline and world == 1:
            import oracle
            kind = "reference" if oracle.have("reference") else "port"
            cb = cpu_baseline(w, kind, budget_s=12.0, step=args.cpu_sample_step)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        emit(line)
    if reducer:
        torch.cuda.synchronize(dev)
        dist.barrier()
        reducer.close()
    if peer:
        torch.cuda.synchronize(dev)
        dist.barrier()
        peer.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def multi_gpu_subrecords(args, rank: int, world: int, local: int) -> dict:
    """The two configs north_star names for sharding, run in the same processes after the headline job:
      c4_samples  3840x2160, 64 spp split into `world` sample ranges + the reduce to rank 0 (strong scaling)
      c5_rows     7680x4320, interleaved 32-row bands, kernel stores straight into rank 0's frame over NVLink
    each with the same job on ONE GPU of the same box (every rank runs it on its own GPU at the same time; rank 0's
    time is reported) for the speed-up."""
    import torch
    import torch.distributed as dist
    from opengl_raytracing_b200 import abi, sharding, synthetic
    from opengl_raytracing_b200.pipeline import RtContext, bind_workload, workload_params
    dev = torch.device("cuda", local)
    stream = torch.cuda.current_stream(dev)
    out = {}

    def timed(fn, steps, warm=1):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize(dev); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize(dev)
        t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=dev)
        t0 = t.clone()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t0[0])

    def pin_variant(ctx, w, **over):
        """Choose the variant on a 1-sample frame of the same shape (seconds, not minutes) and pin it."""
        ctx.set_autotune(False)
        t = ctx.tune(workload_params(w, num_samples=1, **over))
        ctx.set_tuning(abi.KERNEL_IDS[t["kernel"]], 1 if t["registers"] == 80 else 0, t["phase_level"])
        return t

    # ---- C4: sample ranges
    w = synthetic.config(4)
    H, W, S = w.height, w.width, w.num_samples
    ctx = RtContext(local)
    bind_workload(ctx, w)
    var = pin_variant(ctx, w)
    first, count = sharding.sample_range(S, rank, world)
    use_peer = args.reduce == "peer"
    reducer = sharding.PeerReducer(ctx, H, W) if use_peer else None
    part0 = None if reducer else torch.zeros(H, W, 4, device=dev)
    pos = torch.zeros(H, W, 4, device=dev) if rank == 0 else None
    nrm = torch.zeros(H, W, 4, dtype=torch.float16, device=dev) if rank == 0 else None
    p = workload_params(w, flags=abi.FLAG_SUM_SAMPLES, frame_count=w.frame_count + first, num_samples=max(count, 1))

    def c4_step():
        part = reducer.partial if reducer else part0   # the reducer rotates its partial buffers by epoch
        if count > 0:
            ctx.render(p, part, pos, nrm, None, stream.cuda_stream)
        else:
            part.zero_()
        if reducer:
            reducer.reduce(stream.cuda_stream, S)
        else:
            sharding.reduce_samples(part, S)
    ms_n, _ = timed(c4_step, steps=2)
    full = torch.zeros(H, W, 4, device=dev)
    p1 = workload_params(w)
    _, ms_1 = timed(lambda: ctx.render(p1, full, None, None, None, stream.cuda_stream), steps=1)
    out["c4_samples"] = {"workload": f"{w.name}: {w.description}", "ms_per_step": ms_n, "ms_per_step_1gpu_same_box": ms_1,
                         "speedup": ms_1 / ms_n, "efficiency": ms_1 / ms_n / world, "samples_per_rank": count,
                         "exchange": ("peer pull-reduce kernel (rank r sums slice r of all partials in rank order, RGB, into rank 0's frame)"
                                      if use_peer else "NCCL reduce of the RGBA32F partial sums to rank 0"),
                         "nvlink_bytes_into_rank0": (world - 1) * H * W * 16 // world if use_peer else None,
                         "variant": var}
    if reducer:
        torch.cuda.synchronize(dev); dist.barrier(); reducer.close()
    del part0, full
    ctx.close()

    # ---- C5: interleaved row bands, fused peer stores
    w = synthetic.config(5)
    H, W = w.height, w.width
    ctx = RtContext(local)
    bind_workload(ctx, w)
    bp = sharding.band_plan(rank, world)
    var = pin_variant(ctx, w, **bp)
    peer = sharding.PeerFrame(ctx, H, W)
    p = workload_params(w, **bp)

    def c5_step():
        ctx.render(p, peer.color, peer.position, peer.normal, None, stream.cuda_stream)
        dist.barrier()
    ms_n, _ = timed(c5_step, steps=5, warm=2)
    c = torch.zeros(H, W, 4, device=dev); po = torch.zeros(H, W, 4, device=dev); no = torch.zeros(H, W, 4, dtype=torch.float16, device=dev)
    ctx.set_tuning(None)
    var1 = pin_variant(ctx, w)
    p1 = workload_params(w)
    _, ms_1 = timed(lambda: ctx.render(p1, c, po, no, None, stream.cuda_stream), steps=3, warm=1)
    out["c5_rows"] = {"workload": f"{w.name}: {w.description}", "ms_per_step": ms_n, "ms_per_step_1gpu_same_box": ms_1,
                      "speedup": ms_1 / ms_n, "efficiency": ms_1 / ms_n / world,
                      "exchange": f"none: interleaved {sharding.BAND_ROWS}-row bands, float4 stores over NVLink into rank 0's frame, one barrier",
                      "nvlink_bytes_into_rank0": (world - 1) * H * W * 40 // world, "variant": var, "variant_1gpu": var1}
    torch.cuda.synchronize(dev); dist.barrier()
    peer.close()
    ctx.close()
    return out


_REAL_STDOUT = None


def guard_stdout() -> None:
    """stdout carries exactly ONE line — the JSON result. Libraries loaded below (NCCL's version
    banner, torch.distributed notices) write to file descriptor 1 on their own: point fd 1 at
    stderr for the duration of the run and keep the real stdout for emit()."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line: dict) -> None:
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    args = parse_args()
    guard_stdout()
    if args.stage == "post":
        import bench_post
        (bench_post.run_reference if args.impl == "reference" else bench_post.run_ours)(args, emit)
        return
    if args.stage == "frame":
        import bench_frame
        (bench_frame.run_reference if args.impl == "reference" else bench_frame.run_ours)(args, emit)
        return
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
