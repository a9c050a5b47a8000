"""Worker of test_peer_reduce_two_processes (launched under torchrun, 2 ranks)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opengl_raytracing_b200 import abi, sharding, synthetic  # noqa: E402
from opengl_raytracing_b200.pipeline import RtContext, bind_workload, render_workload, workload_params  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    gpus = int(os.environ.get("PEER_GPUS", "1"))
    device = rank % gpus
    torch.cuda.set_device(device)
    dist.init_process_group("gloo")
    ctx = RtContext(device)
    w = synthetic.config(4, 320, 180)
    w.num_samples = 6
    bind_workload(ctx, w)
    red = sharding.PeerReducer(ctx, w.height, w.width)
    stream = torch.cuda.current_stream().cuda_stream
    first, count = sharding.sample_range(w.num_samples, rank, world)
    ok = True
    for frame in range(3):
        fc = w.frame_count + 10 * frame
        p = workload_params(w, flags=abi.FLAG_SUM_SAMPLES, frame_count=fc + first, num_samples=count)
        ctx.render(p, red.partial, None, None, None, stream)
        red.reduce(stream, w.num_samples)
        torch.cuda.synchronize()
        red.check()
        dist.barrier()
        if rank == 0:
            w.frame_count = fc
            full = render_workload(ctx, w, bind=False)["color"]
            w.frame_count -= 10 * frame
            got = red.result.cpu().numpy()
            fin = np.isfinite(full).all(axis=-1)
            err = float(np.abs(got[fin] - full[fin]).max())
            good = err <= 2e-5 * max(1.0, float(np.abs(full[fin]).max())) and bool((got[..., 3] == 1.0).all())
            print(f"frame {frame}: max |mean - one-GPU mean| = {err:.3g} {'ok' if good else 'MISMATCH'}", flush=True)
            ok = ok and good
        dist.barrier()
    # ---- pipelined: the reduce of frame i-1 on a second stream beside the render of frame i (three rotating partials,
    # the render of frame i ordered behind this rank's reduce of frame i-2), every frame different
    ctx.set_sm_reserve(4)
    main, side = torch.cuda.current_stream(), torch.cuda.Stream()
    frames, results, red_done, pending = 6, [], [], None
    for frame in range(frames + 1):
        if pending is not None:
            side.wait_event(pending)
            red.reduce(side.cuda_stream, w.num_samples)
            if rank == 0:
                with torch.cuda.stream(side):
                    results.append(red.result.clone())
            e = torch.cuda.Event(); e.record(side); red_done.append(e)
            pending = None
        if frame == frames:
            break
        if len(red_done) >= 2:
            main.wait_event(red_done[-2])
        fc = w.frame_count + 7 * (frame + 1)
        p = workload_params(w, flags=abi.FLAG_SUM_SAMPLES, frame_count=fc + first, num_samples=count)
        ctx.render(p, red.partial, None, None, None, main.cuda_stream)
        pending = torch.cuda.Event(); pending.record(main)
    torch.cuda.synchronize()
    red.check()
    dist.barrier()
    ctx.set_sm_reserve(0)
    if rank == 0:
        base = w.frame_count
        for frame, got_t in enumerate(results):
            w.frame_count = base + 7 * (frame + 1)
            full = render_workload(ctx, w, bind=False)["color"]
            got = got_t.cpu().numpy()
            fin = np.isfinite(full).all(axis=-1)
            err = float(np.abs(got[fin] - full[fin]).max())
            good = err <= 2e-5 * max(1.0, float(np.abs(full[fin]).max())) and bool((got[..., 3] == 1.0).all())
            print(f"pipelined frame {frame}: max |mean - one-GPU mean| = {err:.3g} {'ok' if good else 'MISMATCH'}", flush=True)
            ok = ok and good
        w.frame_count = base
        ok = ok and len(results) == frames
    dist.barrier()
    if rank == 0:
        print("PEER_REDUCE_OK" if ok else "PEER_REDUCE_MISMATCH", flush=True)
    red.close()
    ctx.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
