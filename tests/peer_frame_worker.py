"""Worker of test_peer_frame_over_cuda_ipc_two_processes (launched under torchrun, 2 ranks)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opengl_raytracing_b200 import sharding, synthetic  # noqa: E402
from opengl_raytracing_b200.pipeline import RtContext, bind_workload, render_workload, workload_params  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    gpus = int(os.environ.get("PEER_GPUS", "1"))
    device = rank % gpus
    torch.cuda.set_device(device)
    dist.init_process_group("gloo")
    ctx = RtContext(device)
    w = synthetic.config(3, 384, 216)
    bind_workload(ctx, w)
    frame = sharding.PeerFrame(ctx, w.height, w.width)
    if frame.is_root:
        for t in frame.tensors():
            t.fill_(-3.0)
        torch.cuda.synchronize()
    dist.barrier()
    p = workload_params(w, **sharding.band_plan(rank, world, band_rows=8))
    ctx.render(p, frame.color, frame.position, frame.normal, None, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    dist.barrier()
    ok = True
    if frame.is_root:
        full = render_workload(ctx, w, bind=False)
        c, pos, n = frame.tensors()
        ok = (np.array_equal(c.cpu().numpy().view(np.uint32), full["color"].view(np.uint32))
              and np.array_equal(pos.cpu().numpy().view(np.uint32), full["position"].view(np.uint32))
              and np.array_equal(n.float().cpu().numpy(), full["normal"]))
        print("PEER_FRAME_OK" if ok else "PEER_FRAME_MISMATCH", flush=True)
    dist.barrier()
    frame.close()
    ctx.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
