"""Dev helper (GPU box, under torchrun): ms per rt_peer_reduce alone (no render) on a W x H frame, N ranks,
vs dist.reduce of the same bytes. usage: torchrun --nproc-per-node N tools/peer_reduce_bench.py [W H]"""
import os, sys
sys.path.insert(0, ".")
import torch, torch.distributed as dist
from opengl_raytracing_b200 import sharding
from opengl_raytracing_b200.pipeline import RtContext

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1080)
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = RtContext(local)
red = sharding.PeerReducer(ctx, H, W)
st = torch.cuda.current_stream()
for e in range(2):                       # both partial buffers hold something
    red.partial.fill_(float(rank + 1)); red.reduce(st.cuda_stream, world)
torch.cuda.synchronize(); dist.barrier()
def timed(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(n): fn()
    e1.record(st); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])
ms_peer = timed(lambda: red.reduce(st.cuda_stream, world))
red.check()
buf = torch.ones(H, W, 4, device="cuda")
ms_nccl = timed(lambda: dist.reduce(buf, dst=0))
if rank == 0:
    ok = bool((red.result[..., 0] == sum(range(1, world + 1)) / world).all())
    print(f"N={world} {W}x{H}: peer reduce {ms_peer*1e3:.1f} us, NCCL reduce {ms_nccl*1e3:.1f} us, result ok {ok}", flush=True)
torch.cuda.synchronize(); dist.barrier()
red.close(); ctx.close(); dist.destroy_process_group()
