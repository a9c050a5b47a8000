"""Dev helper (GPU box, under ncu): renders one BASELINE config with a pinned variant.
usage: python tools/render_once.py CONFIG megakernel|wavefront|regroup [frames] [registers] [phase_level] [WxH]"""
import sys
sys.path.insert(0, ".")
import torch
from opengl_raytracing_b200 import abi, synthetic
from opengl_raytracing_b200.pipeline import RtContext, bind_workload, workload_params

n, kern = int(sys.argv[1]), sys.argv[2]
frames = int(sys.argv[3]) if len(sys.argv) > 3 else 2
regs = int(sys.argv[4]) if len(sys.argv) > 4 else -1
phase = int(sys.argv[5]) if len(sys.argv) > 5 else -1
size = [int(x) for x in sys.argv[6].split("x")] if len(sys.argv) > 6 else [None, None]
w = synthetic.config(n, *size)
if n == 4:
    w.num_samples = 1
ctx = RtContext(0)
bind_workload(ctx, w)
ctx.set_autotune(False)
ctx.set_tuning(abi.KERNEL_IDS[kern], regs, phase)
H, W = w.height, w.width
c = torch.zeros(H, W, 4, device="cuda"); p = torch.zeros(H, W, 4, device="cuda"); nn = torch.zeros(H, W, 4, dtype=torch.float16, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for f in range(frames):
    e0.record()
    ctx.render(workload_params(w), c, p, nn, None, torch.cuda.current_stream().cuda_stream)
    e1.record()
    torch.cuda.synchronize()
    print(f"frame {f}: {e0.elapsed_time(e1):.4f} ms", ctx.get_tuning(), flush=True)
ctx.close()
