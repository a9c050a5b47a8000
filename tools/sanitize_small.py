"""Dev tool (GPU box): tiny renders of every config + the edge cases through both kernels, for
`compute-sanitizer --tool memcheck` (one tool per gpurun call)."""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from opengl_raytracing_b200 import synthetic, abi
from opengl_raytracing_b200.pipeline import RtContext, render_workload
from edge_cases import edge_workloads
ctx = RtContext(0)
n = 0
for cfg, (W, H) in {1: (64, 36), 2: (64, 36), 3: (48, 28), 4: (32, 20), 5: (48, 28)}.items():
    w = synthetic.config(cfg, W, H)
    if cfg == 4: w.num_samples = 2
    for fl in (0, abi.FLAG_PARITY_MATH, abi.FLAG_LITERAL_KERNEL):
        render_workload(ctx, w, flags=fl); n += 1
for name, w in edge_workloads(size=(24, 16)).items():
    render_workload(ctx, w); n += 1
ctx.close()
print("rendered", n, "frames")
