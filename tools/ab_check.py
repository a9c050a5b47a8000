"""Dev helper (GPU box): images with a dev switch of the library off / on (env var read by rt_create),
e.g. `python tools/ab_check.py RT_PK_GRID 3 5 4` — differing pixels per config, both math builds."""
import os, sys
import numpy as np
sys.path.insert(0, ".")
from opengl_raytracing_b200 import abi, synthetic
from opengl_raytracing_b200.pipeline import RtContext, render_workload

var = sys.argv[1]
SIZES = {1: (1280, 720), 2: (1920, 1080), 3: (960, 540), 4: (480, 270), 5: (1920, 1080)}
for n in [int(a) for a in sys.argv[2:]] or [3, 5]:
    W, H = SIZES[n]
    w = synthetic.config(n, W, H)
    if n == 4:
        w.num_samples = 4
    for name, fl in (("fast", abi.FLAG_NORMAL_FP32), ("parity", abi.FLAG_NORMAL_FP32 | abi.FLAG_PARITY_MATH)):
        out = {}
        for v in ("0", "1"):
            os.environ[var] = v
            ctx = RtContext(0)
            out[v] = render_workload(ctx, w, flags=fl)
            ctx.close()
        same = all(np.array_equal(out["0"][k].view(np.uint32), out["1"][k].view(np.uint32)) for k in ("color", "position", "normal"))
        d = np.abs(out["0"]["color"] - out["1"]["color"]).max(axis=2)
        print(f"C{n} {W}x{H} {name}: bit-identical {same}; pixels differing {int((np.isnan(d) | (d > 0)).sum())}")
