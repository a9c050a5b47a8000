"""Dev helper (GPU box): the fast build's analytic PCSS probe classification (RT_PK_FAN) against the
exact per-ray probes (RT_PK_FAN=0) — pixels whose colour differs, per config."""
import os, sys, json
import numpy as np
sys.path.insert(0, ".")
from opengl_raytracing_b200 import synthetic
from opengl_raytracing_b200.pipeline import RtContext, render_workload

for n, (W, H) in [(2, (1920, 1080)), (3, (960, 540)), (4, (480, 270)), (1, (1280, 720))]:
    w = synthetic.config(n, W, H)
    if n == 4:
        w.num_samples = 4
    out = {}
    for fan in ("0", "1"):
        os.environ["RT_PK_FAN"] = fan
        ctx = RtContext(0)
        out[fan] = render_workload(ctx, w)
        ctx.close()
    a, b = out["0"]["color"], out["1"]["color"]
    d = np.abs(a - b).max(axis=2)
    bad = np.isnan(d) | (d > 0)
    print(f"C{n} {W}x{H}: pixels differing {int(bad.sum())} of {W * H} ({100.0 * bad.mean():.5f} %), max |diff| {float(np.nanmax(d)):.4g}, "
          f"> 1e-3: {int((d > 1e-3).sum())}; P/N identical: {np.array_equal(out['0']['position'], out['1']['position'], equal_nan=True)}")
