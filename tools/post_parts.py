"""Dev tool (GPU box): times the parts of the post chain (chain, bloom only, TAA only, one fused pair, literal passes), L2 flushed and warm.
usage: python tools/post_parts.py [width height]"""
import sys, numpy as np, torch
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
from opengl_raytracing_b200 import post as P, synthetic
from opengl_raytracing_b200.pipeline import RtContext, render_workload
W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1080)
ctx = RtContext(0)
k = render_workload(ctx, synthetic.config(2, W, H))
color = torch.from_numpy(k["color"]).cuda(); normal = torch.from_numpy(k["normal"]).cuda().half()
hist = torch.from_numpy(np.roll(k["color"], 3, axis=1).copy()).cuda()
disp = torch.empty_like(color); hout = torch.empty_like(color)
pc = P.PostChain(ctx)
flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
def timeit(fn, flush_l2=True, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        if flush_l2: flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); tot += a.elapsed_time(b)
    return tot / n * 1e3
jx, jy = P.taa_jitter(0, W, H)
p = P.post_params(W, H, jitter=(jx, jy))
p_nt = P.post_params(W, H, jitter=(jx, jy), taa=False)
p_lit = P.post_params(W, H, jitter=(jx, jy), flags=1)
p2 = P.post_params(W, H, jitter=(jx, jy), taa=False, passes=2)
for fl in (True, False):
    print("flush" if fl else "warm",
          "chain %.1f" % timeit(lambda: pc.post_chain(p, color, normal, hist, display=disp, history_out=hout), fl),
          "bloom-only %.1f" % timeit(lambda: pc.post_chain(p_nt, color, None, None, display=disp), fl),
          "taa-only %.1f" % timeit(lambda: pc.taa_resolve(p, color, hist, normal), fl),
          "1 pair(first+last) %.1f" % timeit(lambda: pc.post_chain(p2, color, None, None, display=disp), fl),
          "literal chain %.1f" % timeit(lambda: pc.post_chain(p_lit, color, normal, hist, display=disp, history_out=hout), fl), "us")
