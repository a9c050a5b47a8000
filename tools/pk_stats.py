"""Dev tool (GPU box): packet-kernel statistics build. usage: RT_B200_LIB=.../librt_b200_stats.so python tools/pk_stats.py <config> [w h]"""
import sys, json
sys.path.insert(0, ".")
from opengl_raytracing_b200 import synthetic, abi
from opengl_raytracing_b200.pipeline import RtContext, render_workload
n = int(sys.argv[1]); wh = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (None, None)
w = synthetic.config(n, *wh)
ctx = RtContext(0)
ctx.counters(reset=True)
render_workload(ctx, w)
c = ctx.counters()
names = ["jobs", "jobs_no_cand", "jobs_overflow", "cands_in_lists", "cone_tests", "packets", "packet_obj_visits", "closest_rays",
         "cluster_slabs", "member_slabs", "closest_linear_fallback"]
vals = list(c.values())[:len(names)]
px = w.width * w.height
print(json.dumps({k: v for k, v in zip(names, vals)}))
print({k: round(v / px, 3) for k, v in zip(names, vals)}, "per pixel")
