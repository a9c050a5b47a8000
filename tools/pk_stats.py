"""Dev tool (GPU box): packet-kernel statistics build. usage: RT_B200_LIB=.../librt_b200_stats.so python tools/pk_stats.py <config> [kernel] [w h]"""
import sys, json
sys.path.insert(0, ".")
from opengl_raytracing_b200 import synthetic, abi
from opengl_raytracing_b200.pipeline import RtContext, render_workload
n = int(sys.argv[1])
kern = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].isdigit() else None
rest = sys.argv[3:] if kern else sys.argv[2:]
wh = (int(rest[0]), int(rest[1])) if len(rest) > 1 else (None, None)
w = synthetic.config(n, *wh)
ctx = RtContext(0)
ctx.set_autotune(False)
if kern:
    ctx.set_tuning(abi.KERNEL_IDS[kern])
ctx.counters(reset=True)
render_workload(ctx, w)
c = ctx.counters()
names = ["jobs", "jobs_no_cand", "jobs_overflow", "cands_in_lists", "cone_tests", "packets", "packet_obj_visits", "closest_rays",
         "cluster_slabs", "member_slabs", "closest_linear_fallback", "coop_warps", "coop_lists", "coop_clusters", "coop_spheres",
         "coop_incoherent", "coop_cluster_overflow", "coop_sphere_overflow", "coop_entered"]
vals = list(c.values())[:len(names)]
px = w.width * w.height
print(json.dumps({k: v for k, v in zip(names, vals)}))
print({k: round(v / px, 3) for k, v in zip(names, vals)}, "per pixel")
d = dict(zip(names, vals))
if d["coop_warps"]:
    print("coop: %.1f%% of converged warps entered, %.1f%% listed; mean clusters %.1f spheres %.1f; incoherent %.1f%% cl-ovf %.1f%% sph-ovf %.1f%%" % (
        100 * d["coop_entered"] / d["coop_warps"], 100 * d["coop_lists"] / d["coop_warps"],
        d["coop_clusters"] / max(d["coop_lists"], 1), d["coop_spheres"] / max(d["coop_lists"], 1),
        100 * d["coop_incoherent"] / d["coop_warps"], 100 * d["coop_cluster_overflow"] / d["coop_warps"], 100 * d["coop_sphere_overflow"] / d["coop_warps"]))
