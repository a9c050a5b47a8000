"""Dev helper (GPU box): ms per frame of every kernel variant (rt_tune: megakernel register budgets x phase levels,
wavefront) on the BASELINE configs at their own sizes, L2-cold style (best of rt_tune's trials).
usage: python tools/variants.py [configs...] [--chunks 1048576,2097152,...] [--json out.json]"""
import json, os, sys
sys.path.insert(0, ".")
from opengl_raytracing_b200 import abi, synthetic
from opengl_raytracing_b200.pipeline import RtContext, bind_workload, workload_params

args = sys.argv[1:]
chunks, out_json = [None], None
cfgs = []
i = 0
while i < len(args):
    if args[i] == "--chunks":
        chunks = [int(x) for x in args[i + 1].split(",")]; i += 2
    elif args[i] == "--json":
        out_json = args[i + 1]; i += 2
    else:
        cfgs.append(int(args[i])); i += 1
cfgs = cfgs or [1, 2, 3, 5]
res = {}
for n in cfgs:
    w = synthetic.config(n)
    if n == 4:
        w.num_samples = 2
    for ch in chunks:
        if ch is not None:
            os.environ["RT_WF_CHUNK"] = str(ch)
        ctx = RtContext(0)
        bind_workload(ctx, w)
        t = ctx.tune(workload_params(w))
        t2 = ctx.tune(workload_params(w))                      # a second pass: warm
        ctx.close()
        key = f"C{n}" + (f" chunk={ch}" if ch else "")
        res[key] = t2
        print(key, " | ".join(f"{v['kernel'][:4]}/{v['registers']}/{v['phase_level']}: {v['ms']:.4f}" for v in t2["variants"]), flush=True)
if out_json:
    json.dump(res, open(out_json, "w"), indent=1)
