"""Dev helper (GPU box): parity of every config at small sizes, both kernel variants."""
import sys, time, json
import numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import oracle
from parity import compare
from opengl_raytracing_b200 import synthetic, abi
from opengl_raytracing_b200.pipeline import RtContext, render_workload, workload_params

ctx = RtContext(0)
print("fp32 peak", ctx.measure_fp32_peak())
oracle.build_port()
for n, (W, H) in [(1, (320, 180)), (2, (320, 180)), (3, (256, 144)), (4, (96, 54)), (5, (256, 144))]:
    w = synthetic.config(n, W, H)
    if n == 4: w.num_samples = 8
    o = oracle.render(w, "port")
    for name, fl in (("parity", abi.FLAG_PARITY_MATH | abi.FLAG_NORMAL_FP32), ("fast", abi.FLAG_NORMAL_FP32), ("fast16", 0)):
        k = render_workload(ctx, w, flags=fl)
        st = compare(k, o, normal_fp16=(fl & abi.FLAG_NORMAL_FP32) == 0)
        print(f"C{n} {name}", json.dumps(st))
    k = render_workload(ctx, w, flags=abi.FLAG_PARITY_MATH | abi.FLAG_COUNTERS)
    kc = ctx.counters()
    diff = {key: (kc[key], o["counters"][key]) for key in kc if kc[key] != o["counters"][key]}
    print(f"C{n} counters diff:", diff)
