#!/usr/bin/env python3
"""Dev tool: executed-work figures of one frame from an .ncu-rep (ncu --set full) -> an entry of profiles/executed.json.
Sums over the kernels of the capture whose name matches (all launches of ONE frame must be in the capture).
usage: python tools/ncu_executed.py report.ncu-rep 'C3@3840x2160' 'profiles/r02_..txt' [name-regex]"""
import csv, io, json, os, re, subprocess, sys

rep, key, source = sys.argv[1], sys.argv[2], sys.argv[3]
pat = re.compile(sys.argv[4] if len(sys.argv) > 4 else r"rt_packet_kernel|rt_regroup_kernel|wf_")
out = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[0]
# ncu's raw page reports the FP32 opcode counters as per-cycle rates: count = <op>.sum.per_cycle_elapsed x smsp__cycles_elapsed.avg
NAMES = ("Kernel Name", "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum.per_cycle_elapsed", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum.per_cycle_elapsed",
         "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum.per_cycle_elapsed", "smsp__cycles_elapsed.avg", "smsp__inst_executed.sum",
         "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum")
col = {k: hdr.index(k) for k in NAMES}
units = rows[1]
def num(r, k):
    v = float(r[col[k]].replace(",", "") or 0)
    u = units[col[k]]
    return v * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3}.get(u, 1.0)
acc = {"fadd": 0.0, "fmul": 0.0, "ffma": 0.0, "warp_inst": 0.0, "dram_read": 0.0, "dram_write": 0.0, "ms_under_ncu": 0.0}
kernels = {}
for r in rows[2:]:
    name = r[col["Kernel Name"]]
    if not pat.search(name):
        continue
    kernels[name] = kernels.get(name, 0) + 1
    cyc = num(r, "smsp__cycles_elapsed.avg")
    acc["fadd"] += num(r, "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum.per_cycle_elapsed") * cyc
    acc["fmul"] += num(r, "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum.per_cycle_elapsed") * cyc
    acc["ffma"] += num(r, "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum.per_cycle_elapsed") * cyc
    acc["warp_inst"] += num(r, "smsp__inst_executed.sum")
    acc["dram_read"] += num(r, "dram__bytes_read.sum")
    acc["dram_write"] += num(r, "dram__bytes_write.sum")
    acc["ms_under_ncu"] += num(r, "gpu__time_duration.sum")
entry = dict(acc, source=source, kernels=kernels)
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles", "executed.json")
doc = json.load(open(path)) if os.path.isfile(path) else {"_doc": "executed work of one frame per config, from the committed ncu --set full captures of this build (tools/ncu_executed.py); bench.py's roofline.executed reads it"}
doc[key] = entry
json.dump(doc, open(path, "w"), indent=1)
print(key, json.dumps(entry))
