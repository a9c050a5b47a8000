"""Dev helper: registers / stack / spills per kernel from the nvcc logs of the last build
(`python tools/ptxas_summary.py [variant]`)."""
import re, subprocess, sys, os
here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "opengl_raytracing_b200", "csrc")
variant = sys.argv[1] if len(sys.argv) > 1 else ""
for f in ("rt_kernel_fast", "rt_kernel_parity", "rt_regroup_fast_small", "rt_regroup_fast_big", "rt_regroup_parity_small", "rt_regroup_parity_big", "rt_post", "rt_multi"):
    path = os.path.join(here, f + variant + ".o.log")
    if not os.path.isfile(path):
        continue
    t = open(path).read()
    print(f + variant)
    for m in re.finditer(r"Compiling entry function '([^']+)' for 'sm_100a'\n(.*?)Used (\d+) registers", t, re.S):
        name, body, regs = m.group(1), m.group(2), m.group(3)
        ms = re.findall(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", body)
        stack, ss, sl = ms[-1] if ms else ("?", "?", "?")
        d = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        d = re.sub(r"\(.*", "", d).replace("rt::fast::", "").replace("rt::parity::", "").replace("rt::fast_rg_small::", "").replace("rt::fast_rg_big::", "").replace("rt::parity_rg_small::", "").replace("rt::parity_rg_big::", "").replace("void ", "")
        print(f"  {d:58s} regs {regs:>3} stack {stack:>4} spill st {ss:>4} ld {sl:>4}")
