#!/usr/bin/env python3
"""Dev tool: attributes a profiled kernel's samples / executed instructions to the major device
functions of rt_packet.cuh / rt_kernel.cuh through nvdisasm's inline call stacks.
usage: python tools/ncu_funcs.py <rep.ncu-rep> <object.o> <kernel-substring> <header.cuh>"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

MAJOR = ["scan_candidates", "probe_fan", "cone_misses_sphere", "cone_misses_plane", "grid_cell", "wf_shadow_kernel", "wf_bounce_kernel",
         "rt_regroup_kernel", "wf_light_vector", "wf_finish_pixel", "wf_stage_scene", "cosine_hemisphere", "refraction_dir", "random2",
         "trace_any_packet", "trace_one", "trace_any_one", "trace_closest_generic", "trace_closest", "shadow_packets",
         "compute_lighting", "shade_pixel", "rt_packet_kernel", "subsurface", "compute_pbr", "fetch_sky", "fetch_noise",
         "trace_any", "pcf_shadow", "pcss_shadow", "rt_render_kernel", "load_light", "load_material"]


def func_ranges(path):
    out, cur = [], None
    for i, ln in enumerate(open(path).read().split("\n"), 1):
        m = re.match(r"^(?:template\s*<[^>]*>\s*)?(?:__global__|__device__|inline).*?\b([A-Za-z_][A-Za-z0-9_]*)\s*\(", ln)
        if m and not ln.startswith(" "):
            if cur: out.append((cur[0], cur[1], i - 1))
            cur = (m.group(1), i)
    if cur: out.append((cur[0], cur[1], 10 ** 9))
    return out


def main():
    rep, obj, kname, header = sys.argv[1:5]
    ranges = {os.path.basename(header): func_ranges(header)}
    mh = os.path.join(os.path.dirname(header), "rt_math.cuh")
    if os.path.isfile(mh): ranges["rt_math.cuh"] = func_ranges(mh)
    for extra in ("rt_packet.cuh", "rt_wavefront.cuh", "rt_regroup.cuh", "rt_kernel.cuh"):
        eh = os.path.join(os.path.dirname(header), extra)
        if os.path.isfile(eh) and extra not in ranges: ranges[extra] = func_ranges(eh)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    names = ["# Samples", "Instructions Executed", "stall_no_inst", "stall_barrier", "stall_wait", "stall_short_sb", "stall_long_sb", "stall_math"]
    col = [hdr.index(n) for n in names]
    prof = []
    for r in rows[2:]:
        try: prof.append([int(r[c]) for c in col])
        except (ValueError, IndexError): pass
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
    cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    dis = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.split("\n")
    stacks, cur, on, fresh = [], [], False, True
    for ln in dis:
        if ln.startswith(".text."):
            on = kname in ln; continue
        if not on: continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m:
            if fresh: cur = []; fresh = False
            cur.append((os.path.basename(m.group(1)), int(m.group(2))))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
            stacks.append(list(cur)); fresh = True
    def fn_of(frame):
        for name, a, b in ranges.get(frame[0], []):
            if a <= frame[1] <= b: return name
        return None
    agg = {}
    n = min(len(stacks), len(prof))
    for i in range(n):
        f = "?"
        for fr in stacks[i]:                 # innermost first
            g = fn_of(fr)
            if g in MAJOR: f = g; break
        a = agg.setdefault(f, [0] * (len(names) + 1))
        for j, v in enumerate(prof[i]): a[j] += v
        a[-1] += 1
    tot = [sum(a[j] for a in agg.values()) or 1 for j in range(len(names))]
    print(f"{'function':24s} {'samp%':>6s} {'exec%':>6s} {'noinst%':>7s} {'barr%':>6s} {'wait%':>6s} {'ssb%':>6s} {'lsb%':>6s} {'math%':>6s} {'#sass':>6s}")
    for f, a in sorted(agg.items(), key=lambda x: -x[1][0]):
        print(f"{f:24s} " + " ".join(f"{100 * a[j] / tot[j]:6.2f}" for j in range(len(names))) + f" {a[-1]:6d}")
    print("total samples", tot[0], "executed warp-inst", tot[1])
    # the unattributed part ("?": kernel bodies, small helpers) by the source line of its OUTERMOST frame inside our headers,
    # so that a whole inlined helper chain counts for the kernel-body line that called it
    lines = {}
    for i in range(n):
        f = "?"
        for fr in stacks[i]:
            g = fn_of(fr)
            if g in MAJOR: f = g; break
        if f != "?": continue
        key = None
        for fr in reversed(stacks[i]):       # outermost first
            if fr[0] in ranges: key = fr; break
        if key is None and stacks[i]: key = stacks[i][-1]
        a = lines.setdefault(key, [0] * (len(names) + 1))
        for j, v in enumerate(prof[i]): a[j] += v
        a[-1] += 1
    top = int(os.environ.get("NCU_FUNCS_LINES", "60"))
    print(f"-- unattributed code by source line (outermost frame), top {top} by samples")
    for key, a in sorted(lines.items(), key=lambda x: -x[1][0])[:top]:
        k = f"{key[0]}:{key[1]}" if key else "(no line info)"
        print(f"{k:28s} " + " ".join(f"{100 * a[j] / tot[j]:6.2f}" for j in range(len(names))) + f" {a[-1]:6d}")


if __name__ == "__main__":
    main()
