#!/usr/bin/env python3
"""Dev tool: per-source-line cost of one kernel from an .ncu-rep captured with --import-source on.
Joins the SASS-level source page (per-instruction samples / executed counts) with nvdisasm -g
line info of the same cubin, instruction by instruction.
usage: python tools/ncu_lines.py <rep.ncu-rep> <object.o> <kernel-name-substring> [top_n]"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile


def main():
    rep, obj, kname = sys.argv[1], sys.argv[2], sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    col = {n: hdr.index(n) for n in ("Address", "Source", "# Samples", "Instructions Executed", "stall_no_inst")}
    prof = []
    for r in rows[2:]:
        try:
            prof.append((r[col["Source"]].strip(), int(r[col["# Samples"]]), int(r[col["Instructions Executed"]]), int(r[col["stall_no_inst"]])))
        except (ValueError, IndexError):
            pass
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
    cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.split("\n")
    lines, cur, on = [], ("?", 0), False
    for ln in dis:
        if ln.startswith(".text."):
            on = kname in ln
            continue
        if not on:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
            lines.append(cur)
    n = min(len(lines), len(prof))
    print(f"# {len(prof)} profiled instructions, {len(lines)} disassembled; joined {n}")
    agg = {}
    for i in range(n):
        k = lines[i]
        a = agg.setdefault(k, [0, 0, 0, 0])
        a[0] += prof[i][1]; a[1] += prof[i][2]; a[2] += prof[i][3]; a[3] += 1
    ts = sum(a[0] for a in agg.values()) or 1
    te = sum(a[1] for a in agg.values()) or 1
    src_cache = {}
    def src(k):
        f = os.path.join(os.path.dirname(os.path.abspath(obj)), k[0])
        if f not in src_cache:
            try: src_cache[f] = open(f).read().split("\n")
            except OSError: src_cache[f] = []
        L = src_cache[f]
        return L[k[1] - 1].strip()[:90] if 0 < k[1] <= len(L) else ""
    print(f"{'file:line':28s} {'samp%':>6s} {'exec%':>6s} {'noinst':>7s} {'#sass':>5s}  source")
    for k, a in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
        print(f"{k[0] + ':' + str(k[1]):28s} {100 * a[0] / ts:6.2f} {100 * a[1] / te:6.2f} {a[2]:7d} {a[3]:5d}  {src(k)}")


if __name__ == "__main__":
    main()
