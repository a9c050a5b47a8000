#!/usr/bin/env python3
"""Dev tool: dynamic instruction mix and hot instruction-cache footprint of one kernel from an
.ncu-rep captured with --import-source on (SASS page: executed count and samples per instruction).
usage: python tools/ncu_mix.py <rep.ncu-rep> > profiles/rNN_<name>_mix.txt"""
import collections
import csv
import io
import re
import subprocess
import sys


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    print("#", rows[0][1] if len(rows[0]) > 1 else rows[0])
    hdr, data = rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    ins = []
    for r in data:
        try:
            ins.append((int(r[ix["Address"]], 16), r[ix["Source"]].strip(), int(r[ix["Instructions Executed"]]), int(r[ix["# Samples"]]),
                        int(r[ix["Thread Instructions Executed"]]), int(r[ix["Predicated-On Thread Instructions Executed"]])))
        except (ValueError, IndexError):
            pass
    tot = sum(i[2] for i in ins)
    samp = sum(i[3] for i in ins)
    thr, pon = sum(i[4] for i in ins), sum(i[5] for i in ins)
    print(f"SASS instructions {len(ins)} ({len(ins) * 16 / 1024:.1f} KB), executed at least once {sum(1 for i in ins if i[2] > 0)}")
    print(f"warp instructions executed {tot}, samples {samp}")
    print(f"active threads per warp instruction {thr / tot:.2f}, predicated-on {pon / tot:.2f}")
    mix, smp = collections.Counter(), collections.Counter()
    for _, src, e, s, _, _ in ins:
        m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+)", src)
        op = m.group(2) if m else src
        mix[op] += e
        smp[op] += s
    print("\n## opcode mix (share of executed warp instructions / of stall samples)")
    for op, e in mix.most_common(32):
        print(f"{op:10s} {100 * e / tot:6.2f} %  {100 * smp[op] / samp:6.2f} %")
    ctrl = sum(mix[o] for o in ("BRA", "BSSY", "BSYNC", "BREAK", "WARPSYNC", "CALL", "RET", "EXIT"))
    print(f"control flow (BRA/BSSY/BSYNC/...) {100 * ctrl / tot:.2f} % of instructions; local-memory (STL+LDL) {100 * (mix['STL'] + mix['LDL']) / tot:.2f} %")
    # hot footprint: 128-byte instruction-cache lines (8 instructions) holding the executed stream
    base = min(i[0] for i in ins)
    line_ex = collections.Counter()
    for a, _, e, _, _, _ in ins:
        line_ex[(a - base) // 128] += e
    order = sorted(line_ex.values(), reverse=True)
    print("\n## instruction-cache footprint (128-B lines)")
    for frac in (0.5, 0.9, 0.99, 0.999):
        acc = 0
        for n, e in enumerate(order, 1):
            acc += e
            if acc >= frac * tot:
                print(f"{100 * frac:5.1f} % of the executed instructions live in {n} lines = {n * 128 / 1024:.1f} KB")
                break
    hot = sorted(k for k, v in line_ex.items() if v > tot * 4e-5)
    print(f"lines with > 0.004 % of the executed stream: {len(hot)} = {len(hot) * 128 / 1024:.1f} KB")
    chunks = collections.Counter(k // 8 for k in hot)
    print("hot lines per KB of code, in address order (0-8):")
    print("".join(str(chunks.get(c, 0)) for c in range(max(line_ex) // 8 + 1)))


if __name__ == "__main__":
    main()
