#!/usr/bin/env python
"""Instruction-cache footprint of a kernel from an ncu report: which SASS instructions are ever executed,
how many bytes of 128-byte lines they touch, and which source lines own the hot instructions.
Usage: ncu_hot.py report.ncu-rep [top N lines]"""
import collections
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    cur_file, cur_line, cur_src, idx = None, None, "", {}
    per_line = collections.defaultdict(lambda: [0, 0, 0, ""])  # static, hot, executed, source
    seen = {}
    for r in csv.reader(out.splitlines()):
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            idx = {}
            for i, h in enumerate(r):
                idx.setdefault(h, i)
            continue
        if r[0] == "Function Name" or not idx:
            continue
        if r[0].isdigit():
            cur_line, cur_src = int(r[0]), r[1].strip()
            continue
        if len(r) > 3 and r[2].startswith("0x"):
            addr = int(r[2], 16)
            ex = int(r[idx["Instructions Executed"]])
            if addr in seen:
                continue
            seen[addr] = ex
            e = per_line[(cur_file, cur_line)]
            e[0] += 1
            e[1] += 1 if ex > 0 else 0
            e[2] += ex
            e[3] = cur_src
    if not seen:
        print("no SASS rows found")
        return
    base = min(seen)
    mx = max(seen.values())
    hot = {a for a, e in seen.items() if e > 1e-3 * mx}
    lines = {(a - base) // 128 for a in hot}
    tot = sum(seen.values())
    print(f"static {len(seen)} instr ({len(seen) * 16 / 1024:.1f} KB); ever executed {sum(1 for e in seen.values() if e > 0)}; "
          f"hot (> 0.1 % of max) {len(hot)} instr in {len(lines)} x 128 B lines = {len(lines) * 128 / 1024:.1f} KB")
    byfile = collections.Counter()
    for (f, _), e in per_line.items():
        byfile[f] += e[1]
    print("hot instr by file:", ", ".join(f"{f} {n}" for f, n in byfile.most_common()))
    for (f, ln), e in sorted(per_line.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"{f}:{ln:4d} hot {e[1]:4d} / static {e[0]:4d}  executed {e[2] / tot * 100:5.2f}%  {e[3][:100]}")


if __name__ == "__main__":
    main()
