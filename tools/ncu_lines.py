#!/usr/bin/env python
"""Aggregate an ncu report's source page (cuda,sass view) per source line: executed warp
instructions, stall samples and the dominant stall reasons.  Usage: ncu_lines.py report.ncu-rep [N]"""
import collections
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    cur, idx = None, {}
    agg = collections.defaultdict(lambda: collections.Counter())
    src = {}
    stall_cols = []
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            continue
        if r[0] == "Line No":
            idx = {}
            for i, h in enumerate(r):
                idx.setdefault(h, i)
            stall_cols = [h for h in r if h.startswith("stall_") and "Not Issued" not in h]
            continue
        if len(r) > 2 and r[2] == "-":
            try:
                ln = int(r[0])
            except ValueError:
                continue
            key = (cur, ln)
            src[key] = r[1]
            a = agg[key]
            a["instr"] += int(r[idx["Instructions Executed"]])
            a["thr"] += int(r[idx["Thread Instructions Executed"]])
            a["samples"] += int(r[idx["# Samples"]])
            for h in stall_cols:
                a[h] += int(r[idx[h]])
    ti = sum(v["instr"] for v in agg.values())
    ts = sum(v["samples"] for v in agg.values())
    tt = sum(v["thr"] for v in agg.values())
    print(f"warp instr {ti}  thread instr {tt}  avg active {tt / max(ti, 1):.1f}  samples {ts}")
    tot = collections.Counter()
    for v in agg.values():
        for h in stall_cols:
            tot[h] += v[h]
    print("stalls:", ", ".join(f"{k[6:]} {v / max(ts, 1) * 100:.1f}%" for k, v in tot.most_common(8)))
    byfile = collections.Counter()
    for (f, _), v in agg.items():
        byfile[f] += v["samples"]
    print("samples by file:", ", ".join(f"{f} {v / ts * 100:.1f}%" for f, v in byfile.most_common()))
    # where the lanes are lost: warp instructions and samples bucketed by the line's mean active lanes
    buckets = [(0, 6), (6, 12), (12, 20), (20, 33)]
    for lo, hi in buckets:
        sel = [v for v in agg.values() if v["instr"] and lo <= v["thr"] / v["instr"] < hi]
        print(f"lines with {lo:2d} <= active < {hi:2d}: instr {sum(v['instr'] for v in sel) / ti * 100:5.1f}%  "
              f"samples {sum(v['samples'] for v in sel) / ts * 100:5.1f}%")
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1]["samples"])[:top]:
        st = ", ".join(f"{h[6:]} {v[h]}" for h in sorted(stall_cols, key=lambda h: -v[h])[:2])
        act = v["thr"] / max(v["instr"], 1)
        print(f"{key[0]}:{key[1]:4d} samp {v['samples'] / ts * 100:5.1f}% instr {v['instr'] / ti * 100:5.1f}% act {act:4.1f} [{st}]  {src[key].strip()[:90]}")


if __name__ == "__main__":
    main()
