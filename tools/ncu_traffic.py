#!/usr/bin/env python
"""Record the measured DRAM traffic of a kernel from an ncu --set full report into profiles/traffic.json
(read by bench.py for roofline.traffic).   tools/ncu_traffic.py <config> <kernel name> <report.ncu-rep>"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
config, kernel, rep = sys.argv[1:4]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
head, units, vals = rows[0], rows[1], rows[2]


def get(name):
    i = head.index(name)
    v = float(vals[i].replace(",", ""))
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[units[i]]
    return v * scale


rec = {"kernel": kernel, "dram_bytes_per_launch": int(get("dram__bytes_read.sum") + get("dram__bytes_write.sum")),
       "dram_read": int(get("dram__bytes_read.sum")), "dram_write": int(get("dram__bytes_write.sum")),
       "report": os.path.basename(rep)}
path = os.path.join(ROOT, "profiles", "traffic.json")
data = {}
if os.path.exists(path):
    data = json.load(open(path))
data[config] = rec
json.dump(data, open(path, "w"), indent=1, sort_keys=True)
print(config, rec)
