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
# samples per pixel of the captured launch when it was not the configuration's own (AB_NS=... for a short replay):
# bench.py scales the bytes to the full frame (DRAM traffic of this kernel is proportional to the paths it runs)
captured_spp = int(sys.argv[4]) if len(sys.argv) > 4 else None
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
       "report": os.path.basename(rep), "captured_spp": captured_spp}
path = os.path.join(ROOT, "profiles", "traffic.json")
data = {}
if os.path.exists(path):
    data = json.load(open(path))
data[config] = rec
json.dump(data, open(path, "w"), indent=1, sort_keys=True)
print(config, rec)
