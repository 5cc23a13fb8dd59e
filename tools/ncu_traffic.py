"""DRAM bytes per launch of every kernel in an .ncu-rep (ncu --set full) -> profiles/r2_ncu_traffic.json, which bench.py
reads for `roofline.traffic`.  Usage: python tools/ncu_traffic.py rep1.ncu-rep [rep2 ...]"""
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out_path = os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")
try:
    out = json.load(open(out_path))
except (OSError, ValueError):
    out = {}
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
for rep in sys.argv[1:]:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], dict(zip(rows[0], rows[1]))
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        name = re.sub(r"\(.*", "", d["Kernel Name"]).replace("void ", "").replace("<unnamed>::", "").replace(" ", "")
        name = name.replace("(bool)", "").replace("(int)", "")
        rd = float(d["dram__bytes_read.sum"]) * SCALE[units["dram__bytes_read.sum"]]
        wr = float(d["dram__bytes_write.sum"]) * SCALE[units["dram__bytes_write.sum"]]
        rec = out.setdefault(name, {"dram_read": 0.0, "dram_write": 0.0, "launches": 0, "grid": d.get("Grid Size"), "source": os.path.basename(rep)})
        n = rec["launches"]
        rec["dram_read"] = (rec["dram_read"] * n + rd) / (n + 1)
        rec["dram_write"] = (rec["dram_write"] * n + wr) / (n + 1)
        rec["launches"] = n + 1
        rec["time_us_under_ncu"] = float(d["gpu__time_duration.sum"]) * (1e-3 if units["gpu__time_duration.sum"] == "ns" else 1.0)
json.dump(out, open(out_path, "w"), indent=1, sort_keys=True)
for k, v in out.items():
    print(f"{k}: read {v['dram_read'] / 1e6:.2f} MB, write {v['dram_write'] / 1e6:.2f} MB per launch ({v['launches']} launches, {v['source']})")
