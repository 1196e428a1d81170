#!/usr/bin/env python
"""ncu metrics CSV (--metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,... --csv) ->
profiles/ncu_traffic.json: DRAM bytes per launch of the three matcher kernels, keyed by bench.py's stage name.
usage: tools/ncu_traffic.py gpurun_out/traffic.csv BATCH [out.json]"""
import csv
import json
import sys

src, batch = sys.argv[1], int(sys.argv[2])
out = sys.argv[3] if len(sys.argv) > 3 else "profiles/ncu_traffic.json"
rows = list(csv.reader(l for l in open(src) if not l.startswith("==")))
h = rows[0]
ki, mi, vi, ui = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "inst": 1.0}
stage_of = {"k_cost": "cost", "k_vert": "paths", "k_rows": "wta"}
acc = {}
for r in rows[1:]:
    name = r[ki]
    st = next((s for k, s in stage_of.items() if k in name), None)
    if st is None:
        continue
    d = acc.setdefault(st, {"kernel": name.split("(")[0].replace("void ", ""), "batch": batch, "n": {}, "sum": {}})
    v = float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0)
    d["sum"][r[mi]] = d["sum"].get(r[mi], 0.0) + v
    d["n"][r[mi]] = d["n"].get(r[mi], 0) + 1
res = {}
for st, d in acc.items():
    avg = {m: d["sum"][m] / d["n"][m] for m in d["sum"]}
    res[st] = {"kernel": d["kernel"], "batch": batch, "launches_captured": max(d["n"].values()),
               "dram_bytes_read_per_launch": avg.get("dram__bytes_read.sum"),
               "dram_bytes_write_per_launch": avg.get("dram__bytes_write.sum"),
               "dram_bytes_per_launch": (avg.get("dram__bytes_read.sum", 0.0) + avg.get("dram__bytes_write.sum", 0.0)),
               "ms_under_ncu": avg.get("gpu__time_duration.sum"),
               "warp_instructions": avg.get("smsp__inst_executed.sum"), "source": src.split("/")[-1]}
json.dump(res, open(out, "w"), indent=1)
print(json.dumps(res, indent=1))
