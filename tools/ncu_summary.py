#!/usr/bin/env python
"""`ncu --set full` report -> the committed summaries under profiles/:

    tools/ncu_summary.py gpurun_out/prof_full.ncu-rep BATCH profiles/r1_vNN_ncu_full_batchB.csv [--traffic]

Writes one CSV (metric, unit, one column per captured kernel) with the figures DESIGN.md / profiles/README.md quote
(duration, DRAM bytes, instruction count, issue / pipe utilisation, L1 / L2 hit rates, occupancy limiters and the
share of every warp-stall reason) and, with --traffic, profiles/ncu_traffic.json (DRAM bytes per launch of the three
matcher kernels, which bench.py reports as roofline.traffic). Needs `ncu` on PATH (reading a report needs no GPU)."""
import csv
import io
import json
import subprocess
import sys

rep, batch, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, units, data = rows[0], rows[1], rows[2:]
ki = h.index("Kernel Name")
names = [r[ki].split("(")[0] for r in data]
keep = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes.sum.per_second",
        "smsp__inst_executed.sum", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
        "launch__grid_size", "launch__block_size", "launch__cluster_size")
lines = [["metric", "unit"] + names]
col = {c: i for i, c in enumerate(h)}
for m in keep:
    if m in col:
        lines.append([m, units[col[m]]] + [r[col[m]] for r in data])
stall = [(c, i) for c, i in col.items() if c.startswith("smsp__pcsamp_warps_issue_stalled_") and not c.endswith("_not_issued")]
tot = [sum(float(r[i].replace(",", "") or 0) for _, i in stall) for r in data]
for c, i in sorted(stall, key=lambda t: -sum(float(r[t[1]].replace(",", "") or 0) for r in data)):
    lines.append(["stall share: " + c.replace("smsp__pcsamp_warps_issue_stalled_", ""), "%"] +
                 ["%.1f" % (100.0 * float(r[i].replace(",", "") or 0) / t) if t else "" for r, t in zip(data, tot)])
csv.writer(open(out, "w", newline="")).writerows(lines)
print(open(out).read())

if "--traffic" in sys.argv:
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0}
    def val(r, m):
        return float(r[col[m]].replace(",", "")) * scale.get(units[col[m]], 1.0)
    res = {}
    for st, key in (("cost", "k_cost"), ("paths", "k_vert"), ("wta", "k_rows")):
        rs = [r for r in data if key in r[ki]]
        if not rs:
            continue
        n = len(rs)
        rd = sum(val(r, "dram__bytes_read.sum") for r in rs) / n
        wr = sum(val(r, "dram__bytes_write.sum") for r in rs) / n
        res[st] = {"kernel": rs[0][ki].split("(")[0].replace("void ", ""), "batch": batch, "launches_captured": n,
                   "dram_bytes_read_per_launch": rd, "dram_bytes_write_per_launch": wr, "dram_bytes_per_launch": rd + wr,
                   "ms_under_ncu": sum(val(r, "gpu__time_duration.sum") for r in rs) / n,
                   "warp_instructions": sum(float(r[col["smsp__inst_executed.sum"]].replace(",", "")) for r in rs) / n,
                   "source": "%s (ncu --set full), summarised in %s" % (rep.split("/")[-1], out)}
    json.dump(res, open("profiles/ncu_traffic.json", "w"), indent=1)
    print(json.dumps(res, indent=1))
