#!/usr/bin/env python
"""Turn the ncu artefacts a scripts/gpu_trip_r2*.sh trip left in gpurun_out/ (launch list + the raw-page CSVs of the
--set full captures, converted on the box) into the tracked summaries under profiles/.
usage: scripts/summarize_profile_r2.py <tag> "<command description>" [prof names...]   (e.g. r2a "..." gf gb wg gn at opt loss misc)"""
import collections
import csv
import json
import os
import re
import sys

tag, what = sys.argv[1], sys.argv[2]
csv.field_size_limit(1 << 30)
short = lambda n: re.sub(r"^void ", "", re.sub(r"\(.*", "", n))[:100]
if os.path.exists("gpurun_out/launches.csv"):
    rows = list(csv.reader(l for l in open("gpurun_out/launches.csv") if l.startswith('"')))
    hdr, rows = rows[0], rows[1:]
    ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        agg[short(r[ki])][0] += 1
        agg[short(r[ki])][1] += float(r[vi])
    tot = sum(v[1] for v in agg.values())
    with open(f"profiles/{tag}_launches_summary.txt", "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES)\n")
        f.write(f"# command: {what} ; {len(rows)} launches, {tot / 1e6:.3f} ms total\n")
        f.write(f"# {'ms':>9} {'share':>6} {'count':>6}  kernel\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{v[1] / 1e6:11.3f} {100 * v[1] / tot:5.1f}% {v[0]:6d}  {k}\n")
    with open(f"profiles/{tag}_launches.csv", "w") as f:
        w = csv.writer(f)
        w.writerow(["id", "kernel", "grid", "block", "ns"])
        for r in rows:
            w.writerow([r[0], short(r[ki])[:80], r[gi], r[bi], r[vi]])
    print(open(f"profiles/{tag}_launches_summary.txt").read()[:2500])

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_sector_hit_rate.pct", "lts__t_sectors_op_write.sum",
        "smsp__sass_average_data_bytes_per_sector_mem_global_op_st.ratio", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max", "smsp__cycles_active.avg"]
out_rows = []
for rep in sys.argv[3:]:
    path = f"gpurun_out/prof_{rep}.csv"
    if not os.path.exists(path):
        print("missing", path)
        continue
    rr = list(csv.reader(open(path)))
    h, u = rr[0], rr[1]
    col = {}
    for x in WANT:   # the raw page prefixes some metrics with their section ("SM_A.TriageCompute.")
        idx = [i for i, n in enumerate(h) if n == x or n.endswith("." + x)]
        if idx:
            col[x] = idx[0]
    for r in rr[2:]:
        d = {"report": rep, "kernel": short(r[h.index("Kernel Name")]), "grid": r[h.index("Grid Size")],
             "block": r[h.index("Block Size")]}
        for x, i in col.items():
            d[x] = (r[i] + " " + u[i]).strip()
        out_rows.append(d)
if out_rows:
    keys = ["report", "kernel", "grid", "block"] + [x for x in WANT if any(x in d for d in out_rows)]
    with open(f"profiles/{tag}_ncu_full.csv", "w") as f:
        w = csv.DictWriter(f, fieldnames=keys)
        w.writeheader()
        for d in out_rows:
            w.writerow({k: d.get(k, "") for k in keys})
    print(f"profiles/{tag}_ncu_full.csv: {len(out_rows)} rows")
