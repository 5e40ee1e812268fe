#!/usr/bin/env python
"""Turn the ncu artefacts a gpu_profile.sh trip left in gpurun_out/ into the tracked summaries under
profiles/: per-kernel totals of the launch list and the headline counters of the --set full captures.
usage: scripts/summarize_profile.py <tag>      (e.g. r1a)"""
import collections
import csv
import re
import subprocess
import sys

tag = sys.argv[1]
rows = list(csv.reader(l for l in open("gpurun_out/launches.csv") if l.startswith('"')))
hdr, rows = rows[0], rows[1:]
ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    name = re.sub(r"^void ", "", re.sub(r"\(.*", "", r[ki]))[:100]
    agg[name][0] += 1
    agg[name][1] += float(r[vi])
tot = sum(v[1] for v in agg.values())
with open(f"profiles/{tag}_launches_summary.txt", "w") as f:
    f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES)\n")
    f.write(f"# command: scripts/gpu_round.sh list (one 256-image C4 micro-batch step, warm-up + 1 step) ; {len(rows)} launches, {tot / 1e6:.3f} ms total\n")
    f.write(f"# {'ms':>9} {'share':>6} {'count':>6}  kernel\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{v[1] / 1e6:11.3f} {100 * v[1] / tot:5.1f}% {v[0]:6d}  {k}\n")
with open(f"profiles/{tag}_launches.csv", "w") as f:
    w = csv.writer(f)
    w.writerow(["id", "kernel", "grid", "block", "ns"])
    for r in rows:
        w.writerow([r[0], re.sub(r"^void ", "", re.sub(r"\(.*", "", r[ki]))[:80], r[gi], r[bi], r[vi]])

WANT = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max"]
with open(f"profiles/{tag}_ncu_full.csv", "w") as f:
    w = csv.writer(f)
    w.writerow(["report"] + WANT)
    for rep in sys.argv[2:]:
        out = subprocess.run(["ncu", "-i", f"gpurun_out/{rep}.ncu-rep", "--page", "raw", "--csv"], capture_output=True,
                             text=True).stdout
        rr = list(csv.reader(out.splitlines()))
        h, u = rr[0], rr[1]
        for r in rr[2:]:
            w.writerow([rep] + [(r[h.index(x)] + " " + u[h.index(x)]).strip() if x in h else "" for x in WANT])
print(open(f"profiles/{tag}_launches_summary.txt").read()[:1500])
