"""Summarise an .ncu-rep: headline metrics, dynamic instruction mix, top stall sites (development aid)."""
import collections
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
iters = float(sys.argv[2]) if len(sys.argv) > 2 else None     # warp-iterations, to normalise the mix
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warp_latency_per_inst_issued.ratio", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]
for r in rows[2:]:
    print("kernel:", r[hdr.index("Kernel Name")][:80])
    for k in keys:
        if k in hdr:
            print(f"  {k:70s} {r[hdr.index(k)]} {rows[1][hdr.index(k)]}")
    for i, h in enumerate(hdr):
        if "issue_stalled" in h and h.endswith("per_issue_active.ratio"):
            try:
                v = float(r[i])
            except ValueError:
                continue
            if v > 0.2:
                print(f"  stall {h.split('issue_stalled_')[1].split('_per_issue')[0]:30s} {v:.2f}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr, data = rows[hi], rows[hi + 1:]
isrc, iex, ismp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
tot, smp = collections.Counter(), collections.Counter()
for r in data:
    if len(r) <= ismp or not r[iex].isdigit():
        continue
    op = re.sub(r"^@!?U?P\d+\s+", "", r[isrc].strip()).split()[0].split(".")[0]
    tot[op] += int(r[iex]); smp[op] += int(r[ismp])
total = sum(tot.values()); ts = sum(smp.values())
print("total warp instr", total, "samples", ts, ("per warp-iter %.1f" % (total / iters)) if iters else "")
for op, c in tot.most_common(18):
    print(f"  {op:10s} {c / (iters or 1):12.2f}  samples {100 * smp[op] / ts:5.1f}%")
print("top stall sites:")
for r in sorted([r for r in data if len(r) > ismp and r[ismp].isdigit()], key=lambda r: -int(r[ismp]))[:14]:
    print(f"  {100 * int(r[ismp]) / ts:5.1f}%  x{r[iex]:>9s}  {r[isrc].strip()[:90]}")
