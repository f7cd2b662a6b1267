#!/usr/bin/env python
"""Summarise an .ncu-rep here (no GPU needed): key raw metrics + profile split at barriers.
usage: tools/ncu_summary.py gpurun_out/prof.ncu-rep"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2:]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "launch__occupancy_limit_warps",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size"]
for v in vals:
    print("----", v[hdr.index("Kernel Name")][:60])
    for k in keys:
        if k in hdr:
            print(f"  {k:75s} {v[hdr.index(k)]} {units[hdr.index(k)]}")
    st = {h.split("smsp__average_warps_issue_stalled_")[1].split("_per_issue_active")[0]: float(v[i] or 0) for i, h in enumerate(hdr)
          if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")}
    tot = sum(st.values()) or 1
    print("  stalls:", ", ".join(f"{k}={100*x/tot:.0f}%" for k, x in sorted(st.items(), key=lambda t: -t[1])[:8]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
ix = {h: i for i, h in enumerate(hdr)}
def f(r, k):
    try: return float(r[ix[k]])
    except Exception: return 0.0
ts = sum(f(r, "# Samples") for r in data) or 1; ti = sum(f(r, "Instructions Executed") for r in data) or 1
tw = sum(f(r, "L1 Wavefronts Shared") for r in data) or 1
print("shared wavefronts %.0fM, ideal %.0fM" % (tw / 1e6, sum(f(r, "L1 Wavefronts Shared Ideal") for r in data) / 1e6))
cur = [0, 0, 0, 0]; first = 0
for n, r in enumerate(data):
    cur[0] += f(r, "# Samples"); cur[1] += f(r, "Instructions Executed"); cur[2] += f(r, "L1 Wavefronts Shared"); cur[3] += f(r, "stall_barrier")
    if "BAR" in r[ix["Source"]] or n == len(data) - 1:
        print("  sass %4d-%4d: samples %5.1f%% inst %5.1f%% shared-waves %5.1f%% of which barrier-stall samples %5.1f%%" %
              (first, n, 100 * cur[0] / ts, 100 * cur[1] / ti, 100 * cur[2] / tw, 100 * cur[3] / ts))
        cur = [0, 0, 0, 0]; first = n + 1
