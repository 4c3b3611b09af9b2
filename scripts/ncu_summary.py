#!/usr/bin/env python
"""Summarise ncu outputs into profiles/ (run here, no GPU needed).

  python scripts/ncu_summary.py launches gpurun_out/launches_X.csv            -> per-kernel time shares
  python scripts/ncu_summary.py full gpurun_out/prof_X.ncu-rep [kernel-regex] -> key metrics of the captured kernel
"""
import collections
import csv
import io
import re
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__occupancy_limit",
    "launch__waves_per_multiprocessor", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.per_cycle_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum ",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64", "sm__pipe_fp64_cycles_active",
    "sm__inst_executed_pipe_fma", "sm__pipe_fma_cycles_active", "sm__inst_executed_pipe_fmaheavy", "sm__inst_executed_pipe_alu",
    "sm__inst_executed_pipe_lsu", "sm__inst_executed_pipe_xu",
    "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ", "lts__t_bytes.sum ", "smsp__average_warp",
    "smsp__average_warps_issue_stalled", "derived__smsp__sass_thread_inst_executed_op_", "sm__cycles_active.avg ",
    "smsp__cycles_active.avg ", "l1tex__t_sectors_pipe_lsu_mem_local", "smsp__inst_executed_op_shared",
]


def launches(path):
    text = "\n".join(l for l in open(path) if l.startswith('"'))
    rows = list(csv.DictReader(io.StringIO(text)))
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        try:
            v = float(r["Metric Value"].replace(",", ""))
        except (ValueError, KeyError):
            continue
        if r.get("Metric Unit") == "us":
            v *= 1e3
        elif r.get("Metric Unit") == "ms":
            v *= 1e6
        k = r["Kernel Name"].split("(")[0]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values()) or 1.0
    print(f"# launch list {path}: {len(rows)} launches, total {tot / 1e6:.3f} ms (cold-cache, serialised: compare shares)")
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k:80s} n={v[0]:4d} total_ms={v[1] / 1e6:10.3f} mean_us={v[1] / v[0] / 1e3:10.1f} share={v[1] / tot:.4f}")


def full(path, pat=None):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units = rows[0], rows[1]
    ki = head.index("Kernel Name")
    for row in rows[2:]:
        if pat and pat not in row[ki]:
            continue
        print(f"# {path}: {row[ki][:120]}")
        for i, k in enumerate(head):
            if any(k.startswith(x.strip()) if x.endswith(" ") else k.startswith(x) for x in KEYS) or ("warp_issue_stalled" in k and k.endswith("_ratio")) \
                    or k.startswith("smsp__pcsamp_warps_issue_stalled") :
                if re.search(r"\.(min|max|sum)\.pct_of_peak", k) or ".peak_sustained" in k:
                    continue
                if row[i] not in ("", "0", "0.000000", "no data"):
                    print(f"{k:95s} {units[i]:14s} {row[i]}")
        print()


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        full(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
