#!/bin/bash
# instruction-cache counters of the hot kernel at 6 / 5 / 4 resident problems per SM (boxer N=30): is the instruction stream the bound?
o=gpurun_out; mkdir -p $o
M="gcc__cache_requests_type_instruction.sum,gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed,gcc__cache_requests_type_instruction_lookup_hit.sum,gcc__cache_requests_type_instruction_lookup_miss.sum,gcc__gcc2xbar_requests_type_instruction.sum,gcc__cycles_active.avg,gcc__cycles_elapsed.avg,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,sm__warps_active.avg.per_cycle_active,gpu__time_duration.sum,sm__cycles_active.avg"
for cap in 9 5 4; do
  CMD="python bench.py --workload boxer_radial --steps 3 --warmup 3 --no-cpu-baseline --no-workloads --latency-samples 0 --streams 1"
  RMPC_MAX_BLOCKS_PER_SM=$cap timeout 600 ncu --metrics $M --clock-control none -k regex:solve_kernel -s 6 -c 2 --csv --log-file $o/icache_cap$cap.csv $CMD > /dev/null 2>&1
  echo "== cap $cap"; python - <<PY
import csv
for r in csv.reader(open("$o/icache_cap$cap.csv")):
    if len(r) > 10 and r[0].isdigit() and "<1>" in r[4]: print("%-85s %s %s" % (r[-3], r[-2], r[-1]))
PY
done
