#!/bin/bash
o=gpurun_out; mkdir -p $o
run() { # workload cap pct
  RMPC_MAX_BLOCKS_PER_SM=$2 RMPC_CARVEOUT_PCT=$3 timeout 300 python bench.py --workload $1 --steps 40 --no-workloads --no-cpu-baseline --latency-samples 0 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 cap $2 carve-out $3 %:', round(d['value']), round(d['ms_per_step'],3), 'alone', round(d['roofline']['one_batch_alone']['kernel_ms'],3))"
}
run panda 9 100; run panda 3 72; run panda 2 58; run panda 2 44; run panda 9 100; run panda 3 72
run pointRobot_sweep 9 100; run pointRobot_sweep 7 86; run pointRobot_sweep 6 72
run boxer_radial 9 100; run boxer_radial 5 86; run boxer_radial 9 100; run boxer_radial 5 86
