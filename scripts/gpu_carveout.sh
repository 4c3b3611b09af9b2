#!/bin/bash
# L1 against resident problems: cap the hot kernel's blocks per SM and shrink its shared-memory carve-out so that the rest becomes L1
o=gpurun_out; mkdir -p $o
run() { # workload cap pct
  RMPC_MAX_BLOCKS_PER_SM=$2 RMPC_CARVEOUT_PCT=$3 timeout 300 python bench.py --workload $1 --steps 20 --no-workloads --no-cpu-baseline --latency-samples 0 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 cap $2 carve-out $3 %:', round(d['value']), round(d['ms_per_step'],3))"
}
run boxer_radial 9 100; run boxer_radial 5 86; run boxer_radial 5 100; run boxer_radial 4 72
run panda 9 100; run panda 3 86; run panda 3 72; run panda 3 100
run boxer_supermarket 9 100; run boxer_supermarket 5 86; run boxer_supermarket 4 72; run boxer_supermarket 5 100
