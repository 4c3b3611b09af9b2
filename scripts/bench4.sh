#!/bin/bash
# the four workloads, device-resident numbers only (quick A/B while optimising)
for w in boxer_radial pointRobot_sweep panda boxer_supermarket; do
  python bench.py --steps ${STEPS:-5} --warmup 3 --workload $w --no-cpu-baseline $BENCH_ARGS 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$w', round(d['value']), 'solves/s', round(d['ms_per_step'],2), 'ms conv', round(d['converged_fraction'],4), 'iters', round(d['mean_iters'],2), 'e2e', round(d['e2e']['value']))"
done
