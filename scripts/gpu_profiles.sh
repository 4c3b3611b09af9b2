#!/bin/bash
# ncu: launch list of the headline command + one --set full capture (source-level) of the hot kernel of every workload + FSD
t=${1:-r2k}; o=gpurun_out; mkdir -p $o
for w in boxer_radial pointRobot_sweep panda boxer_supermarket; do
  CMD="python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline --no-workloads --latency-samples 0 --streams 1"
  timeout 300 $CMD > $o/${t}_plain_$w.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -s 6 -c 2 -f -o $o/prof_${t}_$w $CMD > $o/${t}_ncu_$w.log 2>&1
  ls -la $o/prof_${t}_$w.ncu-rep
done
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-workloads --latency-samples 0"
timeout 300 $CMD > $o/${t}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/launches_${t}.csv $CMD > $o/${t}_ncu1.log 2>&1
timeout 300 python scripts/bench_fsd.py > $o/${t}_bench_fsd.json 2>$o/${t}_fsd.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fsd_kernel -s 3 -c 1 -f -o $o/prof_${t}_fsd python scripts/bench_fsd.py --steps 3 > $o/${t}_ncu_fsd.log 2>&1
# the infeasible-start kernel on a batch that has such starts (the mixed batch of the GPU tests)
ls -la $o | tail -12
