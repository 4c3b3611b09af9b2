#!/bin/bash
# one --set full capture (source-level) of both solve kernels per workload, summarised on the box (the reports are too big to bring back all)
t=${1:-r2k}; o=gpurun_out; mkdir -p $o
for w in boxer_radial pointRobot_sweep panda boxer_supermarket; do
  CMD="python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline --no-workloads --latency-samples 0 --streams 1"
  timeout 300 $CMD > $o/${t}_plain_$w.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -s 6 -c 2 -f -o $o/prof_${t}_$w $CMD > $o/${t}_ncu_$w.log 2>&1
  python scripts/ncu_summary.py full $o/prof_${t}_$w.ncu-rep "solve_kernel<1" > $o/${t}_ncu_${w}_feasible_kernel.txt 2>&1
  python scripts/ncu_summary.py full $o/prof_${t}_$w.ncu-rep "solve_kernel<2" > $o/${t}_ncu_${w}_shifted_kernel.txt 2>&1
  python scripts/ncu_lines.py $o/prof_${t}_$w.ncu-rep "solve_kernel<(int)1" 40 > $o/${t}_ncu_${w}_lines.txt 2>&1
  [ $w = boxer_radial ] || rm -f $o/prof_${t}_$w.ncu-rep
  wc -l $o/${t}_ncu_${w}_*.txt
done
