#!/bin/bash
# final build of a round: launch list of the headline command + one --set full capture of both solve kernels per workload, summarised on the box
t=${1:-r2f}; o=gpurun_out; mkdir -p $o
bash scripts/gpu_profiles2.sh $t
rm -f $o/prof_${t}_boxer_radial.ncu-rep
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-workloads --latency-samples 0"
timeout 300 $CMD > $o/${t}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/launches_${t}.csv $CMD > $o/${t}_ncu1.log 2>&1
ls -la $o | tail -20
