#!/bin/bash
# usage (on the GPU box): profile_round.sh <tag>   -> gpurun_out/<tag>_*: the round's measurement set
# bench lines first (never under ncu), then the ncu launch list of the same command, then one --set full capture.
t=$1; o=gpurun_out; mkdir -p $o
python bench.py > $o/${t}_bench_boxer_radial.json 2> $o/${t}_bench_boxer_radial.err || exit 1
tail -1 $o/${t}_bench_boxer_radial.json | cut -c1-400
for w in pointRobot_sweep panda boxer_supermarket; do
  python bench.py --workload $w > $o/${t}_bench_$w.json 2> $o/${t}_bench_$w.err; tail -1 $o/${t}_bench_$w.json | cut -c1-200
done
python bench.py --impl reference > $o/${t}_bench_reference_arm.json 2> $o/${t}_ref.err; tail -1 $o/${t}_bench_reference_arm.json | cut -c1-300
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --latency-samples 0"
$CMD > $o/${t}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/launches_${t}.csv $CMD > $o/${t}_ncu1.log 2>&1
$CMD > $o/${t}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:solve_kernel -s 6 -c 2 -o $o/prof_${t}_boxer $CMD > $o/${t}_ncu2.log 2>&1
ls -la $o | tail -15
