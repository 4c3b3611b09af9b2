#!/bin/bash
# round-2 first GPU call: tests, headline bench, ncu launch list + full captures (source-level) of every workload's kernel
t=r2a; o=gpurun_out; mkdir -p $o
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > $o/${t}_smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $o/${t}_pytest.txt 2>&1; tail -3 $o/${t}_pytest.txt
timeout 600 python bench.py --steps 20 > $o/${t}_bench_boxer_radial.json 2> $o/${t}_bench.err; cut -c1-300 $o/${t}_bench_boxer_radial.json
python scripts/basin_agreement.py --gpu --json $o/${t}_basin_gpu.json > $o/${t}_basin_gpu.txt 2>&1; grep -v "other basin" $o/${t}_basin_gpu.txt
for w in boxer_radial pointRobot_sweep panda boxer_supermarket; do
  CMD="python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline --latency-samples 0 --streams 1"
  timeout 300 $CMD > $o/${t}_plain_$w.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -s 12 -c 2 -f -o $o/prof_${t}_$w $CMD > $o/${t}_ncu_$w.log 2>&1
  ls -la $o/prof_${t}_$w.ncu-rep
done
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --latency-samples 0"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/launches_${t}.csv $CMD > $o/${t}_ncu1.log 2>&1
timeout 300 python scripts/bench_fsd.py > $o/${t}_bench_fsd.json 2>$o/${t}_fsd.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fsd_kernel -s 3 -c 1 -f -o $o/prof_${t}_fsd python scripts/bench_fsd.py --steps 3 > $o/${t}_ncu_fsd.log 2>&1
ls -la $o | tail -20
