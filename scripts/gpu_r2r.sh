#!/bin/bash
o=gpurun_out; t=r2r; mkdir -p $o
python -m pytest tests/test_gpu_features.py -m gpu -q -x -k vel_mode 2>&1 | tail -2
for w in pointRobot boxer_fsd; do timeout 600 python scripts/bench_rollout.py --workload $w > $o/${t}_rollout_$w.json 2> $o/${t}_rollout_$w.err; cut -c1-600 $o/${t}_rollout_$w.json; tail -2 $o/${t}_rollout_$w.err | cut -c1-200; done
# ncu of the infeasible-start kernel on a batch that has infeasible starts (the mixed batch of the GPU tests), and of the mixed-precision kernels
timeout 600 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -c 8 -f -o $o/prof_${t}_shifted python -m pytest tests/test_gpu_features.py -m gpu -q -x -k mixed_batch > $o/${t}_ncu_shifted.log 2>&1
python scripts/ncu_summary.py full $o/prof_${t}_shifted.ncu-rep "solve_kernel<2" > $o/${t}_ncu_shifted_kernel_mixed_batch.txt 2>&1; rm -f $o/prof_${t}_shifted.ncu-rep
CMD="python bench.py --workload pointRobot_sweep --steps 3 --warmup 3 --no-cpu-baseline --no-workloads --latency-samples 0 --streams 1 --precision 1"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"lowp_kernel|solve_kernel" -s 9 -c 3 -f -o $o/prof_${t}_lowp $CMD > $o/${t}_ncu_lowp.log 2>&1
python scripts/ncu_summary.py full $o/prof_${t}_lowp.ncu-rep "lowp_kernel" > $o/${t}_ncu_pointRobot_lowp_kernel.txt 2>&1
python scripts/ncu_summary.py full $o/prof_${t}_lowp.ncu-rep "solve_kernel<3" > $o/${t}_ncu_pointRobot_resume_kernel.txt 2>&1; rm -f $o/prof_${t}_lowp.ncu-rep
wc -l $o/${t}_ncu_*.txt
