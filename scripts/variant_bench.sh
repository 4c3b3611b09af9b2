#!/bin/bash
# usage: variant_bench.sh <solver dir name> <workload> <batch> "<nvcc -D flags>" ...  : rebuild one solver with flags, bench it
d=$1; w=$2; b=$3; shift 3
cd robot_mpcs_b200/solvers/$d
for flags in "$@"; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --expt-relaxed-constexpr -Xcompiler -fPIC -shared $flags -I . -I ../../csrc -I ../../../include ../../csrc/rmpc_solver.cu -o lib$d.so || exit 1
  (cd ../../.. && python bench.py --steps 5 --warmup 3 --workload $w --batch $b --no-cpu-baseline 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$d [$flags]', round(d['value']), 'solves/s', round(d['ms_per_step'],2), 'ms', d['converged_fraction'], d['mean_iters'])")
done
