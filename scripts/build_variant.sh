#!/bin/bash
# usage: build_variant.sh <suffix> "<extra nvcc flags>" : the four bench models into variants/<suffix>/ (for scripts/ab_bench.sh)
v=$1; shift
mkdir -p variants/$v
for m in boxer_n3_01_H30_noSlack pointRobot_n3_005_H20_noSlack panda_n7_005_H20_noSlack boxer_n3_01_H40_noSlack; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -lineinfo -shared -Xcompiler -fPIC $@ \
    -I robot_mpcs_b200/solvers/$m -I robot_mpcs_b200/csrc -I include robot_mpcs_b200/csrc/rmpc_solver.cu -o variants/$v/lib$m.so &
done
wait
