#!/bin/bash
# usage: gpu_ab.sh <tag> <workload> <variant dir name or "base"> ...   : bench one workload with the default build and with variants/<name>
o=gpurun_out; t=$1; w=$2; shift 2; mkdir -p $o
for v in "$@"; do
  if [ $v = base ]; then unset RMPC_VARIANT_DIR; else export RMPC_VARIANT_DIR=$PWD/variants/$v; fi
  timeout 600 python bench.py --workload $w --steps 20 --no-workloads --no-cpu-baseline --latency-samples 300 > $o/${t}_${w}_$v.json 2> $o/${t}_${w}_$v.err
  python - <<PY
import json
try:
    d=json.loads(open("$o/${t}_${w}_$v.json").read().strip().splitlines()[-1])
    l=d.get("single_solve_latency",{})
    print("$w [$v]: value %.0f e2e %.0f conv %.4f iters %.2f ms/step %.3f alone %.2f ms  B=1 p50 %.3f p99 %.3f" % (d["value"], d["e2e"]["value"], d["converged_fraction"], d["mean_iters"], d["ms_per_step"], d["roofline"]["one_batch_alone"]["kernel_ms"], l.get("p50",0), l.get("p99",0)))
except Exception as e:
    print("$w $v failed", e); print(open("$o/${t}_${w}_$v.err").read()[-1200:])
PY
done
