#!/bin/bash
t=${1:-r2c}; o=gpurun_out; mkdir -p $o
timeout 600 python -m pytest tests/test_gpu_features.py -m gpu -q -s -x > $o/${t}_pytest.txt 2>&1; tail -12 $o/${t}_pytest.txt | cut -c1-300
run() { # label, env...
  lbl=$1; shift
  env "$@" timeout 600 python bench.py --workload boxer_radial --steps 20 --no-workloads --no-cpu-baseline --latency-samples 0 > $o/${t}_$lbl.json 2> $o/${t}_$lbl.err
  python - <<PY
import json
try:
    d=json.loads(open("$o/${t}_$lbl.json").read().strip().splitlines()[-1])
    print("$lbl", "value %.0f e2e %.0f ms/step %.2f alone_ms %.2f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["one_batch_alone"]["kernel_ms"]))
except Exception as e:
    print("$lbl failed", e); print(open("$o/${t}_$lbl.err").read()[-1500:])
PY
}
run base A=1
run max7 RMPC_MAX_BLOCKS_PER_SM=7
run max6 RMPC_MAX_BLOCKS_PER_SM=6
CMD="python bench.py --workload boxer_radial --steps 3 --warmup 3 --no-cpu-baseline --no-workloads --latency-samples 0 --streams 1"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -s 12 -c 2 -f -o $o/prof_${t}_boxer_radial $CMD > $o/${t}_ncu.log 2>&1
ls -la $o/prof_${t}_boxer_radial.ncu-rep
