#!/bin/bash
# precision 0 (float64) against precision 1 (float32 iterations while mu is large, float64 to finish), same final tolerances
o=gpurun_out; t=${1:-r2k}; mkdir -p $o
timeout 900 python -m pytest tests -m gpu -q -x -s > $o/${t}_pytest.txt 2>&1; grep -E "precision 1 vs 0|passed|failed|Error" $o/${t}_pytest.txt | cut -c1-250
: > $o/${t}_mixed_ab.txt
for w in boxer_radial pointRobot_sweep panda boxer_supermarket; do
  for p in 0 1; do
    timeout 600 python bench.py --workload $w --steps 20 --no-workloads --no-cpu-baseline --latency-samples 0 --precision $p > $o/${t}_${w}_p$p.json 2> $o/${t}_${w}_p$p.err
    python - <<PY | tee -a $o/${t}_mixed_ab.txt
import json
try:
    d=json.loads(open("$o/${t}_${w}_p$p.json").read().strip().splitlines()[-1])
    print("$w precision $p: value %.0f e2e %.0f conv %.4f iters %.2f ms/step %.3f alone %.2f ms launches %d" % (d["value"], d["e2e"]["value"], d["converged_fraction"], d["mean_iters"], d["ms_per_step"], d["roofline"]["one_batch_alone"]["kernel_ms"], d["gpu_launches"]))
except Exception as e:
    print("$w $p failed", e); print(open("$o/${t}_${w}_p$p.err").read()[-1500:])
PY
  done
done
