#!/bin/bash
t=${1:-r2b}; o=gpurun_out; mkdir -p $o
timeout 1800 python -m pytest tests -m gpu -q -s > $o/${t}_pytest.txt 2>&1; tail -25 $o/${t}_pytest.txt | cut -c1-300
for w in boxer_radial pointRobot_sweep panda boxer_supermarket; do
  timeout 600 python bench.py --workload $w --steps 20 > $o/${t}_bench_$w.json 2> $o/${t}_bench_$w.err
  python - <<PY
import json
try:
    d=json.loads(open("$o/${t}_bench_$w.json").read().strip().splitlines()[-1])
    print("$w", "value %.0f e2e %.0f conv %.4f iters %.1f ms/step %.2f alone_ms %.2f frac %.4f" % (d["value"], d["e2e"]["value"], d["converged_fraction"], d["mean_iters"], d["ms_per_step"], d["roofline"]["one_batch_alone"]["kernel_ms"], d["roofline"]["frac"]), "lat", d.get("single_solve_latency"))
except Exception as e:
    print("$w failed", e); print(open("$o/${t}_bench_$w.err").read()[-800:])
PY
done
