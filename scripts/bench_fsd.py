#!/usr/bin/env python
"""Free-space decomposition throughput on one B200: B problems x N=40 stages x K=8 planes from 64 lidar
returns, planes written into the parameter array.  Prints one JSON line: decompositions/s, HBM roofline (algorithmic bytes
= points + reference states read, planes written), and the oracle (numpy restatement of the reference's Python loop) timed
on a bounded sample.  usage: python scripts/bench_fsd.py [--batch 16384] [--steps 20]"""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, yaml
from robot_mpcs_b200 import scenarios as sc
from robot_mpcs_b200.planner.solver import Solver

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=16384)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--cpu-sample", type=int, default=16)
a = ap.parse_args()
sdir = os.path.join(ROOT, "robot_mpcs_b200", "solvers", "boxer_n3_01_H40_noSlack")
solver = Solver.from_directory(sdir, device=0, max_batch=a.batch)
pm = yaml.safe_load(open(os.path.join(sdir, "paramMap.yaml")))
K, P, B, N = 8, 64, a.batch, solver.N
idx = [pm[f"lin_constrs_{i}"][0] for i in range(K)]
pts, ref = sc.lidar_scene(B, N, P, seed=0)
dev = torch.device("cuda:0")
dp, dr = torch.as_tensor(pts, device=dev), torch.as_tensor(ref, device=dev)
params = torch.zeros(B, N, solver.npar, dtype=torch.float64, device=dev)
flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
ms = []
for i in range(a.warmup + a.steps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    solver.fsd_constraints(dp, dr, K, 5.0, params=params, idx=idx, want_out=False)
    e1.record()
    torch.cuda.synchronize()
    if i >= a.warmup:
        ms.append(e0.elapsed_time(e1))
t = float(np.mean(ms)) * 1e-3
bytes_alg = 8.0 * (B * P * 3 + B * N * 3 + B * N * K * 4)
peaks = {}
try:
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:
    pass
hbm = peaks.get("hbm_gbs", 6650.0)
from oracle import fsd as ofsd
bs = min(B, a.cpu_sample)
t0 = time.perf_counter()
ofsd.batch(pts[:bs], ref[:bs], K, 5.0)
tc = time.perf_counter() - t0
print(json.dumps({"metric": "free-space decompositions/s (one per problem and stage)", "value": B * N / t, "unit": "decompositions/s",
                  "ms_per_step": t * 1e3, "config": {"workload": f"lidar_scene B={B} N={N} P={P} K={K}, planes scattered into params", "l2": "flushed between steps"},
                  "dtype": "f64", "gpu_launches": a.steps,
                  "roofline": {"bound": "hbm", "achieved": bytes_alg / t / 1e9, "peak": hbm, "unit": "GB/s", "frac": bytes_alg / t / 1e9 / hbm,
                               "algorithmic_bytes": bytes_alg, "traffic": None},
                  "cpu_baseline": {"value": bs * N / tc, "unit": "decompositions/s", "cores": 1, "kind": "port",
                                   "sample": f"{bs} problems x {N} stages, oracle/fsd.py (numpy, the reference's own loop structure), {tc:.2f} s"}}))
