#!/usr/bin/env python
"""Closed-loop rollout throughput on one B200: thin wrapper around bench.py's rollout_record.
usage: python scripts/bench_rollout.py [--workload pointRobot|boxer_fsd] [--batch B] [--steps T]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="boxer_fsd", choices=["pointRobot", "boxer_fsd"])
ap.add_argument("--batch", type=int, default=4096)
ap.add_argument("--steps", type=int, default=60)
a = ap.parse_args()
out = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)                      # the planner prints like the reference does
print(json.dumps(bench.rollout_record("rollout_" + a.workload, a.batch, a.steps)), file=out, flush=True)
