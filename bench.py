#!/usr/bin/env python
"""bench.py — converged MPC solves/s of the batched solve (BASELINE.json metric) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--batch B] [--impl ours|reference]

A "step" is one batched solve of `batch` synthetic scenarios per GPU (weak scaling).  Default workload at
N=1 is BASELINE.json configs[1]: boxer diff-drive, N=30, 5 sphere obstacles, batch 4096.
Prints ONE JSON line (rank 0).  `--impl reference` times the CPU path (oracle/cpu_twin.cpp, all host threads)
on the same workload; the reference's own solver (FORCESPRO-generated C) is proprietary and absent.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from robot_mpcs_b200 import scenarios as sc  # noqa: E402

SOLVERS = os.path.join(ROOT, "robot_mpcs_b200", "solvers")

# name -> (solver dir, generator, default batch, F_iter flop/iteration (SURVEY §8d convention), description)
WORKLOADS = {
    "boxer_radial": ("boxer_n3_01_H30_noSlack", sc.boxer_radial, 4096, 94e3,
                     "boxer diff-drive N=30, 5 sphere obstacles (BASELINE configs[1])"),
    "pointRobot_sweep": ("pointRobot_n3_005_H20_noSlack", sc.point_robot_sweep, 131072, 34e3,
                         "pointRobot N=20, goal + 1 sphere obstacle, random scenarios (BASELINE configs[4])"),
    "panda": ("panda_n7_005_H20_noSlack", sc.panda, 1024, 351e3,
              "panda 7-DoF N=20, FK-sphere self-collision + joint limits (BASELINE configs[2])"),
    "boxer_supermarket": ("boxer_n3_01_H40_noSlack", sc.boxer_supermarket, 16384, 127e3,
                          "boxer N=40, 8 stage-varying half-planes (BASELINE configs[3])"),
}


# dram__bytes_read.sum + dram__bytes_write.sum of one solve at the default batch (ncu --set full, each kernel captured
# alone with a cold L2), and where it is kept.  A solve is two concurrent launches: solve_kernel<1> (feasible starts,
# 13.3 ms alone) and solve_kernel<2> (infeasible starts: for this workload it reads every problem once, finds none and
# exits after 38 us).  Both read the inputs, hence twice the algorithmic bytes when neither finds them in L2.
NCU_TRAFFIC = {"boxer_radial": ((52.855296e6 + 0.603392e6) + (52.525568e6 + 0.399104e6), "profiles/r1m_ncu_boxer_radial_B4096.txt")}


_OUT = None


def emit(line):
    """The one JSON line of the contract, on the process's original stdout."""
    print(json.dumps(line), file=_OUT or sys.stdout, flush=True)


def load_paramMap(sdir):
    import yaml
    with open(os.path.join(SOLVERS, sdir, "paramMap.yaml")) as f:
        return yaml.safe_load(f)


def make_inputs(workload, B, seed):
    sdir, gen, _, _, _ = WORKLOADS[workload]
    return gen(load_paramMap(sdir), B, seed=seed)


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index=0):
        self.proc = None
        self.rows = []
        self.index = index

    def start(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def measure_fma_peak(torch, dev, is_double=True):
    so = os.path.join(SOLVERS, "libpeak_fma.so")
    if not os.path.isfile(so):
        return None
    lib = ctypes.CDLL(so)
    lib.peak_fma_launch.restype = ctypes.c_double
    lib.peak_fma_launch.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    blocks, threads, iters = 148 * 8, 256, 20000
    out = torch.empty(blocks * threads, dtype=torch.float64 if is_double else torch.float32, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    best = 0.0
    for i in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fl = lib.peak_fma_launch(out.data_ptr(), blocks, threads, iters, 1 if is_double else 0, ctypes.c_void_p(st))
        e1.record()
        torch.cuda.synchronize(dev)
        if fl > 0 and i > 0:
            best = max(best, fl / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best


def cpu_baseline(workload, B_sample, seed, threads=0):
    """The CPU path (oracle/cpu_twin.cpp: same algorithm, float64, OpenMP over problems) on a bounded sample."""
    from oracle.build_twin import Twin
    sdir = WORKLOADS[workload][0]
    tw = Twin(os.path.join(SOLVERS, sdir))
    s = make_inputs(workload, B_sample, seed)
    if threads <= 0:   # every host core this process may use, whatever OMP_NUM_THREADS says (torchrun sets it to 1)
        threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    tw.solve(s["xinit"][:64], s["x0"][:64], s["params"][:64], nthreads=threads)  # warm-up
    t = time.perf_counter()
    r = tw.solve(s["xinit"], s["x0"], s["params"], nthreads=threads)
    dt = time.perf_counter() - t
    conv = int((r["exitflag"] == 1).sum())
    cores = threads
    return {"value": conv / dt, "unit": "converged solves/s", "cores": cores, "kind": "port",
            "sample": f"{B_sample} scenarios of workload {workload} (seed {seed}), one pass, {dt:.2f} s, converged {conv}/{B_sample}, "
                      f"mean iters {float(r['iters'].mean()):.1f}"}, dt, conv


def single_solve_latency(workload, n, solve_one):
    """p50/p99 wall time of `n` single solves (B = 1) on `n` different scenarios of the workload; solve_one(xinit, x0, params)
    takes host arrays and returns once the action is back on the host (BASELINE.json metric, second half)."""
    s = make_inputs(workload, n, seed=4242)
    xs, x0s, ps = (np.ascontiguousarray(s[k]) for k in ("xinit", "x0", "params"))
    for i in range(min(n, 10)):
        solve_one(xs[i:i + 1], x0s[i:i + 1], ps[i:i + 1])
    ts, conv = [], 0
    for i in range(n):
        t = time.perf_counter()
        ef = solve_one(xs[i:i + 1], x0s[i:i + 1], ps[i:i + 1])
        ts.append(time.perf_counter() - t)
        conv += int(ef == 1)
    ts = np.array(ts) * 1e3
    return {"unit": "ms", "p50": float(np.percentile(ts, 50)), "p99": float(np.percentile(ts, 99)), "mean": float(ts.mean()),
            "n": n, "converged": conv}


def cpu_single_solve_latency(workload, n):
    from oracle.build_twin import Twin
    tw = Twin(os.path.join(SOLVERS, WORKLOADS[workload][0]))
    return single_solve_latency(workload, n, lambda x, x0, p: int(tw.solve(x, x0, p, nthreads=1)["exitflag"][0]))


# ------------------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    _, _, defB, _, desc = WORKLOADS[args.workload]
    B = args.batch or defB
    Bs = min(B, args.ref_sample)
    times, convs = [], []
    base = None
    for i in range(args.warmup + args.steps):
        base, dt, conv = cpu_baseline(args.workload, Bs, seed=100 + i)
        if i >= args.warmup:
            times.append(dt); convs.append(conv)
    val = sum(convs) / sum(times)
    base["value"] = val
    line = {"impl": "reference", "metric": "converged MPC solves/sec", "value": val, "unit": "converged solves/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {desc}", "batch_per_step": Bs, "note": "CPU twin (port) of the same IPM/Riccati; FORCESPRO/IPOPT unavailable on this box"},
            "cpu_baseline": base, "e2e": {"value": val, "unit": "converged solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if args.latency_samples > 0:
        line["single_solve_latency"] = dict(cpu_single_solve_latency(args.workload, args.latency_samples), path="CPU twin, 1 thread, B=1")
    emit(line)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from robot_mpcs_b200.planner.solver import Solver
    from robot_mpcs_b200.parallel import gather_results, init_process_group

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the solver has no CPU fallback (use --impl reference for the CPU path)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        init_process_group("nccl", device_id=dev)   # one NCCL channel: see robot_mpcs_b200/parallel.py
    sdir, _, defB, f_iter, desc = WORKLOADS[args.workload]
    B = args.batch or defB
    solver = Solver.from_directory(os.path.join(SOLVERS, sdir), device=local_rank, max_batch=B)
    solver.set_options(precision=args.precision, hessian=args.hessian, stall_window=args.stall_window)
    set_bytes = 8 * B * (solver.nx + solver.N * solver.nvar + solver.N * solver.npar)
    # rotating input sets whose total exceeds the 126 MB L2, so that no step finds its inputs cached
    nsets = int(min(16, max(2, -(-160e6 // set_bytes))))
    sets = []
    for i in range(nsets):
        s = make_inputs(args.workload, B, seed=1000 * rank + i)
        host = {k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for k, v in s.items()}
        devt = {k: v.to(dev) for k, v in host.items()}
        if args.params == "compressed":
            p3 = devt["params"].view(B, solver.N, solver.npar)
            if not bool((p3 == p3[:, :1]).all()):
                raise SystemExit(f"--params compressed: workload {args.workload} has stage-varying parameters")
            devt["params"] = p3[:, 0].contiguous()
        sets.append((host, devt))
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)  # 256 MB > 126 MB L2
    nstreams = max(1, args.streams)
    streams = [torch.cuda.Stream(dev) for _ in range(nstreams)]

    def step_device(i):
        _, d = sets[i % nsets]
        r = solver.solve_batch(d["xinit"], d["x0"], d["params"], want_z=False, want_info=True)
        # the gather runs on NCCL's stream behind this step's solve; the step's stream is not held until every rank arrives
        g = gather_results(r["action"], r["exitflag"], r["iters"], B * world, async_op=True) if world > 1 else None
        return r, g

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def run_steps(n):
        """n batched solves, step i on stream i % nstreams (consecutive batches overlap: the stragglers of one batch no
        longer idle the GPU); returns device time from the first launch to the last completion, and the per-step results."""
        main = torch.cuda.current_stream(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(main)
        for st in streams:
            st.wait_event(e0)
        res, gathers = [], []
        for i in range(n):
            with torch.cuda.stream(streams[i % nstreams]):
                r, g = step_device(i)
                res.append(r)
                gathers.append(g)
        for st in streams:
            ev = torch.cuda.Event()
            ev.record(st)
            main.wait_event(ev)
        for g in gathers:   # every step's all_gather has delivered before the clock stops
            if g is not None:
                g.wait()
        e1.record(main)
        return e0, e1, res

    # ---- device-resident throughput ("value")
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()  # nvidia-smi needs ~100 ms per sample: it watches the warm-up steps as well as the timed ones
    run_steps(args.warmup)
    barrier()
    launches0 = solver.launch_count()
    t_wall = time.perf_counter()
    e0, e1, results = run_steps(args.steps)
    barrier()
    t_wall = time.perf_counter() - t_wall
    launches = solver.launch_count() - launches0
    conv_total = sum(int((r["exitflag"] == 1).sum()) for r in results)
    iters_total = sum(int(r["iters"].sum()) for r in results)
    nsolved = B * len(results)
    total_ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    conv_t = torch.tensor([conv_total, iters_total, nsolved], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(conv_t, op=dist.ReduceOp.SUM)
    clocks = sampler.stop() if rank == 0 else None
    total_s = float(total_ms) * 1e-3
    value = float(conv_t[0]) / total_s

    # ---- one batch alone (roofline of the solve kernel, and the latency of a whole batch): L2 flushed, events around the launch
    kern_ms = []
    _, d = sets[0]
    for i in range(max(3, min(args.steps, 10))):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rr = solver.solve_batch(d["xinit"], d["x0"], d["params"], want_z=False, want_info=True)
        e1.record()
        torch.cuda.synchronize(dev)
        kern_ms.append(e0.elapsed_time(e1))
    k_ms = float(np.mean(kern_ms))
    flops = float(rr["iters"].sum()) * f_iter
    bytes_alg = 8.0 * B * (solver.nx + solver.N * solver.nvar + solver.N * solver.npar) + 8.0 * B * (solver.nu + 1) + 4.0 * B * 2

    # ---- end to end through the host entry points (pinned host buffers -> H2D -> solve -> D2H), two handles in flight
    nh = min(nstreams, 3)
    hsolvers = [solver] + [Solver.from_directory(os.path.join(SOLVERS, sdir), device=local_rank, max_batch=B) for _ in range(nh - 1)]
    for hs in hsolvers[1:]:
        hs.set_options(precision=args.precision, hessian=args.hessian, stall_window=args.stall_window)
    hnp = [{k: v.numpy() for k, v in h.items()} for h, _ in sets]

    def e2e_run(n):
        pend, conv = [None] * nh, 0
        for i in range(n + nh):
            j = i % nh
            if pend[j] is not None:
                hsolvers[j].wait()
                conv += int((pend[j]["exitflag"] == 1).sum())
                pend[j] = None
            if i < n:
                hh = hnp[i % nsets]
                pend[j] = hsolvers[j].solve_batch(hh["xinit"], hh["x0"], hh["params"], want_z=False, want_info=True, block=False)
        return conv

    e2e_run(nh + 1)  # every handle allocates its staging before the clock starts
    barrier()
    e2e_steps = max(4, args.steps)   # same K as the device-resident run: a 10-step run left a fifth of its time to the last stragglers
    t0 = time.perf_counter()
    e2e_conv = e2e_run(e2e_steps)
    barrier()
    e2e_s = time.perf_counter() - t0
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    e2e_c = torch.tensor([e2e_conv], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(e2e_c, op=dist.ReduceOp.SUM)
    h2d = set_bytes
    d2h = 8 * B * (solver.nu + 5) + 4 * B * 2

    lat = None
    if rank == 0 and args.latency_samples > 0:
        def one(x, x0, p):
            return int(solver.solve_batch(x, x0, p, want_z=False, want_info=True)["exitflag"][0])
        lat = dict(single_solve_latency(args.workload, args.latency_samples, one),
                   path="rmpc_solve_batch_host, B=1: H2D of xinit/x0/params + one-warp solve + D2H of action/info, wall clock")

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        fp64_peak = measure_fma_peak(torch, dev, True)
        fp32_peak = measure_fma_peak(torch, dev, False)
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        # average launch duration over the timed region: K launches share total_s (they overlap on the streams)
        rank_iters = float(iters_total)
        launch_ms = 1e3 * total_s / args.steps
        flops_launch = rank_iters * f_iter / args.steps
        achieved_tf = flops_launch / (launch_ms * 1e-3) / 1e12
        alone_tf = flops / (k_ms * 1e-3) / 1e12
        base = None
        if world == 1 and not args.no_cpu_baseline:
            base, _, _ = cpu_baseline(args.workload, min(B, args.ref_sample), seed=0)
        line = {
            "metric": "converged MPC solves/sec", "value": value, "unit": "converged solves/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total_s / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64" if args.precision == 0 else "f64+f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {desc}", "batch_per_gpu": B, "global_batch": B * world, "params": args.params,
                       "l2": f"{nsets} rotating input sets of {set_bytes / 1e6:.0f} MB (> 126 MB L2 in total); single-batch timing with L2 flushed (256 MB write)",
                       "streams": nstreams, "stall_window": args.stall_window, "hessian": "exact+GN fallback" if args.hessian else "gauss-newton",
                       "tolerances": {k: v for k, v in solver.options().items() if k.startswith("tol")},
                       "parallelism": f"scenario-sharded x{world}, all_gather of action/exitflag/iters per step"},
            "converged_fraction": float(conv_t[0] / conv_t[2]), "mean_iters": float(conv_t[1] / conv_t[2]),
            "wall_s": t_wall, "gpu_launches": int(launches),
            "roofline": {"bound": "fp64", "achieved": achieved_tf, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": (achieved_tf / fp64_peak) if fp64_peak else None,
                         "traffic": NCU_TRAFFIC[args.workload][0] if args.workload in NCU_TRAFFIC and not args.batch else None,
                         "traffic_source": NCU_TRAFFIC.get(args.workload, (None, None))[1],
                         "peak_source": "measured in this run (csrc/peak_fma.cu DFMA chains); MEASURED_PEAKS.json has no FP64 figure",
                         "fp32_peak_measured": fp32_peak, "kernel_ms": launch_ms, "flops_per_launch": flops_launch,
                         "basis": "algorithmic flops = iterations x F_iter (SURVEY 8d convention); kernel_ms = timed region / launches on this rank (launches overlap on the streams)",
                         "one_batch_alone": {"kernel_ms": k_ms, "achieved": alone_tf, "frac": (alone_tf / fp64_peak) if fp64_peak else None,
                                             "solves_per_s": float((rr["exitflag"] == 1).sum()) / (k_ms * 1e-3),
                                             "note": "a single batch, L2 flushed: bounded by its slowest problem (iterations x per-iteration latency of one warp)"},
                         "hbm": {"achieved": bytes_alg / (launch_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                 "frac": bytes_alg / (launch_ms * 1e-3) / 1e9 / hbm_peak, "algorithmic_bytes": bytes_alg}},
            "e2e": {"value": float(e2e_c) / float(e2e_t), "unit": "converged solves/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "api": f"rmpc_solve_batch_host_async + rmpc_wait on {nh} handle(s), pinned host buffers"},
            "clocks": clocks,
        }
        if base is not None:
            line["cpu_baseline"] = base
        if lat is not None:
            line["single_solve_latency"] = lat
            if world == 1 and not args.no_cpu_baseline:
                line["cpu_baseline"]["single_solve_latency"] = cpu_single_solve_latency(args.workload, args.latency_samples)
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="boxer_radial", choices=list(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", type=int, default=0)
    ap.add_argument("--hessian", type=int, default=1)
    ap.add_argument("--ref-sample", type=int, default=4096, help="scenarios per CPU-baseline pass")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--params", default="reference", choices=["reference", "compressed"],
                    help="device-resident parameter format: the reference's all_parameters[N*npar] per problem, or one npar block per problem "
                         "for workloads whose parameters do not vary over the horizon (rmpc_solve_batch_strided, stage stride 0)")
    ap.add_argument("--streams", type=int, default=4, help="CUDA streams the consecutive batches are spread over (1 = strictly one batch at a time)")
    ap.add_argument("--stall-window", type=int, default=0, help="rmpc_options.stall_window (0 = off, the default)")
    ap.add_argument("--latency-samples", type=int, default=300, help="single solves (B=1) timed for the p50/p99 latency; 0 = skip")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    # Libraries write to file descriptor 1 behind Python's back (NCCL's version banner at NCCL_DEBUG=VERSION ignores
    # NCCL_DEBUG_FILE): from here on fd 1 is stderr, and the JSON line goes to a private copy of the real stdout.
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
