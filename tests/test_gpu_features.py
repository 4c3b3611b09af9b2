"""GPU tests of the paths round 1 left untested: classification and output coverage of a mixed batch, z_out
aliasing x0, the small-batch graph path, dynamic-obstacle prediction, the +-100 box, VelLimitConstraints with
control_mode vel through the planner, BASELINE configs 3 and 4 at full size, and the boxer closed loop with per-step
free-space decomposition and goal switching."""
import ctypes
import json
import os

import numpy as np
import pytest
import yaml

from conftest import CONFIG, GOLDEN, SOLVERS
from robot_mpcs_b200 import scenarios as sc

pytestmark = pytest.mark.gpu


def _solver(sdir, B):
    from robot_mpcs_b200.planner.solver import Solver
    return Solver.from_directory(os.path.join(SOLVERS, sdir), device=0, max_batch=B)


def _pm(sdir):
    return yaml.safe_load(open(os.path.join(SOLVERS, sdir, "paramMap.yaml")))


def _dev(a):
    import torch
    return torch.as_tensor(a, device="cuda:0")


def _mixed_batch(B=256):
    """boxer batch whose odd problems warm-start from a plan an obstacle has moved onto (rows start infeasible)."""
    sdir = "boxer_n3_01_H30_noSlack"
    pm = _pm(sdir)
    s = sc.boxer_radial(pm, B, seed=21)
    solver = _solver(sdir, B)
    first = solver.solve_batch(_dev(s["xinit"]), _dev(s["x0"]), _dev(s["params"]))
    z1 = first["z"].cpu().numpy()
    ee = z1[:, :, 0:2] + 0.4 * np.stack([np.cos(z1[:, :, 2]), np.sin(z1[:, :, 2])], 2)
    t = ee[:, 19] - ee[:, 17]
    t /= np.linalg.norm(t, axis=1, keepdims=True) + 1e-12
    c = ee[:, 18] + 0.85 * np.stack([-t[:, 1], t[:, 0]], 1)
    P = s["params"].copy()
    oi = pm["obst"][0]
    odd = np.arange(B) % 2 == 1
    P[odd, :, oi:oi + 2] = c[odd, None, :]
    P[odd, :, oi + 2] = 0.0
    P[odd, :, oi + 3] = 0.3
    x0 = s["x0"].copy()
    x0[odd] = z1[odd]
    return solver, s["xinit"], x0, P, odd


def test_mixed_batch_every_problem_written_once_and_independent_of_batch_composition(built):
    import torch
    solver, xinit, x0, P, odd = _mixed_batch()
    B = xinit.shape[0]
    r = solver.solve_batch(_dev(xinit), _dev(x0), _dev(P))
    ef = r["exitflag"].cpu().numpy()
    assert (ef != -1).all()                                           # -1 = never written (Solver prefills the array)
    assert int(r["iters"].min()) >= 1
    assert (ef[~odd] == 1).all() and (ef[odd] == 1).mean() > 0.8
    # each list alone: the same bits, whatever else is in the batch and whichever kernel's queue order served it
    for sel in (odd, ~odd):
        rs = solver.solve_batch(_dev(xinit[sel]), _dev(x0[sel]), _dev(P[sel]))
        for k in ("z", "action", "exitflag", "iters", "obj"):
            assert torch.equal(rs[k], r[k][torch.as_tensor(sel, device="cuda:0")]), k
    # z_out aliasing x0 (include/rmpc.h): the plan overwrites the guess it was started from
    buf = _dev(x0).clone()
    act = torch.empty(B, solver.nu, dtype=torch.float64, device="cuda:0")
    rc = solver.lib.rmpc_solve_batch(solver._h, B, _dev(xinit).data_ptr(), buf.data_ptr(), _dev(P).data_ptr(), buf.data_ptr(), act.data_ptr(), None, None)
    torch.cuda.synchronize()
    assert rc == 0 and torch.equal(buf, r["z"]) and torch.equal(act, r["action"])


def test_small_batch_graph_path_matches_device_path(built):
    """B <= 32 on host buffers: packed copy + captured graph + single-launch kernel; the same bits as the batched device path,
    for every output selection, and re-captured when the options change."""
    sdir = "pointRobot_n3_005_H20_noSlack"
    s = sc.point_robot_sweep(_pm(sdir), 64, seed=17)
    solver = _solver(sdir, 64)
    ref = solver.solve_batch(_dev(s["xinit"]), _dev(s["x0"]), _dev(s["params"]))          # classification pass + list kernels
    small = solver.solve_batch(_dev(s["xinit"][:32]), _dev(s["x0"][:32]), _dev(s["params"][:32]))   # single-launch kernel, device pointers
    assert (ref["exitflag"] == 1).all() and np.array_equal(small["iters"].cpu().numpy(), ref["iters"][:32].cpu().numpy())
    assert np.allclose(small["z"].cpu().numpy(), ref["z"][:32].cpu().numpy(), rtol=0, atol=1e-9)   # two builds of the same iteration
    for B in (1, 7, 32, 1):
        r = solver.solve_batch(s["xinit"][:B], s["x0"][:B], s["params"][:B])
        assert np.array_equal(r["z"], small["z"][:B].cpu().numpy()) and np.array_equal(r["action"], small["action"][:B].cpu().numpy())
        assert np.array_equal(r["exitflag"], small["exitflag"][:B].cpu().numpy()) and np.array_equal(r["obj"], small["obj"][:B].cpu().numpy())
    r = solver.solve_batch(s["xinit"][:3], s["x0"][:3], s["params"][:3], want_z=False)
    assert r["z"] is None and np.array_equal(r["action"], small["action"][:3].cpu().numpy())
    out, flag, info = solver.solve({"xinit": s["xinit"][5], "x0": s["x0"][5].reshape(-1), "all_parameters": s["params"][5].reshape(-1)})
    assert flag == 1 and np.array_equal(np.stack([out[k] for k in sorted(out)]), small["z"][5].cpu().numpy())
    solver.set_options(max_iter=4)
    r = solver.solve_batch(s["xinit"][:1], s["x0"][:1], s["params"][:1])
    assert int(r["iters"][0]) == 4 and int(r["exitflag"][0]) == 0     # the captured launch was rebuilt with the new options


def test_dynamic_obstacle_prediction(built):
    """updateDynamicObstacles (mpcPlanner.py:144-161; the reference's own method raises, tests/test_planner_reference.py):
    the formula of mpcPlanner.py:158 written out as the reference's loops (with the per-obstacle slice its intent implies),
    the host planner and the device planner (rmpc_predict_obstacles) agree bit for bit for two moving obstacles and three
    parked slots; the prediction then reaches the solver as stage-varying parameters."""
    import torch
    from robot_mpcs_b200.planner.mpcPlanner import MPCPlanner
    from robot_mpcs_b200.utils.utils import parse_setup
    setup = parse_setup(os.path.join(CONFIG, "boxerRadialMpc.yaml"))["mpc"]
    B, N, K, m = 5, setup["time_horizon"], setup["number_obstacles"], 3
    pm = _pm("boxer_n3_01_H30_noSlack")
    g = np.random.default_rng(3)
    arr = g.uniform(-2, 2, (B, 2 * 3 * m))                            # 2 moving obstacles: (pos, vel, acc) each
    host = MPCPlanner("boxer", SOLVERS + os.sep, batch_size=B, **setup)
    dev = MPCPlanner("boxer", SOLVERS + os.sep, batch_size=B, device_resident=True, **setup)
    for pl in (host, dev):
        pl.concretize(); pl.reset()
        pl.updateDynamicObstacles(arr if pl is host else _dev(arr))
    dt, r = setup["time_step"], 0.1
    want = np.zeros((B, N, 4 * K))
    for b in range(B):                                                # mpcPlanner.py:144-161, loop for loop
        for j in range(K):
            if j < 2:
                blk = arr[b, j * 3 * m:(j + 1) * 3 * m]
                pos, vel, acc = blk[:m], blk[m:2 * m], blk[2 * m:3 * m]
            else:
                pos, vel, acc = np.ones(m) * -100, np.zeros(m), np.zeros(m)
            for i in range(N):
                for mi in range(m):
                    want[b, i, j * (m + 1) + mi] = pos[mi] + vel[mi] * dt * i + 0.5 * (dt * i) ** 2 * acc[mi]
                want[b, i, j * (m + 1) + m] = r
    assert np.array_equal(host._params3[:, :, pm["obst"]], want)
    assert np.array_equal(dev._params3[:, :, pm["obst"]].cpu().numpy(), want)
    assert float(dev._params3[:, :, : pm["obst"][0]].abs().sum()) == float(torch.as_tensor(host._params3[:, :, : pm["obst"][0]]).abs().sum())
    # through solve(): an observation longer than nx carries the obstacle states (mpcPlanner.py:243-244)
    s = sc.boxer_radial(pm, B, seed=4)
    for pl in (host, dev):
        pl.setRadialConstraints(s["params"][:, 0, pm["obst"]].reshape(B, K, 4) if pl is host else _dev(s["params"][:, 0, pm["obst"]].reshape(B, K, 4)), 0.6)
        pl.setJointLimits(np.array([[-10.0] * 3, [10.0] * 3])); pl.setInputLimits(np.array([[-10.0] * 2, [10.0] * 2]))
        pl.setGoalReaching(s["params"][:, 0, pm["goal"]] if pl is host else _dev(s["params"][:, 0, pm["goal"]])); pl.setConstraintAvoidance()
    far = np.concatenate([np.tile([30.0, 30.0, 0.0], (B, 1)), np.tile([0.5, 0.0, 0.0], (B, 1)), np.zeros((B, 3))], 1)   # one obstacle drifting far away
    ob = np.concatenate([s["xinit"], far], 1)
    ah, zh, _, fh = host.solve(ob)
    ad, zd, _, fd = dev.solve(_dev(ob))
    assert np.array_equal(host._params3[:, 7, pm["obst"][0]], 30.0 + 0.5 * 0.1 * 7 * np.ones(B))
    assert np.array_equal(np.asarray(fh), fd.cpu().numpy()) and np.array_equal(np.asarray(zh), zd.cpu().numpy())


def test_box_of_the_nlp_is_reported(built):
    """The reference's NLP bounds every stage variable by +-100 (mpcModel.py:23-28,98-104).  The solver does not carry the
    box as rows; a converged plan that touches it is not a solution of that NLP and comes back with exit flag -9, on the
    GPU and on the host twin alike; the same scenario moved inside the box converges normally."""
    from oracle.build_twin import Twin
    sdir = "pointRobot_n3_005_H20_noSlack"
    pm = _pm(sdir)
    s = sc.point_robot_sweep(pm, 2, seed=1)
    P = s["params"]
    P[:, :, pm["lower_limits"]] = -1000.0
    P[:, :, pm["upper_limits"]] = 1000.0
    P[:, :, pm["obst"][0]] = -500.0                                    # obstacle out of the way
    P[:, :, pm["goal"][1]] = 0.0
    for b, x in enumerate((99.5, 49.5)):                               # 6 m/s towards +x, |u| <= 1: 5.5 m further after the 1 s horizon
        P[b, :, pm["goal"][0]] = x + 50.0
        s["xinit"][b] = [x, 0, 0, 6.0, 0, 0]
        s["x0"][b, :, :6] = s["xinit"][b]
    solver = _solver(sdir, 2)
    r = solver.solve_batch(_dev(s["xinit"]), _dev(s["x0"]), _dev(P))
    tw = Twin(os.path.join(SOLVERS, sdir)).solve(s["xinit"], s["x0"], P)
    assert r["exitflag"].tolist() == [-9, 1] and tw["exitflag"].tolist() == [-9, 1]
    assert float(r["z"][0].abs().max()) > 100.0 and float(r["z"][1].abs().max()) < 100.0


def test_vel_limits_and_vel_mode_through_the_planner(built):
    """VelLimitConstraints (inequalities/VelLimitConstraints.py:19-31) and control_mode vel (mpcPlanner.py:275-278) through
    MPCPlanner's own set* calls and the real kernel: the parameters equal the scenario generator's, the action is
    z_2[nvar-2nu : nvar-nu] = (v, omega) of stage 2, the limits hold and are active, and the plan is the oracle's."""
    from robot_mpcs_b200.planner.mpcPlanner import MPCPlanner
    from robot_mpcs_b200.utils.utils import parse_setup
    g = json.load(open(os.path.join(GOLDEN, "boxer_vel.json")))
    pm = _pm(g["solver"])
    s = sc.boxer_vel(pm, **g["kwargs"])
    setup = parse_setup(os.path.join(CONFIG, "boxerVelMpc.yaml"))["mpc"]
    assert setup["control_mode"] == "vel" and "VelLimitConstraints" in setup["constraints"]
    B = s["xinit"].shape[0]
    pl = MPCPlanner("boxerVel", SOLVERS + os.sep, batch_size=B, **setup)
    pl.concretize(); pl.reset()
    P0 = s["params"][:, 0]
    pl.setRadialConstraints(P0[:, pm["obst"]].reshape(B, 2, 4), 0.6)
    pl.setSelfCollisionAvoidanceConstraints(0.6)
    pl.setJointLimits(np.array([[-10.0] * 3, [10.0] * 3]))
    pl.setInputLimits(np.array([[-10.0] * 2, [10.0] * 2]))
    pl.setVelLimits(np.array([[-0.3, -1.0], [1.0, 1.0]]))
    pl.setGoalReaching(P0[:, pm["goal"]])
    pl.setConstraintAvoidance()
    assert np.array_equal(pl._params3, s["params"])
    action, z, flag = pl.computeAction(s["xinit"][:, :3], s["xinit"][:, 3:6], s["xinit"][:, 6:])
    assert (flag == 1).all()
    assert np.array_equal(action, z[:, 1, 6:8])                        # "vel": x02[-2nu:-nu]
    v, om = z[:, :, 6], z[:, :, 7]
    assert v.max() <= 1.0 + 1e-9 and v.min() >= -0.3 - 1e-9 and np.abs(om).max() <= 1.0 + 1e-9
    assert (v.max(1) > 0.999).mean() > 0.5                             # the forward-speed limit is active in most scenarios
    same = 0
    for c in g["cases"]:
        zo = np.array(c["Z"])
        same += bool(np.abs(z[c["index"]] - zo).max() / np.abs(zo).max() < 1e-4)
    assert same >= 0.93 * len(g["cases"])
    # interval > 1 (mpcPlanner.py:293-301): the solve runs every `interval` calls, the action is held in between
    setup["interval"] = 3
    pl = MPCPlanner("boxerVel", SOLVERS + os.sep, batch_size=B, **setup)
    pl.concretize(); pl.reset()
    pl._params3[:] = s["params"]
    calls = []
    inner = pl._solver.solve_batch
    pl._solver.solve_batch = lambda *a, **k: (calls.append(1), inner(*a, **k))[1]
    acts = [pl.computeAction(s["xinit"][:, :3], s["xinit"][:, 3:6], s["xinit"][:, 6:])[0] for _ in range(7)]
    assert len(calls) == 3 and all(np.array_equal(acts[0], a) for a in acts[1:3]) and np.array_equal(acts[3], acts[5])


@pytest.mark.parametrize("sdir,cfg,gen,B", [
    ("panda_n7_005_H20_noSlack", "pandaMpc.yaml", "panda", 1024),
    ("boxer_n3_01_H40_noSlack", "boxerSupermarketMpc.yaml", "boxer_supermarket", 16384),
])
def test_full_size_config3_config4(built, sdir, cfg, gen, B):
    """BASELINE configs 3 (panda, B = 1024) and 4 (supermarket, B = 16384) at full size: converged-fraction floor and the
    size-independent properties of every converged plan (x_1 = xinit, dynamics defect, feasibility, box, same bits again)."""
    import torch
    from oracle.nlp import load_model
    s = getattr(sc, gen)(_pm(sdir), B)
    d = {k: _dev(v) for k, v in s.items()}
    solver = _solver(sdir, B)
    r = solver.solve_batch(d["xinit"], d["x0"], d["params"])
    r2 = solver.solve_batch(d["xinit"], d["x0"], d["params"])
    assert torch.equal(r["z"], r2["z"]) and torch.equal(r["exitflag"], r2["exitflag"])
    ef = r["exitflag"].cpu().numpy()
    print(sdir, "converged", (ef == 1).mean(), dict(zip(*np.unique(ef, return_counts=True))), "iters mean", float(r["iters"].double().mean()), "max", int(r["iters"].max()))
    assert (ef == 1).mean() >= 0.995
    ok = torch.as_tensor(ef == 1)
    m = load_model(os.path.join(CONFIG, cfg))
    Z, Pm = r["z"].cpu()[ok], torch.as_tensor(s["params"])[ok]
    X, U = Z[:, :, : m.nx], Z[:, :, m.nx:]
    assert float((X[:, 0] - torch.as_tensor(s["xinit"])[ok]).abs().max()) == 0.0
    assert float((X[:, 1:] - m.step(X[:, :-1], U[:, :-1])).abs().max()) < 1e-7
    assert float(m.ineq(Z, Pm).min()) > -1e-7
    assert float(Z.abs().max()) < 100.0


def _corridor_scene(B, seed=0):
    """Boxer scenarios in a corridor: start pose, a goal 4-6 m ahead, and 64 lidar returns on two walls 1.6 m to either side
    of the start->goal line (the stand-in for urdfenvs' lidar, boxer_example.py:153-161)."""
    g = np.random.default_rng(seed)
    th = g.uniform(-np.pi, np.pi, B)
    xy = g.uniform(-1, 1, (B, 2))
    L = g.uniform(4, 6, B)
    dirv = np.stack([np.cos(th), np.sin(th)], 1)
    nrm = np.stack([-np.sin(th), np.cos(th)], 1)
    goal = xy + L[:, None] * dirv
    along = np.linspace(-1.5, 7.5, 32)
    pts = np.concatenate([xy[:, None, :] + along[None, :, None] * dirv[:, None, :] + sgn * 1.6 * nrm[:, None, :] for sgn in (1.0, -1.0)], 1)
    pts = np.concatenate([pts, np.full((B, 64, 1), 0.02)], 2)
    xinit = np.concatenate([xy, (th + g.uniform(-0.4, 0.4, B))[:, None], np.zeros((B, 3)), g.uniform(0, 0.3, (B, 1)), np.zeros((B, 1))], 1)
    return xinit, goal, np.ascontiguousarray(pts)


def test_boxer_closed_loop_with_per_step_free_space_decomposition(built):
    """examples/boxer_example.py:152-224 without the simulator: per control step N free-space decompositions per scenario
    (seeded from the previous plan, :191-201) write the half-planes, the batched solve plans, the model is the plant, the goal
    check is :219-224.  Properties: the ee point keeps r_body from every lidar return it was constrained by, scenarios make
    their way down the corridor, warm starts converge."""
    import torch
    from robot_mpcs_b200.planner.mpcPlanner import MPCPlanner
    from robot_mpcs_b200.rollout import LocalGoalPicker, closed_loop
    from robot_mpcs_b200.utils.utils import parse_setup
    B, T = 128, 90
    setup = parse_setup(os.path.join(CONFIG, "boxerSupermarketMpc.yaml"))["mpc"]
    setup["initialization"] = "previous_plan"
    xinit, goal, pts = _corridor_scene(B)
    pl = MPCPlanner("boxer", SOLVERS + os.sep, batch_size=B, device_resident=True, **setup)
    pl.concretize(); pl.reset()
    K = setup["number_obstacles"]
    pl.setLinearConstraints(np.tile([1.0, 0.0, 0.0, 100.0], (setup["time_horizon"], K, 1)), 0.6)   # filler planes until the first FSD (boxer_example.py:93)
    pl.setJointLimits(np.array([[-30.0] * 3, [30.0] * 3])); pl.setInputLimits(np.array([[-10.0] * 2, [10.0] * 2]))
    pl.setGoalReaching(_dev(goal)); pl.setConstraintAvoidance()
    P = _dev(pts)
    clear, flags = [], []

    def on_step(t, x, z, flag):
        ee = x[:, :2] + 0.4 * torch.stack([torch.cos(x[:, 2]), torch.sin(x[:, 2])], 1)
        clear.append(torch.linalg.norm(ee[:, None, :] - P[:, :, :2], dim=2).min(1).values)
        flags.append(flag.clone())

    r = closed_loop(pl, _dev(xinit), _dev(goal), T, epsilon=0.4, fsd_points=P, on_step=on_step)
    cl = torch.stack(clear)
    ok_frac = float((torch.stack(flags) == 1).double().mean())
    d0 = np.linalg.norm(xinit[:, :2] - goal, axis=1)
    print("FSD loop: solves ok", ok_frac, "min clearance", float(cl.min()), "reached", float(r["reached"].double().mean()),
          "mean progress", float((_dev(d0) - r["goal_distance"]).mean()), "warm iters", float(r["mean_iters"][5:].mean()), "first", float(r["mean_iters"][0]))
    assert ok_frac > 0.97
    assert float(cl.min()) > 0.6 - 5e-3                               # r_body from every return, up to the plant/plan mismatch of a failed step
    assert float(r["reached"].double().mean()) > 0.8
    # goal switching on the same scene (boxer_example_supermarket.py:235-237): three waypoints down the corridor
    pl.reset()
    pl.setLinearConstraints(np.tile([1.0, 0.0, 0.0, 100.0], (setup["time_horizon"], K, 1)), 0.6)
    pl.setJointLimits(np.array([[-30.0] * 3, [30.0] * 3])); pl.setInputLimits(np.array([[-10.0] * 2, [10.0] * 2])); pl.setConstraintAvoidance()
    way = np.stack([xinit[:, :2] + f * (goal - xinit[:, :2]) for f in (0.35, 0.7, 1.0)], 1)
    picker = LocalGoalPicker(_dev(way), threshold_local_goal=1.3)
    r2 = closed_loop(pl, _dev(xinit), _dev(goal), T, epsilon=0.4, fsd_points=P, global_path=picker)
    print("FSD + global path: reached", float(r2["reached"].double().mean()), "last waypoint", float((picker.idx_local == 2).double().mean()))
    assert float((picker.idx_local == 2).double().mean()) > 0.9 and float(r2["reached"].double().mean()) > 0.8


@pytest.mark.parametrize("name", ["pointRobot_sweep", "boxer_radial", "panda_jitter", "boxer_supermarket"])
def test_mixed_precision_against_the_fixtures_and_float64(built, name):
    """rmpc_options.precision = 1 (north_star: "fp32 plus fp64 where the barrier conditioning demands it"): float32
    iterations while the barrier parameter is large, float64 to the same final tolerances.  Against the oracle's from-x0
    fixtures it must reach the basin-agreement floor of precision 0 (less 5 points) at the same bars (1e-4 / 1e-5), converge
    wherever float64 does, and agree with the float64 kernel to 1e-6 wherever both stop in the same basin."""
    import torch
    from test_gpu_parity import FIXTURES, GEN
    g = json.load(open(os.path.join(GOLDEN, name + ".json")))
    s = GEN[g["generator"]](_pm(g["solver"]), **g["kwargs"])
    d = {k: _dev(v) for k, v in s.items()}
    B = d["xinit"].shape[0]
    solver = _solver(g["solver"], B)
    r0 = solver.solve_batch(d["xinit"], d["x0"], d["params"])
    solver.set_options(precision=1)
    r1 = solver.solve_batch(d["xinit"], d["x0"], d["params"])
    r1b = solver.solve_batch(d["xinit"], d["x0"], d["params"])
    assert torch.equal(r1["z"], r1b["z"])                                       # deterministic
    assert not torch.equal(r1["z"], r0["z"])                                    # and really another arithmetic
    e0, e1 = r0["exitflag"].cpu().numpy(), r1["exitflag"].cpu().numpy()
    assert (e1 == 1).sum() >= (e0 == 1).sum() - 1
    z0, z1 = r0["z"].cpu().numpy(), r1["z"].cpu().numpy()
    both = (e0 == 1) & (e1 == 1)
    err = np.abs(z1[both] - z0[both]).max(axis=(1, 2)) / np.abs(z0[both]).max(axis=(1, 2))
    print(name, "precision 1 vs 0: same basin", float((err < 1e-5).mean()), "median", float(np.median(err)), "iters", float(r1["iters"].double().mean()), float(r0["iters"].double().mean()))
    assert np.median(err) < 1e-6 and (err < 1e-5).mean() > 0.85
    same = valid = 0
    for c in g["cases"]:
        k = c["kkt"]
        if not (k["eq"] < 1e-6 and k["ineq"] < 1e-6 and k["stat"] < 1e-5):
            continue
        valid += 1
        b, zo = c["index"], np.array(c["Z"])
        assert e1[b] == 1
        same += bool(np.abs(z1[b] - zo).max() / np.abs(zo).max() < 1e-4 and abs(float(r1["obj"][b]) - c["obj"]) <= 1e-5 * abs(c["obj"]))
    assert same >= (FIXTURES[name] - 0.05) * valid, (name, same, valid)


def test_closed_loop_in_vel_mode_follows_its_own_plan(built):
    """rollout.closed_loop with control_mode vel: the velocity action is turned back into the input that realises it, so the
    plant lands on the plan's own stage-2 state, and the velocity limits hold along the whole run."""
    import torch
    from robot_mpcs_b200.planner.mpcPlanner import MPCPlanner
    from robot_mpcs_b200.rollout import closed_loop
    from robot_mpcs_b200.utils.utils import parse_setup
    B, T = 64, 25
    setup = parse_setup(os.path.join(CONFIG, "boxerVelMpc.yaml"))["mpc"]
    setup["initialization"] = "previous_plan"
    pm = _pm("boxerVel_n3_01_H15_noSlack")
    s = sc.boxer_vel(pm, B, seed=11)
    pl = MPCPlanner("boxerVel", SOLVERS + os.sep, batch_size=B, device_resident=True, **setup)
    pl.concretize(); pl.reset()
    pl._params3[:] = _dev(s["params"])
    err, vmax = [], []

    def on_step(t, x, z, flag):
        ok = flag == 1
        err.append((z[ok, 1, :8] - x[ok]).abs().max())
        vmax.append(torch.stack([x[:, 6].max(), x[:, 7].abs().max()]))

    goal = _dev(s["params"][:, 0, pm["goal"]])
    r = closed_loop(pl, _dev(s["xinit"]), goal, T, epsilon=0.3, on_step=on_step)
    assert float(torch.stack(err).max()) < 1e-8
    vm = torch.stack(vmax).max(0).values
    assert float(vm[0]) <= 1.0 + 1e-6 and float(vm[1]) <= 1.0 + 1e-6
    d0 = torch.linalg.norm(_dev(s["xinit"])[:, :2] - goal[:, :2], dim=1)
    assert float((d0 - r["goal_distance"]).mean()) > 1.0 and int(r["failures"].sum()) <= 2


@pytest.mark.parametrize("B", [8, 96])          # single-launch kernel (B <= 32) and classify + two kernels
@pytest.mark.parametrize("precision", [0, 1])
def test_nan_and_inf_inputs_are_refused_without_touching_their_neighbours(built, B, precision):
    """A state, guess or parameter that is not a number (a sensor glitch): exitflag -6 before the first iteration, iters 0, and
    every other problem of the batch is solved exactly as without the bad ones.  FORCESPRO answers such inputs with a negative
    exitflag as well; the reference prints it and applies no action (mpcPlanner.py:263-264, boxer_example.py:194)."""
    import torch
    sdir = "boxer_n3_01_H30_noSlack"
    pm = _pm(sdir)
    s = sc.boxer_radial(pm, B, seed=33)
    solver = _solver(sdir, B)
    solver.set_options(precision=precision)
    clean = solver.solve_batch(_dev(s["xinit"]), _dev(s["x0"]), _dev(s["params"]))
    x, x0, P = s["xinit"].copy(), s["x0"].copy(), s["params"].copy()
    x[1, 0] = np.nan
    x0[2, 3, 1] = np.nan
    P[3, :, pm["goal"][0]] = np.inf
    P[4, 5, pm["wu"][0]] = np.nan
    x[5, 1] = -np.inf
    bad = np.zeros(B, bool)
    bad[1:6] = True
    r = solver.solve_batch(_dev(x), _dev(x0), _dev(P))
    flag, it = r["exitflag"].cpu().numpy(), r["iters"].cpu().numpy()
    assert (flag[bad] == -6).all() and (it[bad] == 0).all(), (flag[:8], it[:8])
    good = torch.as_tensor(~bad, device="cuda:0")
    assert torch.equal(r["exitflag"][good], clean["exitflag"][good]) and torch.equal(r["iters"][good], clean["iters"][good])
    assert torch.equal(r["z"][good], clean["z"][good]) and torch.equal(r["action"][good], clean["action"][good])
    assert (clean["exitflag"] == 1).float().mean() > 0.95
