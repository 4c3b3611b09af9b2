/* rmpc.h — C ABI of one generated batched-MPC solver library (lib<solverName>.so).
 *
 * Every solver directory written by MpcModel.generateSolver() holds one shared library that exports exactly
 * these symbols, specialised for that model/horizon.  It is the drop-in for the FORCESPRO-generated
 * solver the reference loads with `forcespro.nlp.Solver.from_directory(dir)` and calls with
 * `solver.solve(problem)`:
 *     /root/reference/robotmpcs/planner/mpcPlanner.py:73   (load)
 *     /root/reference/robotmpcs/planner/mpcPlanner.py:246-250,262  (problem = {xinit, x0, all_parameters}; solve)
 * with one change of shape: a call solves B independent problems.  Array layouts are the reference's own,
 * with a leading batch dimension:
 *     xinit  [B][nx]            problem["xinit"]            float64
 *     x0     [B][N][nvar]       problem["x0"]               float64   (initial guess, stage-major)
 *     params [B][N][npar]       problem["all_parameters"]   float64   (stage-major, paramMap.yaml indices)
 *     z_out  [B][N][nvar]       output["x01".."xNN"]        float64
 *     action [B][nu]            what mpcPlanner.py:275-281 slices out of the output
 * Plain pointers and sizes only; no torch types.  All functions return 0 on success, negative on API error
 * (message via rmpc_last_error()).  Per-problem solver status is reported only through info.exitflag[]
 * with the FORCESPRO sign convention the reference tests (mpcPlanner.py:263): 1 optimal, 0 iteration limit,
 * <0 failure (-5 factorisation, -6 NaN/Inf — in the iteration, or in xinit / x0 / params (then no iteration is run: iters 0) —, -7 no progress / locally infeasible, -8 initial guess on the wrong side
 * of a ConstraintAvoidance pole, -9 the converged point touches or leaves the -100 <= z <= 100 box of the reference's NLP
 * (models/mpcModel.py:23-28,98-104), which this solver checks but does not carry as rows: not a solution of that NLP;
 * -1 the problem was never written: an internal error).
 *
 * Threading: a handle may be driven from several host threads and on any mix of streams; the host side of every entry
 * point is serialised per handle by a mutex, the launches themselves are asynchronous.  Up to 8 batched solves of one
 * handle can be in flight at a time on different streams (each owns one of 8 slots of queue heads and problem lists);
 * the ninth waits, in stream order and without blocking the host, for the first to finish.  Every entry point runs on
 * the handle's device and restores the calling thread's current device before it returns.
 */
#ifndef RMPC_H
#define RMPC_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rmpc_solver rmpc_solver;

typedef struct {
  int N, nx, nu, nvar, npar, nh;
  double dt;
} rmpc_dims;

typedef struct {
  int max_iter;     /* iteration cap, default 200 (FORCESPRO default) */
  int hessian;      /* 0 Gauss-Newton, 1 exact Lagrangian Hessian with Gauss-Newton fallback */
  int precision;    /* 0 (default, the parity mode: the reference's generated code is all-double): float64 throughout.
                       1: mixed — batches of more than 32 problems iterate in float32 while the barrier parameter is large
                       (down to mu = 1e-4, loose tolerances), then a float64 kernel continues from that state to the
                       tolerances below; a problem float32 cannot handle is solved in float64 from the caller's guess.
                       Same exit flags and tolerances as 0; may stop in another local minimum of a non-convex problem */
  int control_mode; /* 0 "acc": action = u of stage 1; 1 "vel": action = z_2[nvar-2nu : nvar-nu] (mpcPlanner.py:275-278) */
  double tol_stat, tol_eq, tol_ineq, tol_comp;
  double mu0, mu_min, reg;
  int creep_trigger; /* consecutive iterations whose first trial step left the interior of a nonlinear row before the
                        barrier parameter is raised again (x10, at most mu0); 0 = monotone barrier only.  Default 3 */
  int stall_window;  /* exit with -7 (no progress) when the optimality error has not dropped tenfold within this many
                        iterations; 0 = never (default): only max_iter bounds the work of a problem */
} rmpc_options;

/* Per-problem diagnostics, structure of arrays of length B.  Device pointers for rmpc_solve_batch, host
 * pointers for rmpc_solve_batch_host.  Any member may be NULL. */
typedef struct {
  int* exitflag;
  int* iters;
  double* obj;
  double* res_stat;
  double* res_eq;
  double* res_ineq;
  double* res_comp;
} rmpc_info;

const char* rmpc_model_name(void);
int rmpc_get_dims(rmpc_dims* out);
void rmpc_default_options(rmpc_options* out);
const char* rmpc_last_error(void);

/* Allocates the device workspace for up to max_batch problems on `device`. */
rmpc_solver* rmpc_create(int device, int max_batch);
void rmpc_destroy(rmpc_solver* s);
int rmpc_set_options(rmpc_solver* s, const rmpc_options* opt);
int rmpc_get_options(const rmpc_solver* s, rmpc_options* out);

/* Solve B problems.  All array arguments are DEVICE pointers owned by the caller; the call is asynchronous
 * on `stream` (a cudaStream_t passed as void*, NULL = default stream).  z_out / action_out / info may be NULL; z_out may
 * alias x0 (every problem's guess is read by the one kernel that then writes its plan).
 * Launches: a 16-byte memset of the call's queue slot, classify_kernel (every problem is classified once: does a row of
 * its initial guess start infeasible? -> two index lists; and do all its stages carry the same parameter block? -> the solve then
 * reads stage 1's block for every stage, one address per warp instead of N, with identical results), then two persistent solve kernels side by side (one warp per
 * problem, each pulling from its own list): feasible starts on `stream`, infeasible starts on a side stream forked from
 * and joined to it.  B <= 32 takes a single launch that classifies inside the solve. */
int rmpc_solve_batch(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params,
                     double* z_out, double* action_out, const rmpc_info* info, void* stream);

/* Same with the parameter blocks addressed by strides (in doubles): problem b, stage k reads
 * params[b * params_batch_stride + k * params_stage_stride + j], j < npar.  The reference layout all_parameters[N * npar]
 * (mpcPlanner.py:94,249) is (N * npar, npar).  params_stage_stride = 0 with params[B][npar] is the compressed form for
 * scenarios whose parameters do not vary over the horizon (SURVEY.md 8f-2, B4 "or compressed"): every set* of the
 * reference except per-stage setLinearConstraints and updateDynamicObstacles writes the same values to all N stages
 * (mpcPlanner.py:120-210); the lanes of a warp then read one address instead of N, and the batch shrinks N-fold. */
int rmpc_solve_batch_strided(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params,
                             long params_batch_stride, long params_stage_stride, double* z_out, double* action_out,
                             const rmpc_info* info, void* stream);

/* Same with HOST pointers: copies inputs to the device, solves, copies results back, synchronises. */
int rmpc_solve_batch_host(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params,
                          double* z_out, double* action_out, const rmpc_info* info);

/* The same without the final synchronisation: enqueues H2D copies, the solve and the D2H copies on the handle's private
 * stream and returns.  The host buffers must stay valid (and should be page-locked, otherwise the copies block) until
 * rmpc_wait(s) returns.  One call in flight per handle; use two handles to overlap the copies of one batch with the
 * solve of another (bench.py's end-to-end leg does). */
int rmpc_solve_batch_host_async(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params,
                                double* z_out, double* action_out, const rmpc_info* info);
int rmpc_wait(rmpc_solver* s);

/* Warm start for `initialization: previous_plan` (mpcPlanner.py:215-226 shiftHorizon): x0[b][k] = z[b][k+1] for
 * k < N-1, x0[b][N-1] = z[b][N-1].  Device pointers, [B][N][nvar] each; in-place (x0 == z) is allowed (one thread walks
 * the horizon of one variable, no scratch copy). */
int rmpc_shift_horizon(rmpc_solver* s, int B, const double* z, double* x0, void* stream);

/* Initial guess for `initialization: current_state` (mpcPlanner.py:228-232 setX0): x0[b][k][0:nx] = xinit[b] for every
 * stage k; the input part of x0 is left as it is.  Device pointers. */
int rmpc_set_x0(rmpc_solver* s, int B, const double* xinit, double* x0, void* stream);

/* One step of the plant for a headless closed loop: x_next[b] = Phi(x[b], u[b]), the discrete model the solver was generated
 * from (models/mpcModel.py:65-69,118-120 and diff_drive_mpc_model.py:24-41: ERK2 with 5 nodes, in closed form).  It stands where
 * the reference's examples call env.step(action) of the pybullet simulation (examples/boxer_example.py:205; unavailable on this
 * box).  x[B][nx], u[B][nu] (the input; for control_mode "vel" the caller converts the velocity action first), exitflag[B] may be
 * NULL: a problem with exitflag < 0 is stepped with u = 0, as the reference applies no action after a failed solve
 * (boxer_example.py:194).  x_next may alias x.  Device pointers, asynchronous on `stream`. */
int rmpc_plant_step(rmpc_solver* s, int B, const double* x, const double* u, const int* exitflag, double* x_next, void* stream);

/* Batched parameter scatter (the set* loops of mpcPlanner.py:120-210): for every problem b and every stage k,
 * params[b][k][idx[j]] = values[b][j] (stage_varying == 0, values is [B][n_idx]) or values[b][k][j]
 * (stage_varying == 1, values is [B][N][n_idx]).  idx is a HOST array of paramMap indices (it travels by value with
 * the launch); the rest are device pointers. */
int rmpc_set_params(rmpc_solver* s, int B, double* params, const int* idx, int n_idx, const double* values,
                    int stage_varying, void* stream);

/* Constant-acceleration prediction of moving obstacles over the horizon (planner/mpcPlanner.py:144-161
 * updateDynamicObstacles): obst_state[B][n_dyn][9] = (position, velocity, acceleration) of each moving obstacle; for every
 * problem, stage k and obstacle slot j < n_slots (= number_obstacles) the parameters
 * params[b][k][p_obst + 4 j .. + 3] become (pos + vel (k dt) + acc (k dt)^2 / 2, radius) for j < n_dyn and
 * (-100, -100, -100, radius) otherwise (mpcPlanner.py:152-154,158-160; `radius` is the reference's self._r).  p_obst is
 * paramMap["obst"][0].  Device pointers, asynchronous on `stream`. */
int rmpc_predict_obstacles(rmpc_solver* s, int B, const double* obst_state, int n_dyn, int n_slots, int p_obst, double radius,
                           double* params, void* stream);

/* A batched solve that allocates nothing.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arrays, layouts and results as rmpc_solve_batch (device pointers owned by the caller, asynchronous on `stream`), but
 * there is no handle: the scratch memory of the solve (queue heads + the two problem lists of the classification pass)
 * is a file-scope device array of 131 328 bytes inside the library, held by every GPU that loads it until the process
 * exits, and the call performs no cudaMalloc / cudaFree / stream or event creation.  Limits and rules:
 *   - B <= rmpc_static_max_batch() = 16384; a larger B is refused with an error (rmpc_last_error);
 *   - runs on the calling thread's CURRENT device; the first call on a device does a one-time set-up (symbol address,
 *     kernel attributes, occupancy query — no allocation), so make that first call outside graph capture;
 *   - every call shares the one array (it is cleared at the start of each call, in stream order): calls that overlap on
 *     the GPU — two streams, or two host threads on two streams — race and give wrong results; serialise them;
 *   - the three kernels run one after the other on `stream` (no side stream), so a batch with many infeasible starts
 *     is slower than through a handle;
 *   - opt may be NULL (defaults); precision must be 0. */
int rmpc_solve_batch_static(int B, const double* xinit, const double* x0, const double* params, double* z_out,
                            double* action_out, const rmpc_info* info, const rmpc_options* opt, void* stream);
int rmpc_static_max_batch(void);

/* ---- Result exchange between the ranks of one node over peer memory (NVLink / NVSwitch) --------------------------------
 * No counterpart in the reference (a single process, one solve at a time).  The scenario batch is sharded over one process
 * per GPU and the solve needs no collective; these calls hand every rank the whole batch's (action, exitflag, iterations)
 * of a step without NCCL: each rank owns one device region which its peers map through CUDA IPC, a small kernel behind the
 * rank's solve stores its rows into every rank's region (posted peer stores) and raises one flag per (step, rank), and the
 * collecting side waits for the world's flags with ONE resident warp, unpacks, and acknowledges the ring slot
 * (csrc/rmpc_exchange.cuh states the protocol and why it cannot deadlock).  Rules:
 *   - publish is called in the same order on every rank (it numbers the steps), collect in step order after its publish;
 *   - at most `depth` steps may be published ahead of the slowest rank's collects: a publish beyond that waits on the device;
 *   - a wait that lasts longer than 20 s gives up and raises the status word (rmpc_exchange_status != 0) instead of
 *     hanging the GPU; results after that are undefined;
 *   - rows_per_rank >= every rank's shard; row i of the collected arrays is row i % rows_per_rank of rank i / rows_per_rank.
 * rmpc_ipc_alloc returns zeroed device memory and its 64-byte cudaIpcMemHandle_t, rmpc_ipc_open maps a peer's handle into
 * this process (same node, peer access enabled lazily); regions[w] of rmpc_exchange_create is rank w's region as seen from
 * this process (regions[rank] is the rank's own allocation).  Inside one process the regions can be passed around as plain
 * pointers, which is how tests/test_gpu_exchange.py runs two ranks on one GPU. */
typedef struct rmpc_exchange rmpc_exchange;
size_t rmpc_exchange_region_bytes(int world, int rows_per_rank, int depth);   /* 0 = bad argument */
int rmpc_ipc_alloc(int device, size_t bytes, void** ptr, unsigned char* handle64 /* may be NULL */);
int rmpc_ipc_open(int device, const unsigned char* handle64, void** ptr);
int rmpc_ipc_close(int device, void* ptr);
int rmpc_ipc_free(int device, void* ptr);
rmpc_exchange* rmpc_exchange_create(int device, int world, int rank, int rows_per_rank, int depth, void* const* regions);
void rmpc_exchange_destroy(rmpc_exchange* x);
/* Stores this rank's B rows action[B][nu], exitflag[B], iters[B] (device pointers; exitflag / iters may be NULL) into every
 * rank's region, on the exchange's own stream behind everything `stream` holds at the time of the call (the solve that
 * produces them); `stream` itself is never held.  The arrays must stay valid until the step has been collected.  Returns
 * the step number, -1 on error. */
long rmpc_exchange_publish(rmpc_exchange* x, int B, const double* action, const int* exitflag, const int* iters, void* stream);
/* Step `seq` of every rank into action_all[world * rows_per_rank][nu], exitflag_all[...], iters_all[...] (device pointers, the
 * int arrays may be NULL).  The wait for the slowest rank and the unpack run on the exchange's own stream, the unpack behind
 * everything `stream` holds at the time of the call (so the output arrays may be fresh from a stream-ordered allocator);
 * `stream` itself is not held.  rmpc_exchange_join makes `stream` wait for every collect enqueued so far. */
int rmpc_exchange_collect(rmpc_exchange* x, long seq, double* action_all, int* exitflag_all, int* iters_all, void* stream);
int rmpc_exchange_join(rmpc_exchange* x, void* stream);
int rmpc_exchange_status(rmpc_exchange* x);   /* 0 ok, 1 a publish timed out waiting for a slot, 2 a collect timed out; synchronises */

/* Number of CUDA kernels this library has launched since it was loaded (bench.py reports it). */
long rmpc_launch_count(void);

/* Batched free-space decomposition (robotmpcs/utils/free_space_decomposition.py:61-116 as driven by
 * examples/boxer_example.py:132-148,191-203): for every problem b and stage k the seed is the lidar position of the
 * reference state ref[b*ref_batch_stride + k*ref_stage_stride + {0,1,2}] = (x, y, theta), i.e. (x, y) + Rot(theta)
 * (lidar_offset, 0) at height lidar_z; up to K half-planes [n0 n1 n2 c] are grown greedily from the nearest points of
 * points[b][n_points][3] within max_radius, unused slots get the reference's dummy plane.  Results go to
 * out[b][k][K][4] (may be NULL) and/or into params[b][k][idx[c] .. idx[c]+3] (may be NULL; idx is a HOST array of the K
 * `lin_constrs_c` parameter indices).  ref may point into a previous plan z (strides N*nvar, nvar) or at one state per
 * problem (strides 3, 0).  Device pointers, asynchronous on `stream`.  K <= 16, n_points <= 128. */
int rmpc_fsd_constraints(rmpc_solver* s, int B, const double* points, int n_points, const double* ref, long ref_batch_stride,
                         long ref_stage_stride, double max_radius, double lidar_offset, double lidar_z, int K, double* out,
                         double* params, const int* idx, void* stream);

#ifdef __cplusplus
}
#endif
#endif
