// rmpc_solver.cu — CUDA kernels (sm_100a) + C ABI (include/rmpc.h) of one generated solver library.
// Compiled once per model against its generated model_traits.h by MpcModel.generateSolver().
//
// Kernels
//   solve_kernel   one warp = one NLP (rmpc_ipm.cuh, G = 32), iterate in shared memory, persistent grid + work queue;
//                  reads the reference layout directly and writes z, action and the info arrays itself
//   shift_kernel   previous-plan warm start;  scatter_kernel  batched set* parameter writes
// There is no CPU fallback: every entry point fails loudly if CUDA fails.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <atomic>
#include <mutex>
#include <type_traits>

#include "model_traits.h"
#ifndef RMPC_WARPS_FIXED
#define RMPC_WARPS_FIXED 1
#endif
#define RMPC_WARPS RMPC_WARPS_FIXED
#include "rmpc_ipm.cuh"
#include "rmpc_fsd.cuh"
#include "rmpc_exchange.cuh"
#include "rmpc.h"

using M = ModelTraits;
using namespace rmpc;

static thread_local char g_err[512] = "";
static std::atomic<long> g_launches{0};

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      snprintf(g_err, sizeof g_err, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      return -1;                                                                                   \
    }                                                                                              \
  } while (0)

// Every entry point runs on the handle's device and leaves the calling thread's current device as it found it.
struct DeviceGuard {
  int prev = -1;
  cudaError_t err;
  explicit DeviceGuard(int dev) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != dev) err = cudaSetDevice(dev);
    else if (err == cudaSuccess) prev = -1;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
#define GUARD(s) DeviceGuard guard_((s)->device); CK(guard_.err)

// Calls in flight per handle (on any mix of streams): each call owns one slot = queue heads + the two problem lists of
// its batch + a side stream; a slot is reused only after the call that used it last has finished (stream-ordered wait
// on the slot's event, no host blocking).
constexpr int kSlots = 8;
constexpr int kSmallBatch = 32;     // up to here a batch goes through the single-launch kernel (no classification pass)
constexpr int kMaxIdx = 128;        // parameter indices per rmpc_set_params launch (passed by value)

struct Slot {
  int* q;               // device: [0] feasible count, [1] shifted count, [2] feasible head, [3] shifted head
  int* list;            // device: [max_batch] feasible problems, [max_batch] shifted problems
  cudaStream_t side;    // the kernel for shifted starts runs here, forked from and joined to the caller's stream
  cudaEvent_t ev_fork, ev_join, ev_done;
  bool used;
  // mixed precision (options.precision = 1): what the float32 kernel hands to the float64 kernel, per problem of the batch
  float *hz, *hlam, *hmu;   // z [N][nvar], lambda [N][nh], mu
  int* hit;                 // iterations done in float32, -1 = solve from the caller's guess in float64
};

struct LatencyGraph {   // captured H2D -> single-launch solve -> D2H for one small batch size (rmpc_solve_batch_host)
  cudaGraphExec_t exec;
  int B, flags;
};

struct rmpc_solver {
  int device;
  int max_batch;
  rmpc_options opt;
  std::mutex mu;        // serialises the host side of every entry point of this handle (launches are asynchronous)
  Slot slot[kSlots];
  int next_slot;
  int blocks_per_sm, blocks_per_sm_sh, blocks_per_sm_all, sm_count;
  int bps_lowp, bps_resume;   // resident blocks per SM of the two kernels of a mixed-precision solve
  size_t smem_bytes;  // dynamic shared memory per block
  size_t smem_bytes_all;   // ... of the single-launch kernel for small batches (state in shared memory for every model)
  // device staging for the host entry points: one packed input block and one packed output block
  double* d_in;         // xinit | x0 | params
  double* d_out;        // z | action | obj res_stat res_eq res_ineq res_comp | exitflag iters (ints)
  double *p_in, *p_out; // page-locked mirrors
  bool staged;
  cudaStream_t hstream;  // private stream of the host entry points
  int pend_B;
  double *u_z, *u_action;
  rmpc_info u_info;
  LatencyGraph lg;
  int* lat_q;           // queue head of the captured small-batch launch
};
// ---------------------------------------------------------------------------------------------
// One warp = one NLP (rmpc_ipm.cuh, G = 32): lane k owns stage k in the stage-parallel phases, the 32 lanes share the
// per-stage dense algebra of the Riccati sweep through the warp's slice of shared memory.  Warps pull problems from a
// device-side queue, so a hard problem (many iterations) only delays its own warp.  Inputs and outputs are read and
// written in the reference's own layout ([B][N][nvar], [B][N][npar]); there is no separate pack/unpack pass.
using Shared = ProblemShared<M, double>;
using I32 = Ipm<M, double, 32>;
constexpr size_t kSharedPerWarp = (sizeof(Shared) + 15) / 16 * 16;
// The single-launch kernel for batches of <= 32 problems (the latency path) keeps multipliers and slacks in shared memory for
// every model: with at most 32 problems on 148 SMs occupancy does not matter, local-memory round trips do (r3s A/B, N = 40:
// B = 1 p50 1.28 -> 0.99 ms, p99 4.98 -> 3.81 ms, while the same form costs a full batch 5 % of its throughput).
struct LatencyTraits : ModelTraits { static constexpr bool STATE_IN_SHARED = true; };
template <int MODE> using ModelOf = typename std::conditional<MODE == I32::kAll, LatencyTraits, M>::type;
using SharedL = ProblemShared<LatencyTraits, double>;
constexpr size_t kSharedLPerWarp = (sizeof(SharedL) + 15) / 16 * 16;
// Warps (= problems) per thread block.  With more than one, the warps of a block meet at the top of every solver
// iteration (Group<32>::align) so that they walk through the same code together and share instruction-cache fills.
// Measured on B200 (boxer, 20 steps): 1 warp/block 702 k solves/s, 2: 622 k, 3: 574 k, 7: 534 k — waiting for the
// slowest iteration of the block costs more than the shared fetches save, so the default stays at 1.
constexpr int kWarps = RMPC_WARPS_FIXED;
static_assert(kWarps >= 1, "one problem does not fit into the shared memory of an SM");

#ifndef RMPC_MIN_BLOCKS
#define RMPC_MIN_BLOCKS 1   // resident blocks per SM the register allocation must allow
#endif
// Bit of a problem-list entry: every stage of the problem carries the same parameter block (true for every set* of the reference
// except per-stage setLinearConstraints and updateDynamicObstacles, mpcPlanner.py:120-210).  The solve then reads stage 1's block
// for all stages: the lanes of a warp load one address instead of N, one L1 line instead of N (what rmpc_solve_batch_strided with
// stage stride 0 gives a caller who knows it: boxer +4 %), with bit-identical results.
constexpr int kStageInvariant = 1 << 30;
struct Handoff { const float *z, *lam, *mu; const int* it; };
struct SolveArgs {
  Handoff ho;        // mixed precision only: state left by the float32 pass
  const double *xinit, *x0, *params;
  long p_bs, p_ss;   // elements between the parameter blocks of consecutive problems / stages (reference layout: N*npar, npar)
  double *z_out, *action;
  rmpc_info info;
  Options opt;
  int control_mode;
};

template <class SharedT>
__device__ __forceinline__ void write_outputs(const SolveArgs& a, const SharedT& S, int lane, int b, const Info& inf) {
  if (a.action && lane < M::NU) {
    // "acc": u of stage 1; "vel": z_2[nvar-2nu : nvar-nu]  (planner/mpcPlanner.py:275-281)
    a.action[(size_t)b * M::NU + lane] =
        a.control_mode == 0 ? S.st[0].z[M::NX + lane] : S.st[1].z[M::NVAR - 2 * M::NU + lane];
  }
  if (lane == 0) {
    if (a.info.exitflag) a.info.exitflag[b] = inf.exitflag;
    if (a.info.iters) a.info.iters[b] = inf.iters;
    if (a.info.obj) a.info.obj[b] = inf.obj;
    if (a.info.res_stat) a.info.res_stat[b] = inf.res_stat;
    if (a.info.res_eq) a.info.res_eq[b] = inf.res_eq;
    if (a.info.res_ineq) a.info.res_ineq[b] = inf.res_ineq;
    if (a.info.res_comp) a.info.res_comp[b] = inf.res_comp;
  }
}

// Pass 1 of a batched solve: every problem is classified ONCE — does a row of its initial guess start below delta? —
// and appended to the list of the kernel that will solve it.  One warp per problem (lane = stage, as in the solve), no
// shared memory.  Everything after this pass reads a problem's inputs only from the kernel that owns it, and before that
// kernel writes the problem's outputs, so z_out may alias x0.
__global__ void __launch_bounds__(256) classify_kernel(int B, const __grid_constant__ SolveArgs a, int* __restrict__ q,
                                                       int* __restrict__ list, int max_batch) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < B; b += warps) {
    const ParamView<double> pv{a.params + (size_t)b * a.p_bs, a.p_ss, 1};
    bool shifted = false, varies = false;
#pragma unroll
    for (int sl = 0; sl < I32::SLOTS; ++sl) {
      const int k = lane + 32 * sl;
      if (k < M::N) {
        double z[M::NVAR], p[M::NPAR];
        const double* zin = a.x0 + ((size_t)b * M::N + k) * M::NVAR;
#pragma unroll
        for (int i = 0; i < M::NVAR; ++i) z[i] = zin[i];
        if (k == 0) {
#pragma unroll
          for (int i = 0; i < M::NX; ++i) z[i] = a.xinit[(size_t)b * M::NX + i];
        }
        load_params<M, double>(pv, k, p);
        // do this stage's parameters differ from stage 1's?  (every lane reads the same address: one transaction per parameter)
#pragma unroll
        for (int j = 0; j < M::NPAR; ++j) varies = varies || !(p[j] == pv.get(0, j));
        StageVal<M, double> sv;
        stage_values<M, double>(z, p, sv);
#pragma unroll
        for (int r = 0; r < M::NH; ++r) shifted = shifted || !(sv.h[r] >= a.opt.delta);
      }
    }
    shifted = __any_sync(0xffffffffu, shifted) != 0;
    varies = __any_sync(0xffffffffu, varies) != 0;
    if (lane == 0) {
      const int at = atomicAdd(q + (shifted ? 1 : 0), 1);
      list[(shifted ? max_batch : 0) + at] = b | (varies ? 0 : kStageInvariant);
    }
  }
}

// Pass 2, two instantiations launched side by side on two streams.  kFeasible is the common case and contains no code for
// rows that start infeasible; kShifted carries both copies of the iteration and solves the problems of the second list.
// With the rare path in the same kernel as the hot one the solver was 4-10 % slower (inlined) or 25 % slower (as a called
// function); B200, r1l A/B in DESIGN.md.  kAll (list == nullptr) is the single-launch form for small batches: it
// classifies inside the solve and takes problems 0 .. B-1 straight from the queue head.
template <int MODE>
__global__ void __launch_bounds__(32 * kWarps, RMPC_MIN_BLOCKS)
solve_kernel(int B, const __grid_constant__ SolveArgs a, const int* __restrict__ list, const int* __restrict__ count,
             int* __restrict__ head) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5;
  using MM = ModelOf<MODE>;
  using SharedM = ProblemShared<MM, double>;
  using IpmM = Ipm<MM, double, 32>;
  SharedM& S = *reinterpret_cast<SharedM*>(smem_raw + warp * ((sizeof(SharedM) + 15) / 16 * 16));
  Group<32> g{int(threadIdx.x & 31)};
  const int n = MODE == I32::kAll ? B : *count;
  for (;;) {
    int b = 0;
    if (g.lane == 0) {
      b = atomicAdd(head, 1);
      if (MODE != I32::kAll) b = b < n ? list[b] : B;   // kFeasible, kShifted, kResume: the problems of a list
    }
    b = __shfl_sync(0xffffffffu, b, 0);
    const bool invariant = MODE != I32::kAll && (b & kStageInvariant) != 0;
    if (MODE != I32::kAll) b &= ~kStageInvariant;
    if (b >= B) {
      // out of work: keep meeting the other warps of the block at their iteration rendezvous until all are out of work
      if (kWarps > 1) while (!__syncthreads_and(1)) {}
      break;
    }
    const ParamView<double> pv{a.params + (size_t)b * a.p_bs, invariant ? 0 : a.p_ss, 1};
    Info inf;
    {
      IpmM ipm(g, S, pv, a.opt);
      if constexpr (MODE == I32::kResume)
        ipm.solve_resume(a.xinit + (size_t)b * M::NX, a.x0 + (size_t)b * M::N * M::NVAR, a.z_out ? a.z_out + (size_t)b * M::N * M::NVAR : nullptr,
                         a.ho.z + (size_t)b * M::N * M::NVAR, a.ho.lam + (size_t)b * M::N * M::NH, a.ho.mu[b], a.ho.it[b], inf);
      else
        ipm.template solve<MODE>(a.xinit + (size_t)b * M::NX, a.x0 + (size_t)b * M::N * M::NVAR,
                                 a.z_out ? a.z_out + (size_t)b * M::N * M::NVAR : nullptr, inf);
    }
    write_outputs(a, S, g.lane, b, inf);
    __syncwarp();
  }
}

// Mixed precision, pass 1: the iteration in float32 (rmpc_ipm.cuh solve_lowp) on the feasible list, while the barrier
// parameter is large; the state goes to global memory, and solve_kernel<kResume> (float64) finishes the solve.  Half the
// shared memory and fewer registers than the float64 kernel: more problems per SM, and shorter dependent chains.
#ifndef RMPC_LOWP_MIN_BLOCKS
#define RMPC_LOWP_MIN_BLOCKS 10
#endif
using SharedF = ProblemShared<M, float>;
constexpr size_t kSharedFPerWarp = (sizeof(SharedF) + 15) / 16 * 16;
__global__ void __launch_bounds__(32, RMPC_LOWP_MIN_BLOCKS)
lowp_kernel(int B, const __grid_constant__ SolveArgs a, const int* __restrict__ list, const int* __restrict__ count, int* __restrict__ head,
            float* __restrict__ hz, float* __restrict__ hlam, float* __restrict__ hmu, int* __restrict__ hit) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SharedF& S = *reinterpret_cast<SharedF*>(smem_raw);
  Group<32> g{int(threadIdx.x & 31)};
  using IF = Ipm<M, float, 32, double>;
  const Options lo = IF::lowp_options(a.opt);
  const int n = *count;
  for (;;) {
    int b = 0;
    if (g.lane == 0) { b = atomicAdd(head, 1); b = b < n ? list[b] : B; }
    b = __shfl_sync(0xffffffffu, b, 0);
    const bool invariant = (b & kStageInvariant) != 0;
    b &= ~kStageInvariant;
    if (b >= B) break;
    const ParamView<float, double> pv{a.params + (size_t)b * a.p_bs, invariant ? 0 : a.p_ss, 1};
    {
      IF ipm(g, S, pv, lo);
      ipm.solve_lowp(a.xinit + (size_t)b * M::NX, a.x0 + (size_t)b * M::N * M::NVAR, hz + (size_t)b * M::N * M::NVAR,
                     hlam + (size_t)b * M::N * M::NH, hmu + b, hit + b);
    }
    __syncwarp();
  }
}

// previous-plan warm start, safe in place: one thread per (problem, variable) walks the horizon, reading stage k+1 before
// anything writes it
__global__ void shift_kernel(long total, const double* __restrict__ z, double* __restrict__ x0) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const long b = i / M::NVAR;
  const int v = int(i - b * M::NVAR);
  const double* src = z + b * M::N * M::NVAR + v;
  double* dst = x0 + b * M::N * M::NVAR + v;
  double nxt = src[0];
#pragma unroll 4
  for (int k = 0; k < M::N - 1; ++k) {
    nxt = src[(k + 1) * M::NVAR];
    dst[k * M::NVAR] = nxt;
  }
  dst[(M::N - 1) * M::NVAR] = nxt;
}

// "current_state" initial guess (planner/mpcPlanner.py:228-232): the state part of every stage <- xinit, inputs untouched
__global__ void guess_kernel(long total, const double* __restrict__ xinit, double* __restrict__ x0) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int v = int(i % M::NX);
  const long bk = i / M::NX;
  x0[bk * M::NVAR + v] = xinit[(bk / M::N) * M::NX + v];
}

// x+ = Phi(x, u) of the model the solver was generated from (rmpc_model.cuh dyn_value), one thread per problem; a problem whose
// solve failed (exitflag < 0) gets u = 0, as the reference's loops do not apply a failed solve (examples/boxer_example.py:194)
__global__ void plant_kernel(int B, const double* __restrict__ x, const double* __restrict__ u, const int* __restrict__ exitflag,
                             double* __restrict__ x_next) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double z[M::NVAR], xn[M::NX];
#pragma unroll
  for (int i = 0; i < M::NX; ++i) z[i] = x[(size_t)b * M::NX + i];
  const bool apply = !exitflag || exitflag[b] >= 0;
#pragma unroll
  for (int i = 0; i < M::NU; ++i) z[M::NX + i] = apply ? u[(size_t)b * M::NU + i] : 0.0;
#pragma unroll
  for (int i = 0; i < M::NVAR - M::NX - M::NU; ++i) z[M::NX + M::NU + i] = 0.0;
  dyn_value<M, double>(z, xn);
#pragma unroll
  for (int i = 0; i < M::NX; ++i) x_next[(size_t)b * M::NX + i] = xn[i];
}

struct IdxList { int n; int v[kMaxIdx]; };

__global__ void scatter_kernel(int B, int n_all, int off, int stage_varying, const __grid_constant__ IdxList idx,
                               const double* __restrict__ values, double* __restrict__ params) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long total = (long)B * M::N * idx.n;
  if (i >= total) return;
  const int j = int(i % idx.n);
  const long bk = i / idx.n;
  const long b = bk / M::N;
  const double v = stage_varying ? values[bk * n_all + off + j] : values[b * n_all + off + j];
  params[bk * M::NPAR + idx.v[j]] = v;
}

// Constant-acceleration prediction of the moving obstacles over the horizon, written into the stage-varying `obst`
// parameter slots (planner/mpcPlanner.py:144-161): stage i sees pos + vel (i dt) + acc (i dt)^2 / 2, radius r; obstacle
// slots without a moving obstacle are parked at -100 with the same radius.  One thread per (problem, stage, slot).
__global__ void predict_kernel(int B, int n_dyn, int K, int p_obst, double radius, const double* __restrict__ state,
                               double* __restrict__ params) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)B * M::N * K) return;
  const int j = int(i % K);
  const long bk = i / K;
  const int k = int(bk % M::N);
  const long b = bk / M::N;
  double* o = params + bk * M::NPAR + p_obst + 4 * j;
  const double t = __dmul_rn(M::DT, double(k));
  if (j < n_dyn) {
    const double* st = state + ((size_t)b * n_dyn + j) * 9;
#pragma unroll
    for (int c = 0; c < 3; ++c)   // the operation order of mpcPlanner.py:158, without fused multiply-adds: the same bits as the Python line
      o[c] = __dadd_rn(__dadd_rn(st[c], __dmul_rn(__dmul_rn(st[3 + c], M::DT), double(k))), __dmul_rn(__dmul_rn(0.5, __dmul_rn(t, t)), st[6 + c]));
  } else {
    o[0] = o[1] = o[2] = -100.0;
  }
  o[3] = radius;
}

// ---------------------------------------------------------------------------------------------
static Options to_core(const rmpc_options& o) {
  Options c = default_options();
  c.max_iter = o.max_iter; c.hessian = o.hessian;
  c.tol_stat = o.tol_stat; c.tol_eq = o.tol_eq; c.tol_ineq = o.tol_ineq; c.tol_comp = o.tol_comp;
  c.mu0 = o.mu0; c.mu_min = o.mu_min; c.reg = o.reg;
  c.creep_trigger = o.creep_trigger; c.stall_window = o.stall_window;
  return c;
}

extern "C" {

const char* rmpc_model_name(void) { return M::NAME; }
const char* rmpc_last_error(void) { return g_err; }
long rmpc_launch_count(void) { return g_launches.load(); }

int rmpc_get_dims(rmpc_dims* d) {
  if (!d) return -1;
  d->N = M::N; d->nx = M::NX; d->nu = M::NU; d->nvar = M::NVAR; d->npar = M::NPAR; d->nh = M::NH; d->dt = M::DT;
  return 0;
}

void rmpc_default_options(rmpc_options* o) {
  const Options c = default_options();
  o->max_iter = c.max_iter; o->hessian = c.hessian; o->precision = 0; o->control_mode = 0;
  o->tol_stat = c.tol_stat; o->tol_eq = c.tol_eq; o->tol_ineq = c.tol_ineq; o->tol_comp = c.tol_comp;
  o->mu0 = c.mu0; o->mu_min = c.mu_min; o->reg = c.reg;
  o->creep_trigger = int(c.creep_trigger); o->stall_window = int(c.stall_window);
}

static void free_all(rmpc_solver* s) {
  for (int i = 0; i < kSlots; ++i) {
    Slot& sl = s->slot[i];
    cudaFree(sl.q); cudaFree(sl.list); cudaFree(sl.hz); cudaFree(sl.hlam); cudaFree(sl.hmu); cudaFree(sl.hit);
    if (sl.side) cudaStreamDestroy(sl.side);
    if (sl.ev_fork) cudaEventDestroy(sl.ev_fork);
    if (sl.ev_join) cudaEventDestroy(sl.ev_join);
    if (sl.ev_done) cudaEventDestroy(sl.ev_done);
  }
  if (s->lg.exec) cudaGraphExecDestroy(s->lg.exec);
  cudaFree(s->d_in); cudaFree(s->d_out); cudaFree(s->lat_q);
  cudaFreeHost(s->p_in); cudaFreeHost(s->p_out);
  if (s->hstream) cudaStreamDestroy(s->hstream);
}

rmpc_solver* rmpc_create(int device, int max_batch) {
  if (max_batch <= 0 || max_batch >= kStageInvariant) { snprintf(g_err, sizeof g_err, "max_batch must be in (0, 2^30)"); return nullptr; }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || device < 0 || device >= ndev) {
    snprintf(g_err, sizeof g_err, "no usable CUDA device %d (%s); this library has no CPU fallback", device,
             e == cudaSuccess ? "out of range" : cudaGetErrorString(e));
    return nullptr;
  }
  DeviceGuard guard(device);
  if (guard.err != cudaSuccess) { snprintf(g_err, sizeof g_err, "cudaSetDevice: %s", cudaGetErrorString(guard.err)); return nullptr; }
  rmpc_solver* s = new rmpc_solver();
  s->device = device;
  s->max_batch = max_batch;
  s->next_slot = 0; s->staged = false; s->pend_B = 0;
  s->d_in = s->d_out = s->p_in = s->p_out = nullptr; s->hstream = nullptr; s->lat_q = nullptr;
  s->u_z = s->u_action = nullptr;
  memset(&s->u_info, 0, sizeof s->u_info);
  memset(&s->lg, 0, sizeof s->lg);
  memset(s->slot, 0, sizeof s->slot);
  rmpc_default_options(&s->opt);
  auto fail = [&](const char* what, cudaError_t err) -> rmpc_solver* {
    snprintf(g_err, sizeof g_err, "%s: %s", what, cudaGetErrorString(err));
    free_all(s);
    delete s;
    return nullptr;
  };
  for (int i = 0; i < kSlots; ++i) {
    Slot& sl = s->slot[i];
    if ((e = cudaMalloc(&sl.q, sizeof(int) * 8)) != cudaSuccess) return fail("cudaMalloc", e);
    if ((e = cudaMalloc(&sl.list, sizeof(int) * 2 * (size_t)max_batch)) != cudaSuccess) return fail("cudaMalloc", e);
    if ((e = cudaStreamCreateWithFlags(&sl.side, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    if ((e = cudaEventCreateWithFlags(&sl.ev_fork, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&sl.ev_join, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&sl.ev_done, cudaEventDisableTiming)) != cudaSuccess)
      return fail("cudaEventCreate", e);
  }
  size_t smem = kSharedPerWarp * kWarps;
  if (const char* pad = getenv("RMPC_PAD_SMEM_KB")) smem += (size_t)atoi(pad) * 1024;  // experiment knob: fewer resident problems per SM
  s->smem_bytes = smem;
  const size_t smem_all = kSharedLPerWarp * kWarps;
  s->smem_bytes_all = smem_all;
  if ((e = cudaFuncSetAttribute(solve_kernel<I32::kFeasible>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(solve_kernel<I32::kShifted>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(solve_kernel<I32::kAll>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_all)) != cudaSuccess)
    return fail("cudaFuncSetAttribute(shared memory)", e);
  if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&s->blocks_per_sm, solve_kernel<I32::kFeasible>, 32 * kWarps, smem)) != cudaSuccess ||
      (e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&s->blocks_per_sm_sh, solve_kernel<I32::kShifted>, 32 * kWarps, smem)) != cudaSuccess ||
      (e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&s->blocks_per_sm_all, solve_kernel<I32::kAll>, 32 * kWarps, smem_all)) != cudaSuccess)
    return fail("cudaOccupancyMaxActiveBlocksPerMultiprocessor", e);
  if ((e = cudaFuncSetAttribute(solve_kernel<I32::kResume>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(lowp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSharedFPerWarp)) != cudaSuccess)
    return fail("cudaFuncSetAttribute(shared memory, mixed precision)", e);
  if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&s->bps_resume, solve_kernel<I32::kResume>, 32 * kWarps, smem)) != cudaSuccess ||
      (e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&s->bps_lowp, lowp_kernel, 32, kSharedFPerWarp)) != cudaSuccess)
    return fail("cudaOccupancyMaxActiveBlocksPerMultiprocessor(mixed precision)", e);
  // L1 against resident problems (r3c, profiles/r3c_l1_carveout_experiment.txt): the driver's default carve-out already leaves panda
  // 60 KB of L1 beside its four problems per SM; one problem fewer with 92 KB of L1 measures the same throughput (474 against 487 k)
  // and a slower lone batch (3.1 against 2.5 ms), and the other models lose 0-3 %.  The two knobs below stay for experiments.
  if (const char* pct = getenv("RMPC_CARVEOUT_PCT")) {  // experiment knob: shared-memory carve-out of the hot kernel in percent (the rest of the 256 KB is L1)
    if ((e = cudaFuncSetAttribute(solve_kernel<I32::kFeasible>, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(pct))) != cudaSuccess)
      return fail("cudaFuncSetAttribute(carve-out)", e);
  }
  if (const char* cap = getenv("RMPC_MAX_BLOCKS_PER_SM")) {  // experiment knob: fewer problems in flight per SM
    const int c = atoi(cap);
    if (c > 0 && c < s->blocks_per_sm) s->blocks_per_sm = c;
  }
  if (s->blocks_per_sm < 1 || s->blocks_per_sm_sh < 1 || s->blocks_per_sm_all < 1) {
    snprintf(g_err, sizeof g_err, "solve kernel does not fit on an SM (%zu B shared)", smem);
    free_all(s); delete s;
    return nullptr;
  }
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return fail("cudaGetDeviceProperties", e);
  s->sm_count = prop.multiProcessorCount;
  if ((e = cudaStreamCreateWithFlags(&s->hstream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
  return s;
}

void rmpc_destroy(rmpc_solver* s) {
  if (!s) return;
  {
    DeviceGuard guard(s->device);
    cudaDeviceSynchronize();
    free_all(s);
  }
  delete s;
}

int rmpc_set_options(rmpc_solver* s, const rmpc_options* o) {
  if (!s || !o) { snprintf(g_err, sizeof g_err, "null argument"); return -1; }
  if (o->precision < 0 || o->precision > 1 || o->hessian < 0 || o->hessian > 1 || o->control_mode < 0 || o->control_mode > 1 ||
      o->creep_trigger < 0 || o->stall_window < 0) {
    snprintf(g_err, sizeof g_err, "option out of range");
    return -1;
  }
  std::lock_guard<std::mutex> lock(s->mu);
  s->opt = *o;
  if (s->lg.exec) { cudaGraphExecDestroy(s->lg.exec); s->lg.exec = nullptr; }   // the captured launch holds the old options
  return 0;
}

int rmpc_get_options(const rmpc_solver* s, rmpc_options* o) {
  if (!s || !o) return -1;
  *o = s->opt;
  return 0;
}

static SolveArgs make_args(const rmpc_solver* s, const double* xinit, const double* x0, const double* params, long p_bs, long p_ss,
                           double* z_out, double* action_out, const rmpc_info* info) {
  SolveArgs a;
  a.ho = Handoff{nullptr, nullptr, nullptr, nullptr};
  a.xinit = xinit; a.x0 = x0; a.params = params; a.z_out = z_out; a.action = action_out;
  a.p_bs = p_bs; a.p_ss = p_ss;
  memset(&a.info, 0, sizeof a.info);
  if (info) a.info = *info;
  a.opt = to_core(s->opt);
  a.control_mode = s->opt.control_mode;
  return a;
}

// options.precision = 1: classification as usual; shifted starts go to the float64 kernel on the side stream; the feasible
// list is worked through twice on the caller's stream, in float32 (lowp_kernel) and then in float64 (solve_kernel<kResume>).
static int launch_solve_mixed(rmpc_solver* s, int B, SolveArgs a, cudaStream_t st) {
  static_assert(kWarps == 1, "the mixed-precision kernels are one-warp blocks");
  if (!s->slot[kSlots - 1].hit) {   // first mixed-precision call on this handle: the hand-over buffers of every slot at once
    const size_t mb = (size_t)s->max_batch;   // (an allocation waits for the device; it must not happen between the batches of a running loop)
    for (int i = 0; i < kSlots; ++i) {
      Slot& a_ = s->slot[i];
      if (!a_.hz) CK(cudaMalloc(&a_.hz, sizeof(float) * mb * M::N * M::NVAR));
      if (!a_.hlam) CK(cudaMalloc(&a_.hlam, sizeof(float) * mb * M::N * M::NH));
      if (!a_.hmu) CK(cudaMalloc(&a_.hmu, sizeof(float) * mb));
      if (!a_.hit) CK(cudaMalloc(&a_.hit, sizeof(int) * mb));
    }
  }
  Slot& sl = s->slot[s->next_slot];
  s->next_slot = (s->next_slot + 1) % kSlots;
  if (sl.used) CK(cudaStreamWaitEvent(st, sl.ev_done, 0));
  sl.used = true;
  a.ho = Handoff{sl.hz, sl.hlam, sl.hmu, sl.hit};
  CK(cudaMemsetAsync(sl.q, 0, 8 * sizeof(int), st));
  {
    const int warps_per_block = 8;
    long blocks = ((long)B + warps_per_block - 1) / warps_per_block;
    const long cap = (long)s->sm_count * 8;
    if (blocks > cap) blocks = cap;
    classify_kernel<<<(unsigned)blocks, 32 * warps_per_block, 0, st>>>(B, a, sl.q, sl.list, s->max_batch);
    g_launches++;
    CK(cudaGetLastError());
  }
  CK(cudaEventRecord(sl.ev_fork, st));
  CK(cudaStreamWaitEvent(sl.side, sl.ev_fork, 0));
  long grid = (long)s->blocks_per_sm_sh * s->sm_count;
  if (grid > B) grid = B;
  solve_kernel<I32::kShifted><<<(unsigned)grid, 32, s->smem_bytes, sl.side>>>(B, a, sl.list + s->max_batch, sl.q + 1, sl.q + 3);
  g_launches++;
  CK(cudaGetLastError());
  CK(cudaEventRecord(sl.ev_join, sl.side));
  grid = (long)s->bps_lowp * s->sm_count;
  if (grid > B) grid = B;
  lowp_kernel<<<(unsigned)grid, 32, kSharedFPerWarp, st>>>(B, a, sl.list, sl.q, sl.q + 2, sl.hz, sl.hlam, sl.hmu, sl.hit);
  g_launches++;
  CK(cudaGetLastError());
  grid = (long)s->bps_resume * s->sm_count;
  if (grid > B) grid = B;
  solve_kernel<I32::kResume><<<(unsigned)grid, 32, s->smem_bytes, st>>>(B, a, sl.list, sl.q, sl.q + 4);
  g_launches++;
  CK(cudaGetLastError());
  CK(cudaStreamWaitEvent(st, sl.ev_join, 0));
  CK(cudaEventRecord(sl.ev_done, st));
  return 0;
}

// the launches of one batched solve on stream `st`; the caller holds s->mu and has set the device
static int launch_solve(rmpc_solver* s, int B, const SolveArgs& a, cudaStream_t st) {
  if (s->opt.precision == 1 && B > kSmallBatch) return launch_solve_mixed(s, B, a, st);   // small batches (latency path): float64 throughout
  Slot& sl = s->slot[s->next_slot];
  s->next_slot = (s->next_slot + 1) % kSlots;
  if (sl.used) CK(cudaStreamWaitEvent(st, sl.ev_done, 0));   // the call that used this slot kSlots calls ago (any stream) has finished
  sl.used = true;
  const long need = ((long)B + kWarps - 1) / kWarps;
  if (B <= kSmallBatch) {   // one launch: classification inside the solve
    CK(cudaMemsetAsync(sl.q, 0, 8 * sizeof(int), st));
    long grid = (long)s->blocks_per_sm_all * s->sm_count;
    if (grid > need) grid = need;
    solve_kernel<I32::kAll><<<(unsigned)grid, 32 * kWarps, s->smem_bytes_all, st>>>(B, a, nullptr, nullptr, sl.q + 2);
    g_launches++;
    CK(cudaGetLastError());
    CK(cudaEventRecord(sl.ev_done, st));
    return 0;
  }
  CK(cudaMemsetAsync(sl.q, 0, 8 * sizeof(int), st));
  {
    const int warps_per_block = 8;
    long blocks = ((long)B + warps_per_block - 1) / warps_per_block;
    const long cap = (long)s->sm_count * 8;
    if (blocks > cap) blocks = cap;
    classify_kernel<<<(unsigned)blocks, 32 * warps_per_block, 0, st>>>(B, a, sl.q, sl.list, s->max_batch);
    g_launches++;
    CK(cudaGetLastError());
  }
  // fork: shifted starts on the side stream, next to ...
  CK(cudaEventRecord(sl.ev_fork, st));
  CK(cudaStreamWaitEvent(sl.side, sl.ev_fork, 0));
  long grid = (long)s->blocks_per_sm_sh * s->sm_count;
  if (grid > need) grid = need;
  solve_kernel<I32::kShifted><<<(unsigned)grid, 32 * kWarps, s->smem_bytes, sl.side>>>(B, a, sl.list + s->max_batch, sl.q + 1, sl.q + 3);
  g_launches++;
  CK(cudaGetLastError());
  CK(cudaEventRecord(sl.ev_join, sl.side));
  // ... the common case on the caller's stream.  Persistent grid: as many blocks as fit at once, never more than the problems need
  grid = (long)s->blocks_per_sm * s->sm_count;
  if (grid > need) grid = need;
  solve_kernel<I32::kFeasible><<<(unsigned)grid, 32 * kWarps, s->smem_bytes, st>>>(B, a, sl.list, sl.q, sl.q + 2);
  g_launches++;
  CK(cudaGetLastError());
  CK(cudaStreamWaitEvent(st, sl.ev_join, 0));
  CK(cudaEventRecord(sl.ev_done, st));
  return 0;
}

int rmpc_solve_batch(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params, double* z_out,
                     double* action_out, const rmpc_info* info, void* stream) {
  return rmpc_solve_batch_strided(s, B, xinit, x0, params, (long)M::N * M::NPAR, (long)M::NPAR, z_out, action_out, info, stream);
}

static int check_solve_args(const rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params) {
  if (!s || !xinit || !x0 || !params) { snprintf(g_err, sizeof g_err, "null argument"); return -1; }
  if (B <= 0 || B > s->max_batch) { snprintf(g_err, sizeof g_err, "B=%d outside (0, max_batch=%d]", B, s->max_batch); return -1; }
  return 0;
}

int rmpc_solve_batch_strided(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params,
                             long params_batch_stride, long params_stage_stride, double* z_out, double* action_out,
                             const rmpc_info* info, void* stream) {
  if (check_solve_args(s, B, xinit, x0, params)) return -1;
  if (params_stage_stride != 0 && params_stage_stride < M::NPAR) { snprintf(g_err, sizeof g_err, "params_stage_stride %ld: 0 (one block for all stages) or >= npar = %d", params_stage_stride, M::NPAR); return -1; }
  if (params_batch_stride < (params_stage_stride ? params_stage_stride * (M::N - 1) + M::NPAR : M::NPAR)) { snprintf(g_err, sizeof g_err, "params_batch_stride %ld too small", params_batch_stride); return -1; }
  std::lock_guard<std::mutex> lock(s->mu);
  GUARD(s);
  const SolveArgs a = make_args(s, xinit, x0, params, params_batch_stride, params_stage_stride, z_out, action_out, info);
  return launch_solve(s, B, a, (cudaStream_t)stream);
}

// ---- host entry points -------------------------------------------------------------------------------------------
// Device staging: [xinit | x0 | params] and [z | action | 5 residual arrays | exitflag, iters], each section sized for
// max_batch, plus page-locked mirrors of the outputs (and, for small batches, of the inputs).
struct Stage {
  size_t xinit, x0, params, in_total;          // offsets in doubles
  size_t z, action, res, flags, out_total;
};
static Stage stage_layout(size_t mb) {
  Stage L;
  L.xinit = 0; L.x0 = mb * M::NX; L.params = L.x0 + mb * M::N * M::NVAR; L.in_total = L.params + mb * M::N * M::NPAR;
  L.z = 0; L.action = mb * M::N * M::NVAR; L.res = L.action + mb * M::NU; L.flags = L.res + 5 * mb; L.out_total = L.flags + mb;   // 2 ints per double
  return L;
}

static int ensure_host_staging(rmpc_solver* s) {
  if (s->staged) return 0;
  const Stage L = stage_layout((size_t)s->max_batch);
  cudaFree(s->d_in); cudaFree(s->d_out); cudaFreeHost(s->p_out); cudaFreeHost(s->p_in);   // leftovers of an attempt that failed half way
  s->d_in = s->d_out = s->p_out = s->p_in = nullptr;
  CK(cudaMalloc(&s->d_in, sizeof(double) * L.in_total));
  CK(cudaMalloc(&s->d_out, sizeof(double) * L.out_total));
  CK(cudaMallocHost(&s->p_out, sizeof(double) * L.out_total));
  const Stage Ls = stage_layout((size_t)(s->max_batch < kSmallBatch ? s->max_batch : kSmallBatch));
  CK(cudaMallocHost(&s->p_in, sizeof(double) * Ls.in_total));
  if (!s->lat_q) CK(cudaMalloc(&s->lat_q, sizeof(int) * 4));
  s->staged = true;
  return 0;
}

static rmpc_info staged_info(const rmpc_solver* s, double* out, size_t mb) {
  const Stage L = stage_layout(mb);
  rmpc_info d;
  d.exitflag = reinterpret_cast<int*>(out + L.flags); d.iters = d.exitflag + mb;
  d.obj = out + L.res; d.res_stat = d.obj + mb; d.res_eq = d.obj + 2 * mb; d.res_ineq = d.obj + 3 * mb; d.res_comp = d.obj + 4 * mb;
  (void)s;
  return d;
}

// Small batches (B <= kSmallBatch, the single-solve latency case): inputs packed into one page-locked block, then ONE
// captured graph = H2D copy, queue reset, single-launch solve, D2H copy.  The graph is rebuilt when B, the set of wanted
// outputs or the options change.
static int latency_solve(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params, bool want_z,
                         bool want_action, bool want_info) {
  const size_t b = (size_t)B;
  const Stage L = stage_layout(b);   // packed for this B: one contiguous copy each way
  memcpy(s->p_in + L.xinit, xinit, sizeof(double) * b * M::NX);
  memcpy(s->p_in + L.x0, x0, sizeof(double) * b * M::N * M::NVAR);
  memcpy(s->p_in + L.params, params, sizeof(double) * b * M::N * M::NPAR);
  const int flags = (want_z ? 1 : 0) | (want_action ? 2 : 0) | (want_info ? 4 : 0);
  if (!s->lg.exec || s->lg.B != B || s->lg.flags != flags) {
    if (s->lg.exec) { cudaGraphExecDestroy(s->lg.exec); s->lg.exec = nullptr; }
    const rmpc_info dinf = staged_info(s, s->d_out, b);
    const SolveArgs a = make_args(s, s->d_in + L.xinit, s->d_in + L.x0, s->d_in + L.params, (long)M::N * M::NPAR, (long)M::NPAR,
                                  want_z ? s->d_out + L.z : nullptr, want_action ? s->d_out + L.action : nullptr, want_info ? &dinf : nullptr);
    cudaGraph_t graph;
    CK(cudaStreamBeginCapture(s->hstream, cudaStreamCaptureModeThreadLocal));
    cudaError_t e = cudaMemcpyAsync(s->d_in, s->p_in, sizeof(double) * L.in_total, cudaMemcpyHostToDevice, s->hstream);
    if (e == cudaSuccess) e = cudaMemsetAsync(s->lat_q, 0, 4 * sizeof(int), s->hstream);
    if (e == cudaSuccess) {
      long grid = (long)s->blocks_per_sm_all * s->sm_count;
      if (grid > B) grid = B;
      solve_kernel<I32::kAll><<<(unsigned)grid, 32 * kWarps, s->smem_bytes_all, s->hstream>>>(B, a, nullptr, nullptr, s->lat_q + 2);
      e = cudaGetLastError();
    }
    // outputs: from the first wanted section to the end of the flags (z is large and first, so it is skipped when unwanted)
    const size_t from = want_z ? L.z : L.action;
    if (e == cudaSuccess) e = cudaMemcpyAsync(s->p_out + from, s->d_out + from, sizeof(double) * (L.out_total - from), cudaMemcpyDeviceToHost, s->hstream);
    cudaError_t e2 = cudaStreamEndCapture(s->hstream, &graph);
    if (e != cudaSuccess || e2 != cudaSuccess) {
      snprintf(g_err, sizeof g_err, "graph capture: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
      return -1;
    }
    e = cudaGraphInstantiate(&s->lg.exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) { s->lg.exec = nullptr; snprintf(g_err, sizeof g_err, "cudaGraphInstantiate: %s", cudaGetErrorString(e)); return -1; }
    s->lg.B = B; s->lg.flags = flags;
  }
  CK(cudaGraphLaunch(s->lg.exec, s->hstream));
  g_launches++;
  return 0;
}

int rmpc_solve_batch_host_async(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params,
                                double* z_out, double* action_out, const rmpc_info* info) {
  if (check_solve_args(s, B, xinit, x0, params)) return -1;
  std::lock_guard<std::mutex> lock(s->mu);
  if (s->pend_B) { snprintf(g_err, sizeof g_err, "a host solve is already in flight on this handle: call rmpc_wait first"); return -1; }
  GUARD(s);
  if (ensure_host_staging(s)) return -1;
  const size_t b = (size_t)B, mb = (size_t)s->max_batch;
  if (B <= kSmallBatch) {
    if (latency_solve(s, B, xinit, x0, params, z_out != nullptr, action_out != nullptr, info != nullptr)) return -1;
  } else {
    const Stage L = stage_layout(mb);
    cudaStream_t st = s->hstream;
    CK(cudaMemcpyAsync(s->d_in + L.xinit, xinit, sizeof(double) * b * M::NX, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(s->d_in + L.x0, x0, sizeof(double) * b * M::N * M::NVAR, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(s->d_in + L.params, params, sizeof(double) * b * M::N * M::NPAR, cudaMemcpyHostToDevice, st));
    const rmpc_info dinf = staged_info(s, s->d_out, mb);
    if (info) CK(cudaMemsetAsync(dinf.exitflag, 0xff, sizeof(int) * b, st));   // -1 = "never written" would show an internal error
    const SolveArgs a = make_args(s, s->d_in + L.xinit, s->d_in + L.x0, s->d_in + L.params, (long)M::N * M::NPAR, (long)M::NPAR,
                                  z_out ? s->d_out + L.z : nullptr, action_out ? s->d_out + L.action : nullptr, info ? &dinf : nullptr);
    if (launch_solve(s, B, a, st)) return -1;
    if (z_out) CK(cudaMemcpyAsync(s->p_out + L.z, s->d_out + L.z, sizeof(double) * b * M::N * M::NVAR, cudaMemcpyDeviceToHost, st));
    if (action_out) CK(cudaMemcpyAsync(s->p_out + L.action, s->d_out + L.action, sizeof(double) * b * M::NU, cudaMemcpyDeviceToHost, st));
    if (info) CK(cudaMemcpyAsync(s->p_out + L.res, s->d_out + L.res, sizeof(double) * (L.out_total - L.res), cudaMemcpyDeviceToHost, st));
  }
  s->pend_B = B; s->u_z = z_out; s->u_action = action_out;
  memset(&s->u_info, 0, sizeof s->u_info);
  if (info) s->u_info = *info;
  return 0;
}

int rmpc_wait(rmpc_solver* s) {
  if (!s) { snprintf(g_err, sizeof g_err, "null argument"); return -1; }
  std::lock_guard<std::mutex> lock(s->mu);
  GUARD(s);
  CK(cudaStreamSynchronize(s->hstream));
  if (!s->pend_B) return 0;
  const size_t b = (size_t)s->pend_B;
  const size_t lb = s->pend_B <= kSmallBatch ? b : (size_t)s->max_batch;   // small batches are packed for their own B
  s->pend_B = 0;
  const Stage L = stage_layout(lb);
  if (s->u_z) memcpy(s->u_z, s->p_out + L.z, sizeof(double) * b * M::N * M::NVAR);
  if (s->u_action) memcpy(s->u_action, s->p_out + L.action, sizeof(double) * b * M::NU);
  const rmpc_info& u = s->u_info;
  const rmpc_info h = staged_info(s, s->p_out, lb);
  if (u.exitflag) memcpy(u.exitflag, h.exitflag, sizeof(int) * b);
  if (u.iters) memcpy(u.iters, h.iters, sizeof(int) * b);
  double* dst[5] = {u.obj, u.res_stat, u.res_eq, u.res_ineq, u.res_comp};
  const double* src[5] = {h.obj, h.res_stat, h.res_eq, h.res_ineq, h.res_comp};
  for (int i = 0; i < 5; ++i)
    if (dst[i]) memcpy(dst[i], src[i], sizeof(double) * b);
  return 0;
}

int rmpc_solve_batch_host(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params,
                          double* z_out, double* action_out, const rmpc_info* info) {
  if (rmpc_solve_batch_host_async(s, B, xinit, x0, params, z_out, action_out, info)) return -1;
  return rmpc_wait(s);
}

int rmpc_shift_horizon(rmpc_solver* s, int B, const double* z, double* x0, void* stream) {
  if (!s || !z || !x0 || B <= 0) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  std::lock_guard<std::mutex> lock(s->mu);
  GUARD(s);
  const long total = (long)B * M::NVAR;
  shift_kernel<<<(unsigned)((total + 127) / 128), 128, 0, (cudaStream_t)stream>>>(total, z, x0);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

int rmpc_set_x0(rmpc_solver* s, int B, const double* xinit, double* x0, void* stream) {
  if (!s || !xinit || !x0 || B <= 0) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  std::lock_guard<std::mutex> lock(s->mu);
  GUARD(s);
  const long total = (long)B * M::N * M::NX;
  guess_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(total, xinit, x0);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

int rmpc_plant_step(rmpc_solver* s, int B, const double* x, const double* u, const int* exitflag, double* x_next, void* stream) {
  if (!s || !x || !u || !x_next || B <= 0) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  std::lock_guard<std::mutex> lock(s->mu);
  GUARD(s);
  plant_kernel<<<(unsigned)((B + 127) / 128), 128, 0, (cudaStream_t)stream>>>(B, x, u, exitflag, x_next);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

int rmpc_set_params(rmpc_solver* s, int B, double* params, const int* idx, int n_idx, const double* values,
                    int stage_varying, void* stream) {
  if (!s || !params || !idx || !values || B <= 0 || n_idx <= 0) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  for (int j = 0; j < n_idx; ++j)
    if (idx[j] < 0 || idx[j] >= M::NPAR) { snprintf(g_err, sizeof g_err, "parameter index %d out of range", idx[j]); return -1; }
  std::lock_guard<std::mutex> lock(s->mu);
  GUARD(s);
  for (int off = 0; off < n_idx; off += kMaxIdx) {   // the indices travel by value with the launch: nothing shared between calls
    IdxList il;
    il.n = n_idx - off < kMaxIdx ? n_idx - off : kMaxIdx;
    for (int j = 0; j < il.n; ++j) il.v[j] = idx[off + j];
    const long total = (long)B * M::N * il.n;
    scatter_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(B, n_idx, off, stage_varying, il, values, params);
    g_launches++;
    CK(cudaGetLastError());
  }
  return 0;
}

int rmpc_predict_obstacles(rmpc_solver* s, int B, const double* obst_state, int n_dyn, int n_slots, int p_obst, double radius,
                           double* params, void* stream) {
  if (!s || !params || B <= 0 || n_dyn < 0 || n_slots <= 0 || (n_dyn > 0 && !obst_state)) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  if (p_obst < 0 || p_obst + 4 * n_slots > M::NPAR) { snprintf(g_err, sizeof g_err, "obstacle slots %d..%d outside the parameter vector (npar = %d)", p_obst, p_obst + 4 * n_slots, M::NPAR); return -1; }
  if (n_dyn > n_slots) n_dyn = n_slots;   // the reference ignores obstacles beyond number_obstacles (mpcPlanner.py:146-147)
  std::lock_guard<std::mutex> lock(s->mu);
  GUARD(s);
  const long total = (long)B * M::N * n_slots;
  predict_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(B, n_dyn, n_slots, p_obst, radius, obst_state, params);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

int rmpc_fsd_constraints(rmpc_solver* s, int B, const double* points, int n_points, const double* ref, long ref_batch_stride,
                         long ref_stage_stride, double max_radius, double lidar_offset, double lidar_z, int K, double* out,
                         double* params, const int* idx, void* stream) {
  if (!s || !points || !ref || B <= 0 || (!out && !params)) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  if (K <= 0 || K > kFsdMaxK || n_points <= 0 || n_points > kFsdMaxPoints) {
    snprintf(g_err, sizeof g_err, "K must be in [1, %d] and n_points in [1, %d]", kFsdMaxK, kFsdMaxPoints);
    return -1;
  }
  FsdArgs a;
  memset(&a, 0, sizeof a);
  a.B = B; a.N = M::N; a.P = n_points; a.K = K; a.points = points; a.ref = ref; a.ref_bs = ref_batch_stride; a.ref_ss = ref_stage_stride;
  a.max_radius = max_radius; a.lidar_offset = lidar_offset; a.lidar_z = lidar_z; a.out = out; a.params = params; a.npar = M::NPAR;
  if (params) {
    if (!idx) { snprintf(g_err, sizeof g_err, "idx is required with params"); return -1; }
    for (int c = 0; c < K; ++c) {
      if (idx[c] < 0 || idx[c] + 4 > M::NPAR) { snprintf(g_err, sizeof g_err, "parameter index %d out of range", idx[c]); return -1; }
      a.idx[c] = idx[c];
    }
  }
  std::lock_guard<std::mutex> lock(s->mu);
  GUARD(s);
  const long warps = (long)B * M::N;
  fsd_kernel<<<(unsigned)((warps + 7) / 8), 256, 0, (cudaStream_t)stream>>>(a);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}


// ---- rmpc_solve_batch_static: a batched solve that allocates nothing --------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch memory of a solve — the two
// queue heads and the two problem lists the classification pass fills — lives in a file-scope device array instead of
// the handle's cudaMalloc'ed slots.  Every GPU that loads this library holds the array until the process exits.
constexpr int kStaticMaxBatch = 16384;                       // problems per call of rmpc_solve_batch_static
constexpr int kStaticHeaderInts = 64;                        // 256 bytes: [0] feasible count, [1] shifted count, [2] feasible head, [3] shifted head
static __device__ __align__(256) int g_static_scratch[kStaticHeaderInts + 2 * kStaticMaxBatch];   // 131 328 bytes

constexpr int kStaticMaxDevices = 64;
struct StaticDevice {
  bool ready;
  int* scratch;          // device address of g_static_scratch on this device (cudaGetSymbolAddress; the symbol itself is a host-side stand-in)
  int sm_count, bps, bps_sh;
};
static StaticDevice g_static_dev[kStaticMaxDevices];
static std::mutex g_static_mu;

// once per device, with that device current (the caller's current device); no allocation
static int static_setup(int dev) {
  StaticDevice& d = g_static_dev[dev];
  if (d.ready) return 0;
  void* ptr = nullptr;
  CK(cudaGetSymbolAddress(&ptr, g_static_scratch));
  d.scratch = static_cast<int*>(ptr);
  const size_t smem = kSharedPerWarp * kWarps;
  CK(cudaFuncSetAttribute(solve_kernel<I32::kFeasible>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(solve_kernel<I32::kShifted>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.bps, solve_kernel<I32::kFeasible>, 32 * kWarps, smem));
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.bps_sh, solve_kernel<I32::kShifted>, 32 * kWarps, smem));
  CK(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev));
  if (d.bps < 1 || d.bps_sh < 1) { snprintf(g_err, sizeof g_err, "solve kernel does not fit on an SM (%zu B shared)", smem); return -1; }
  d.ready = true;
  return 0;
}

int rmpc_solve_batch_static(int B, const double* xinit, const double* x0, const double* params, double* z_out, double* action_out,
                            const rmpc_info* info, const rmpc_options* opt, void* stream) {
  if (!xinit || !x0 || !params) { snprintf(g_err, sizeof g_err, "null argument"); return -1; }
  if (B <= 0 || B > kStaticMaxBatch) {
    snprintf(g_err, sizeof g_err, "rmpc_solve_batch_static: B=%d outside (0, %d] (the static scratch array holds two lists of %d problems)", B, kStaticMaxBatch, kStaticMaxBatch);
    return -1;
  }
  rmpc_options o;
  rmpc_default_options(&o);
  if (opt) {
    if (opt->precision != 0 || opt->hessian < 0 || opt->hessian > 1 || opt->control_mode < 0 || opt->control_mode > 1 || opt->creep_trigger < 0 || opt->stall_window < 0) {
      snprintf(g_err, sizeof g_err, "option out of range");
      return -1;
    }
    o = *opt;
  }
  int dev = -1;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kStaticMaxDevices) { snprintf(g_err, sizeof g_err, "device %d outside the static table", dev); return -1; }
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (static_setup(dev)) return -1;
  const StaticDevice& d = g_static_dev[dev];
  cudaStream_t st = (cudaStream_t)stream;
  SolveArgs a;
  a.ho = Handoff{nullptr, nullptr, nullptr, nullptr};
  a.xinit = xinit; a.x0 = x0; a.params = params; a.z_out = z_out; a.action = action_out;
  a.p_bs = (long)M::N * M::NPAR; a.p_ss = (long)M::NPAR;
  memset(&a.info, 0, sizeof a.info);
  if (info) a.info = *info;
  a.opt = to_core(o);
  a.control_mode = o.control_mode;
  int* q = d.scratch;
  int* list = d.scratch + kStaticHeaderInts;
  // the array keeps the last call's contents: counts and heads must start at zero; the lists are rewritten up to the new counts
  CK(cudaMemsetAsync(q, 0, kStaticHeaderInts * sizeof(int), st));
  {
    const int warps_per_block = 8;
    long blocks = ((long)B + warps_per_block - 1) / warps_per_block;
    const long cap = (long)d.sm_count * 8;
    if (blocks > cap) blocks = cap;
    classify_kernel<<<(unsigned)blocks, 32 * warps_per_block, 0, st>>>(B, a, q, list, kStaticMaxBatch);
    g_launches++;
    CK(cudaGetLastError());
  }
  const long need = ((long)B + kWarps - 1) / kWarps;
  const size_t smem = kSharedPerWarp * kWarps;
  // one stream, no fork (no streams or events of a handle here): feasible starts, then the (rare) shifted starts
  long grid = (long)d.bps * d.sm_count;
  if (grid > need) grid = need;
  solve_kernel<I32::kFeasible><<<(unsigned)grid, 32 * kWarps, smem, st>>>(B, a, list, q, q + 2);
  g_launches++;
  CK(cudaGetLastError());
  grid = (long)d.bps_sh * d.sm_count;
  if (grid > need) grid = need;
  solve_kernel<I32::kShifted><<<(unsigned)grid, 32 * kWarps, smem, st>>>(B, a, list + kStaticMaxBatch, q + 1, q + 3);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Result exchange over peer memory (rmpc_exchange.cuh): host side.
struct rmpc_exchange {
  int device;
  rmpc_x::View view;
  std::mutex mu;
  long next_publish, next_collect;
  cudaStream_t pstream;   // the publishes of this rank, in step order, each behind the solve it belongs to
  cudaStream_t xstream;   // the collects of this rank, in step order
  cudaEvent_t ev;
};

static_assert(M::NU + 1 <= 8, "publish_kernel stages a row of at most 8 doubles");
size_t rmpc_exchange_region_bytes(int world, int rows_per_rank, int depth) {
  if (world < 1 || world > rmpc_x::kMaxWorld || rows_per_rank < 1 || depth < 1) return 0;
  return rmpc_x::Layout{world, rows_per_rank, M::NU + 1, depth}.bytes();
}

int rmpc_ipc_alloc(int device, size_t bytes, void** ptr, unsigned char* handle64) {
  if (!ptr || bytes == 0) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  DeviceGuard guard(device); CK(guard.err);
  CK(cudaMalloc(ptr, bytes));
  CK(cudaMemset(*ptr, 0, bytes));
  CK(cudaDeviceSynchronize());
  if (handle64) {
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, *ptr));
    memcpy(handle64, &h, 64);
  }
  return 0;
}

int rmpc_ipc_open(int device, const unsigned char* handle64, void** ptr) {
  if (!ptr || !handle64) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  DeviceGuard guard(device); CK(guard.err);
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  CK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int rmpc_ipc_close(int device, void* ptr) {
  DeviceGuard guard(device); CK(guard.err);
  CK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

int rmpc_ipc_free(int device, void* ptr) {
  DeviceGuard guard(device); CK(guard.err);
  CK(cudaFree(ptr));
  return 0;
}

rmpc_exchange* rmpc_exchange_create(int device, int world, int rank, int rows_per_rank, int depth, void* const* regions) {
  if (world < 1 || world > rmpc_x::kMaxWorld || rank < 0 || rank >= world || rows_per_rank < 1 || depth < 1 || !regions) {
    snprintf(g_err, sizeof g_err, "rmpc_exchange_create: world 1..%d, 0 <= rank < world, rows_per_rank >= 1, depth >= 1", rmpc_x::kMaxWorld);
    return nullptr;
  }
  for (int w = 0; w < world; ++w)
    if (!regions[w]) { snprintf(g_err, sizeof g_err, "rmpc_exchange_create: region %d is NULL", w); return nullptr; }
  DeviceGuard guard(device);
  if (guard.err != cudaSuccess) { snprintf(g_err, sizeof g_err, "cudaSetDevice(%d): %s", device, cudaGetErrorString(guard.err)); return nullptr; }
  auto* x = new rmpc_exchange();
  x->device = device;
  x->view.l = rmpc_x::Layout{world, rows_per_rank, M::NU + 1, depth};
  x->view.rank = rank;
  for (int w = 0; w < rmpc_x::kMaxWorld; ++w) x->view.region[w] = w < world ? static_cast<char*>(regions[w]) : nullptr;
  x->next_publish = x->next_collect = 0;
  if (cudaStreamCreateWithFlags(&x->pstream, cudaStreamNonBlocking) != cudaSuccess || cudaStreamCreateWithFlags(&x->xstream, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&x->ev, cudaEventDisableTiming) != cudaSuccess) {
    snprintf(g_err, sizeof g_err, "rmpc_exchange_create: %s", cudaGetErrorString(cudaGetLastError()));
    delete x;
    return nullptr;
  }
  return x;
}

void rmpc_exchange_destroy(rmpc_exchange* x) {
  if (!x) return;
  DeviceGuard guard(x->device);
  cudaStreamSynchronize(x->pstream);
  cudaStreamSynchronize(x->xstream);
  cudaStreamDestroy(x->pstream);
  cudaStreamDestroy(x->xstream);
  cudaEventDestroy(x->ev);
  delete x;
}

long rmpc_exchange_publish(rmpc_exchange* x, int B, const double* action, const int* exitflag, const int* iters, void* stream) {
  if (!x || !action || B < 0 || B > x->view.l.rows) { snprintf(g_err, sizeof g_err, "rmpc_exchange_publish: 0 <= B <= rows_per_rank (%d)", x ? x->view.l.rows : 0); return -1; }
  std::lock_guard<std::mutex> lock(x->mu);
  GUARD(x);
  const long seq = x->next_publish;
  const int depth = x->view.l.depth, slot = int(seq % depth);
  // The stores run on the exchange's own stream behind what `stream` holds now (the solve): a publish that has to wait for
  // a free ring slot, i.e. for the slowest rank, then holds up later publishes but never a solve.
  cudaStream_t st = x->pstream;
  CK(cudaEventRecord(x->ev, (cudaStream_t)stream));
  CK(cudaStreamWaitEvent(st, x->ev, 0));
  if (seq >= depth) {   // the slot is free once every rank has collected the step that used it last
    rmpc_x::wait_kernel<<<1, 32, 0, st>>>(x->view, 0, 0, (unsigned long long)(seq - depth + 1));
    g_launches++;
  }
  const int blocks = B > 0 ? (B + 255) / 256 : 1;
  rmpc_x::publish_kernel<<<blocks, 256, 0, st>>>(x->view, slot, (unsigned long long)(seq + 1), B, M::NU, action, exitflag, iters);
  g_launches++;
  CK(cudaGetLastError());
  x->next_publish = seq + 1;
  return seq;
}

int rmpc_exchange_collect(rmpc_exchange* x, long seq, double* action_all, int* exitflag_all, int* iters_all, void* stream) {
  if (!x || !action_all) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  std::lock_guard<std::mutex> lock(x->mu);
  if (seq != x->next_collect || seq >= x->next_publish) {
    snprintf(g_err, sizeof g_err, "rmpc_exchange_collect: steps are collected in order, after their publish (next %ld, published %ld, asked %ld)",
             x->next_collect, x->next_publish, seq);
    return -1;
  }
  GUARD(x);
  const int slot = int(seq % x->view.l.depth);
  const int total = x->view.l.world * x->view.l.rows;
  // the waiting warp and the unpack run on the exchange's own stream: no solve stream is held until the slowest rank arrives.
  // The unpack is ordered behind what `stream` holds so far (the output buffers may come from a stream-ordered allocator).
  rmpc_x::wait_kernel<<<1, 32, 0, x->xstream>>>(x->view, 1, slot, (unsigned long long)(seq + 1));
  CK(cudaEventRecord(x->ev, (cudaStream_t)stream));
  CK(cudaStreamWaitEvent(x->xstream, x->ev, 0));
  rmpc_x::unpack_kernel<<<(total + 255) / 256, 256, 0, x->xstream>>>(x->view, slot, (unsigned long long)(seq + 1), M::NU, action_all, exitflag_all, iters_all);
  g_launches += 2;
  CK(cudaGetLastError());
  x->next_collect = seq + 1;
  return 0;
}

int rmpc_exchange_join(rmpc_exchange* x, void* stream) {
  if (!x) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  std::lock_guard<std::mutex> lock(x->mu);
  GUARD(x);
  CK(cudaEventRecord(x->ev, x->xstream));
  CK(cudaStreamWaitEvent((cudaStream_t)stream, x->ev, 0));
  return 0;
}

int rmpc_exchange_status(rmpc_exchange* x) {
  if (!x) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  std::lock_guard<std::mutex> lock(x->mu);
  GUARD(x);
  unsigned long long st = 0;
  CK(cudaMemcpy(&st, x->view.region[x->view.rank], sizeof st, cudaMemcpyDeviceToHost));   // synchronises with the device
  return int(st);
}

int rmpc_static_max_batch(void) { return kStaticMaxBatch; }

}  // extern "C"
