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

#include "model_traits.h"
#ifndef RMPC_WARPS_FIXED
#define RMPC_WARPS_FIXED 1
#endif
#define RMPC_WARPS RMPC_WARPS_FIXED
#include "rmpc_ipm.cuh"
#include "rmpc_fsd.cuh"
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

constexpr int kSideStreams = 4;

struct rmpc_solver {
  int device;
  int max_batch;
  rmpc_options opt;
  int* counter;       // device work-queue heads, one per in-flight call slot (round robin)
  int counter_slot;
  int blocks_per_sm, blocks_per_sm_sh, sm_count;
  size_t smem_bytes;  // dynamic shared memory per block
  // the kernel for infeasible starts runs on a side stream, forked from and joined to the caller's stream; a few of
  // them in rotation, so that calls on different streams do not queue behind each other here
  cudaStream_t side[kSideStreams];
  cudaEvent_t ev_fork[kSideStreams], ev_join[kSideStreams];
  int side_slot;
  double* z_tmp;      // scratch plan for the in-place horizon shift
  int* d_idx;         // scratch for rmpc_set_params
  int d_idx_cap;
  // device staging for the host entry point
  double *h_xinit, *h_x0, *h_params, *h_z, *h_action, *h_res;
  int* h_flag;
  cudaStream_t hstream;  // private stream of the host entry points
  // page-locked landing zone of the asynchronous D2H copies, and the caller's buffers they are handed to in rmpc_wait
  double *p_z, *p_action, *p_res;
  int* p_flag;
  int pend_B;
  double *u_z, *u_action;
  rmpc_info u_info;
};

// ---------------------------------------------------------------------------------------------
// One warp = one NLP (rmpc_ipm.cuh, G = 32): lane k owns stage k in the stage-parallel phases, the 32 lanes share the
// per-stage dense algebra of the Riccati sweep through the warp's slice of shared memory.  Warps pull problems from a
// device-side counter, so a hard problem (many iterations) only delays its own warp.  Inputs and outputs are read and
// written in the reference's own layout ([B][N][nvar], [B][N][npar]); there is no separate pack/unpack pass.
using Shared = ProblemShared<M, double>;
constexpr size_t kSharedPerWarp = (sizeof(Shared) + 15) / 16 * 16;
// Warps (= problems) per thread block.  With more than one, the warps of a block meet at the top of every solver
// iteration (Group<32>::align) so that they walk through the same code together and share instruction-cache fills.
// Measured on B200 (boxer, 20 steps): 1 warp/block 702 k solves/s, 2: 622 k, 3: 574 k, 7: 534 k — waiting for the
// slowest iteration of the block costs more than the shared fetches save, so the default stays at 1.
#ifndef RMPC_WARPS_FIXED
#define RMPC_WARPS_FIXED 1
#endif
constexpr int kWarps = RMPC_WARPS_FIXED;
static_assert(kWarps >= 1, "one problem does not fit into the shared memory of an SM");

#ifndef RMPC_MIN_BLOCKS
#define RMPC_MIN_BLOCKS 1   // resident blocks per SM the register allocation must allow
#endif
struct SolveArgs {
  const double *xinit, *x0, *params;
  long p_bs, p_ss;   // elements between the parameter blocks of consecutive problems / stages (reference layout: N*npar, npar)
  double *z_out, *action;
  rmpc_info info;
  Options opt;
  int control_mode;
};

__device__ __forceinline__ void write_outputs(const SolveArgs& a, const Shared& S, int lane, int b, const Info& inf) {
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

// Two instantiations, launched side by side on two streams.  kFeasible is the common case and contains no code for
// rows that start infeasible; kShifted carries both copies of the iteration and solves only problems whose initial guess
// violates a row (for a feasible batch it evaluates each problem's rows once and exits).  Each problem is solved and
// written by exactly one of the two.  With the rare path in the same kernel as the hot one the solver was 4-10 %
// slower (inlined) or 25 % slower (as a called function); B200, r1l A/B in DESIGN.md.
template <int MODE>
__global__ void __launch_bounds__(32 * kWarps, RMPC_MIN_BLOCKS)
solve_kernel(int B, const __grid_constant__ SolveArgs a, int* __restrict__ counter) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5;
  Shared& S = *reinterpret_cast<Shared*>(smem_raw + warp * kSharedPerWarp);
  Group<32> g{int(threadIdx.x & 31)};
  for (;;) {
    int b = 0;
    if (g.lane == 0) b = atomicAdd(counter, 1);
    b = __shfl_sync(0xffffffffu, b, 0);
    if (b >= B) {
      // out of work: keep meeting the other warps of the block at their iteration rendezvous until all are out of work
      if (kWarps > 1) while (!__syncthreads_and(1)) {}
      break;
    }
    const ParamView<double> pv{a.params + (size_t)b * a.p_bs, a.p_ss, 1};
    Info inf;
    bool mine;
    {
      Ipm<M, double, 32> ipm(g, S, pv, a.opt);
      mine = ipm.template solve<MODE>(a.xinit + (size_t)b * M::NX, a.x0 + (size_t)b * M::N * M::NVAR,
                                      a.z_out ? a.z_out + (size_t)b * M::N * M::NVAR : nullptr, inf);
    }
    if (mine) write_outputs(a, S, g.lane, b, inf);
    __syncwarp();
  }
}

__global__ void shift_kernel(long total, const double* __restrict__ z, double* __restrict__ x0) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int per = M::N * M::NVAR;
  const long b = i / per;
  const int r = int(i - b * per);
  const int k = r / M::NVAR;
  const int ks = k < M::N - 1 ? k + 1 : k;
  // in-place safe: element (k) reads (k+1) which no other thread writes before ... so stage through a register and
  // rely on one thread per element reading strictly-later data only; x0 == z aliasing is handled by the host (copy).
  x0[i] = z[b * per + (long)ks * M::NVAR + (r - k * M::NVAR)];
}

__global__ void scatter_kernel(int B, int n_idx, int stage_varying, const int* __restrict__ idx,
                               const double* __restrict__ values, double* __restrict__ params) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long total = (long)B * M::N * n_idx;
  if (i >= total) return;
  const int j = int(i % n_idx);
  const long bk = i / n_idx;
  const long b = bk / M::N;
  const double v = stage_varying ? values[i] : values[b * n_idx + j];
  params[bk * M::NPAR + idx[j]] = v;
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

constexpr int kCounterSlots = 64;

static int free_all(rmpc_solver* s) {
  cudaFree(s->counter); cudaFree(s->z_tmp);
  cudaFree(s->d_idx); cudaFree(s->h_xinit); cudaFree(s->h_x0); cudaFree(s->h_params); cudaFree(s->h_z);
  cudaFree(s->h_action); cudaFree(s->h_res); cudaFree(s->h_flag);
  if (s->hstream) cudaStreamDestroy(s->hstream);
  for (int i = 0; i < kSideStreams; ++i) {
    if (s->side[i]) cudaStreamDestroy(s->side[i]);
    if (s->ev_fork[i]) cudaEventDestroy(s->ev_fork[i]);
    if (s->ev_join[i]) cudaEventDestroy(s->ev_join[i]);
  }
  cudaFreeHost(s->p_z); cudaFreeHost(s->p_action); cudaFreeHost(s->p_res); cudaFreeHost(s->p_flag);
  return 0;
}

rmpc_solver* rmpc_create(int device, int max_batch) {
  if (max_batch <= 0) { snprintf(g_err, sizeof g_err, "max_batch must be positive"); return nullptr; }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || device < 0 || device >= ndev) {
    snprintf(g_err, sizeof g_err, "no usable CUDA device %d (%s); this library has no CPU fallback", device,
             e == cudaSuccess ? "out of range" : cudaGetErrorString(e));
    return nullptr;
  }
  rmpc_solver* s = new rmpc_solver();
  memset(s, 0, sizeof *s);
  s->device = device;
  s->max_batch = max_batch;
  rmpc_default_options(&s->opt);
  auto fail = [&](const char* what, cudaError_t err) -> rmpc_solver* {
    snprintf(g_err, sizeof g_err, "%s: %s", what, cudaGetErrorString(err));
    free_all(s);
    delete s;
    return nullptr;
  };
  if ((e = cudaSetDevice(device)) != cudaSuccess) return fail("cudaSetDevice", e);
  if ((e = cudaMalloc(&s->counter, sizeof(int) * kCounterSlots)) != cudaSuccess) return fail("cudaMalloc", e);
  size_t smem = kSharedPerWarp * kWarps;
  if (const char* pad = getenv("RMPC_PAD_SMEM_KB")) smem += (size_t)atoi(pad) * 1024;  // experiment knob: fewer resident problems per SM
  s->smem_bytes = smem;
  using I = Ipm<M, double, 32>;
  if ((e = cudaFuncSetAttribute(solve_kernel<I::kFeasible>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(solve_kernel<I::kShifted>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess)
    return fail("cudaFuncSetAttribute(shared memory)", e);
  if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&s->blocks_per_sm, solve_kernel<I::kFeasible>, 32 * kWarps, smem)) != cudaSuccess ||
      (e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&s->blocks_per_sm_sh, solve_kernel<I::kShifted>, 32 * kWarps, smem)) != cudaSuccess)
    return fail("cudaOccupancyMaxActiveBlocksPerMultiprocessor", e);
  for (int i = 0; i < kSideStreams; ++i) {
    if ((e = cudaStreamCreateWithFlags(&s->side[i], cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    if ((e = cudaEventCreateWithFlags(&s->ev_fork[i], cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&s->ev_join[i], cudaEventDisableTiming)) != cudaSuccess)
      return fail("cudaEventCreate", e);
  }
  if (const char* cap = getenv("RMPC_MAX_BLOCKS_PER_SM")) {  // experiment knob: fewer problems in flight per SM
    const int c = atoi(cap);
    if (c > 0 && c < s->blocks_per_sm) s->blocks_per_sm = c;
  }
  if (s->blocks_per_sm < 1 || s->blocks_per_sm_sh < 1) { snprintf(g_err, sizeof g_err, "solve kernel does not fit on an SM (%zu B shared)", smem); free_all(s); delete s; return nullptr; }
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return fail("cudaGetDeviceProperties", e);
  s->sm_count = prop.multiProcessorCount;
  if ((e = cudaStreamCreateWithFlags(&s->hstream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
  return s;
}

void rmpc_destroy(rmpc_solver* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  free_all(s);
  delete s;
}

int rmpc_set_options(rmpc_solver* s, const rmpc_options* o) {
  if (!s || !o) { snprintf(g_err, sizeof g_err, "null argument"); return -1; }
  if (o->precision < 0 || o->precision > 1 || o->hessian < 0 || o->hessian > 1 || o->control_mode < 0 || o->control_mode > 1 ||
      o->creep_trigger < 0 || o->stall_window < 0) {
    snprintf(g_err, sizeof g_err, "option out of range");
    return -1;
  }
  s->opt = *o;
  return 0;
}

int rmpc_get_options(const rmpc_solver* s, rmpc_options* o) {
  if (!s || !o) return -1;
  *o = s->opt;
  return 0;
}

int rmpc_solve_batch(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params, double* z_out,
                     double* action_out, const rmpc_info* info, void* stream) {
  return rmpc_solve_batch_strided(s, B, xinit, x0, params, (long)M::N * M::NPAR, (long)M::NPAR, z_out, action_out, info, stream);
}

int rmpc_solve_batch_strided(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params,
                             long params_batch_stride, long params_stage_stride, double* z_out, double* action_out,
                             const rmpc_info* info, void* stream) {
  if (!s || !xinit || !x0 || !params) { snprintf(g_err, sizeof g_err, "null argument"); return -1; }
  if (B <= 0 || B > s->max_batch) { snprintf(g_err, sizeof g_err, "B=%d outside (0, max_batch=%d]", B, s->max_batch); return -1; }
  if (params_stage_stride != 0 && params_stage_stride < M::NPAR) { snprintf(g_err, sizeof g_err, "params_stage_stride %ld: 0 (one block for all stages) or >= npar = %d", params_stage_stride, M::NPAR); return -1; }
  if (params_batch_stride < (params_stage_stride ? params_stage_stride * (M::N - 1) + M::NPAR : M::NPAR)) { snprintf(g_err, sizeof g_err, "params_batch_stride %ld too small", params_batch_stride); return -1; }
  if (s->opt.precision != 0) { snprintf(g_err, sizeof g_err, "precision %d is not available in this build (float64 only)", s->opt.precision); return -1; }
  CK(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  int* counter = s->counter + s->counter_slot;   // two queue heads per call
  s->counter_slot = (s->counter_slot + 2) % kCounterSlots;
  CK(cudaMemsetAsync(counter, 0, 2 * sizeof(int), st));
  SolveArgs a;
  a.xinit = xinit; a.x0 = x0; a.params = params; a.z_out = z_out; a.action = action_out;
  a.p_bs = params_batch_stride; a.p_ss = params_stage_stride;
  memset(&a.info, 0, sizeof a.info);
  if (info) a.info = *info;
  a.opt = to_core(s->opt);
  a.control_mode = s->opt.control_mode;
  using I = Ipm<M, double, 32>;
  const long need = ((long)B + kWarps - 1) / kWarps;
  // fork: infeasible starts on the side stream, next to ...
  const int ss = s->side_slot;
  s->side_slot = (ss + 1) % kSideStreams;
  CK(cudaEventRecord(s->ev_fork[ss], st));
  CK(cudaStreamWaitEvent(s->side[ss], s->ev_fork[ss], 0));
  long grid = (long)s->blocks_per_sm_sh * s->sm_count;
  if (grid > need) grid = need;
  solve_kernel<I::kShifted><<<(unsigned)grid, 32 * kWarps, s->smem_bytes, s->side[ss]>>>(B, a, counter + 1);
  g_launches++;
  CK(cudaGetLastError());
  CK(cudaEventRecord(s->ev_join[ss], s->side[ss]));
  // ... the common case on the caller's stream.  Persistent grid: as many blocks as fit at once, never more than the problems need
  grid = (long)s->blocks_per_sm * s->sm_count;
  if (grid > need) grid = need;
  solve_kernel<I::kFeasible><<<(unsigned)grid, 32 * kWarps, s->smem_bytes, st>>>(B, a, counter);
  CK(cudaStreamWaitEvent(st, s->ev_join[ss], 0));
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

static int ensure_host_staging(rmpc_solver* s) {
  if (s->h_xinit) return 0;
  const size_t mb = (size_t)s->max_batch;
  CK(cudaMalloc(&s->h_xinit, sizeof(double) * mb * M::NX));
  CK(cudaMalloc(&s->h_x0, sizeof(double) * mb * M::N * M::NVAR));
  CK(cudaMalloc(&s->h_params, sizeof(double) * mb * M::N * M::NPAR));
  CK(cudaMalloc(&s->h_z, sizeof(double) * mb * M::N * M::NVAR));
  CK(cudaMalloc(&s->h_action, sizeof(double) * mb * M::NU));
  CK(cudaMalloc(&s->h_res, sizeof(double) * mb * 5));
  CK(cudaMalloc(&s->h_flag, sizeof(int) * mb * 2));
  CK(cudaMallocHost(&s->p_action, sizeof(double) * mb * M::NU));
  CK(cudaMallocHost(&s->p_res, sizeof(double) * mb * 5));
  CK(cudaMallocHost(&s->p_flag, sizeof(int) * mb * 2));
  return 0;
}

int rmpc_solve_batch_host_async(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params,
                                double* z_out, double* action_out, const rmpc_info* info) {
  if (!s || !xinit || !x0 || !params) { snprintf(g_err, sizeof g_err, "null argument"); return -1; }
  if (B <= 0 || B > s->max_batch) { snprintf(g_err, sizeof g_err, "B=%d outside (0, max_batch=%d]", B, s->max_batch); return -1; }
  if (s->pend_B) { snprintf(g_err, sizeof g_err, "a host solve is already in flight on this handle: call rmpc_wait first"); return -1; }
  CK(cudaSetDevice(s->device));
  if (ensure_host_staging(s)) return -1;
  const size_t b = (size_t)B, mb = (size_t)s->max_batch;
  cudaStream_t st = s->hstream;
  CK(cudaMemcpyAsync(s->h_xinit, xinit, sizeof(double) * b * M::NX, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(s->h_x0, x0, sizeof(double) * b * M::N * M::NVAR, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(s->h_params, params, sizeof(double) * b * M::N * M::NPAR, cudaMemcpyHostToDevice, st));
  rmpc_info dinf;
  memset(&dinf, 0, sizeof dinf);
  if (info) {
    dinf.exitflag = s->h_flag; dinf.iters = s->h_flag + mb;
    dinf.obj = s->h_res; dinf.res_stat = s->h_res + mb; dinf.res_eq = s->h_res + 2 * mb;
    dinf.res_ineq = s->h_res + 3 * mb; dinf.res_comp = s->h_res + 4 * mb;
  }
  if (rmpc_solve_batch(s, B, s->h_xinit, s->h_x0, s->h_params, z_out ? s->h_z : nullptr,
                       action_out ? s->h_action : nullptr, info ? &dinf : nullptr, st))
    return -1;
  if (z_out) {
    if (!s->p_z) CK(cudaMallocHost(&s->p_z, sizeof(double) * mb * M::N * M::NVAR));
    CK(cudaMemcpyAsync(s->p_z, s->h_z, sizeof(double) * b * M::N * M::NVAR, cudaMemcpyDeviceToHost, st));
  }
  if (action_out) CK(cudaMemcpyAsync(s->p_action, s->h_action, sizeof(double) * b * M::NU, cudaMemcpyDeviceToHost, st));
  if (info) {
    CK(cudaMemcpyAsync(s->p_flag, s->h_flag, sizeof(int) * (mb + b), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(s->p_res, s->h_res, sizeof(double) * (4 * mb + b), cudaMemcpyDeviceToHost, st));
  }
  s->pend_B = B; s->u_z = z_out; s->u_action = action_out;
  memset(&s->u_info, 0, sizeof s->u_info);
  if (info) s->u_info = *info;
  return 0;
}

int rmpc_wait(rmpc_solver* s) {
  if (!s) { snprintf(g_err, sizeof g_err, "null argument"); return -1; }
  CK(cudaSetDevice(s->device));
  CK(cudaStreamSynchronize(s->hstream));
  if (!s->pend_B) return 0;
  const size_t b = (size_t)s->pend_B, mb = (size_t)s->max_batch;
  s->pend_B = 0;
  if (s->u_z) memcpy(s->u_z, s->p_z, sizeof(double) * b * M::N * M::NVAR);
  if (s->u_action) memcpy(s->u_action, s->p_action, sizeof(double) * b * M::NU);
  const rmpc_info& u = s->u_info;
  if (u.exitflag) memcpy(u.exitflag, s->p_flag, sizeof(int) * b);
  if (u.iters) memcpy(u.iters, s->p_flag + mb, sizeof(int) * b);
  double* dst[5] = {u.obj, u.res_stat, u.res_eq, u.res_ineq, u.res_comp};
  for (int i = 0; i < 5; ++i)
    if (dst[i]) memcpy(dst[i], s->p_res + i * mb, sizeof(double) * b);
  return 0;
}

int rmpc_solve_batch_host(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params,
                          double* z_out, double* action_out, const rmpc_info* info) {
  if (rmpc_solve_batch_host_async(s, B, xinit, x0, params, z_out, action_out, info)) return -1;
  return rmpc_wait(s);
}

int rmpc_shift_horizon(rmpc_solver* s, int B, const double* z, double* x0, void* stream) {
  if (!s || !z || !x0 || B <= 0) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  CK(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  const long total = (long)B * M::N * M::NVAR;
  const double* src = z;
  if (z == x0) {  // in-place: stage the plan in the (free) z workspace first
    if (B > s->max_batch) { snprintf(g_err, sizeof g_err, "B exceeds max_batch for in-place shift"); return -1; }
    if (!s->z_tmp) CK(cudaMalloc(&s->z_tmp, sizeof(double) * (size_t)s->max_batch * M::N * M::NVAR));
    CK(cudaMemcpyAsync(s->z_tmp, z, sizeof(double) * total, cudaMemcpyDeviceToDevice, st));
    src = s->z_tmp;
  }
  shift_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(total, src, x0);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

int rmpc_set_params(rmpc_solver* s, int B, double* params, const int* idx, int n_idx, const double* values,
                    int stage_varying, void* stream) {
  if (!s || !params || !idx || !values || B <= 0 || n_idx <= 0) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  for (int j = 0; j < n_idx; ++j)
    if (idx[j] < 0 || idx[j] >= M::NPAR) { snprintf(g_err, sizeof g_err, "parameter index %d out of range", idx[j]); return -1; }
  CK(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  if (n_idx > s->d_idx_cap) {
    if (s->d_idx) CK(cudaFree(s->d_idx));
    CK(cudaMalloc(&s->d_idx, sizeof(int) * n_idx));
    s->d_idx_cap = n_idx;
  }
  CK(cudaMemcpyAsync(s->d_idx, idx, sizeof(int) * n_idx, cudaMemcpyHostToDevice, st));
  const long total = (long)B * M::N * n_idx;
  scatter_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(B, n_idx, stage_varying, s->d_idx, values, params);
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
  CK(cudaSetDevice(s->device));
  const long warps = (long)B * M::N;
  fsd_kernel<<<(unsigned)((warps + 7) / 8), 256, 0, (cudaStream_t)stream>>>(a);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

}  // extern "C"
