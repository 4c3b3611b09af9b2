// rmpc_solver.cu — CUDA kernels (sm_100a) + C ABI (include/rmpc.h) of one generated solver library.
// Compiled once per model against its generated model_traits.h by MpcModel.generateSolver().
//
// Kernels
//   pack_kernel    [B][E] float64 (reference layout)  ->  [E][Bpad] lane-interleaved (SoA), tiled transpose
//   solve_kernel   one lane = one NLP (rmpc_ipm.cuh); iterates in lane-interleaved local memory
//   unpack_kernel  SoA results -> [B][N][nvar], action[B][nu], info arrays
//   shift_kernel   previous-plan warm start;  scatter_kernel  batched set* parameter writes
// There is no CPU fallback: every entry point fails loudly if CUDA fails.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <atomic>

#include "model_traits.h"
#include "rmpc_ipm.cuh"
#include "rmpc.h"

using M = ModelTraits;
using namespace rmpc;

#ifndef RMPC_BLOCK
#define RMPC_BLOCK 64
#endif

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

struct rmpc_solver {
  int device;
  int max_batch;
  long bpad;
  rmpc_options opt;
  double* xinit_soa;  // [NX][bpad]
  double* z_soa;      // [N*NVAR][bpad]  (in: initial guess, out: solution)
  double* p_soa;      // [N*NPAR][bpad]
  double* res_soa;    // [5][bpad]  obj, stat, eq, ineq, comp
  int* flag_soa;      // [2][bpad]  exitflag, iters
  int* d_idx;         // scratch for rmpc_set_params
  int d_idx_cap;
  // device staging for the host entry point
  double *h_xinit, *h_x0, *h_params, *h_z, *h_action, *h_res;
  int* h_flag;
};

// ---------------------------------------------------------------------------------------------
// [B][E] -> [E][bpad]  (32x32 tiles through shared memory: coalesced on both sides)
__global__ void pack_kernel(int B, long bpad, int E, const double* __restrict__ src, double* __restrict__ dst) {
  __shared__ double tile[32][33];
  const int b0 = blockIdx.x * 32, e0 = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int b = b0 + r, e = e0 + threadIdx.x;
    if (b < B && e < E) tile[r][threadIdx.x] = src[(long)b * E + e];
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int e = e0 + r, b = b0 + threadIdx.x;
    if (b < B && e < E) dst[(long)e * bpad + b] = tile[threadIdx.x][r];
  }
}

// [E][bpad] -> [B][E]
__global__ void unpack_kernel(int B, long bpad, int E, const double* __restrict__ src, double* __restrict__ dst) {
  __shared__ double tile[32][33];
  const int b0 = blockIdx.x * 32, e0 = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int e = e0 + r, b = b0 + threadIdx.x;
    if (b < B && e < E) tile[r][threadIdx.x] = src[(long)e * bpad + b];
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int b = b0 + r, e = e0 + threadIdx.x;
    if (b < B && e < E) dst[(long)b * E + e] = tile[threadIdx.x][r];
  }
}

// action + info (all coalesced reads of SoA; small strided writes)
__global__ void finish_kernel(int B, long bpad, int control_mode, const double* __restrict__ z_soa,
                              const double* __restrict__ res_soa, const int* __restrict__ flag_soa, double* action,
                              rmpc_info info) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  if (action) {
    for (int j = 0; j < M::NU; ++j) {
      const int e = control_mode == 0 ? (M::NX + j) : (M::NVAR + M::NVAR - 2 * M::NU + j);
      action[(long)b * M::NU + j] = z_soa[(long)e * bpad + b];
    }
  }
  if (info.exitflag) info.exitflag[b] = flag_soa[b];
  if (info.iters) info.iters[b] = flag_soa[bpad + b];
  if (info.obj) info.obj[b] = res_soa[b];
  if (info.res_stat) info.res_stat[b] = res_soa[bpad + b];
  if (info.res_eq) info.res_eq[b] = res_soa[2 * bpad + b];
  if (info.res_ineq) info.res_ineq[b] = res_soa[3 * bpad + b];
  if (info.res_comp) info.res_comp[b] = res_soa[4 * bpad + b];
}

template <class TF>
__global__ void __launch_bounds__(RMPC_BLOCK)
solve_kernel(int B, long bpad, const double* __restrict__ xinit_soa, double* __restrict__ z_soa,
             const double* __restrict__ p_soa, double* __restrict__ res_soa, int* __restrict__ flag_soa, Options opt) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  Work<M, double> w;
  for (int k = 0; k < M::N; ++k)
    RMPC_UNROLL
    for (int i = 0; i < M::NVAR; ++i) w.z[k][i] = z_soa[(long)(k * M::NVAR + i) * bpad + b];
  double xi[M::NX];
  RMPC_UNROLL
  for (int i = 0; i < M::NX; ++i) xi[i] = xinit_soa[(long)i * bpad + b];
  ParamView<double> pv{p_soa + b, (long)M::NPAR * bpad, bpad};
  Info info;
  solve_one<M, double, TF>(w, xi, pv, opt, info);
  for (int k = 0; k < M::N; ++k)
    RMPC_UNROLL
    for (int i = 0; i < M::NVAR; ++i) z_soa[(long)(k * M::NVAR + i) * bpad + b] = w.z[k][i];
  flag_soa[b] = info.exitflag;
  flag_soa[bpad + b] = info.iters;
  res_soa[b] = info.obj;
  res_soa[bpad + b] = info.res_stat;
  res_soa[2 * bpad + b] = info.res_eq;
  res_soa[3 * bpad + b] = info.res_ineq;
  res_soa[4 * bpad + b] = info.res_comp;
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
  Options c;
  c.max_iter = o.max_iter; c.hessian = o.hessian;
  c.tol_stat = o.tol_stat; c.tol_eq = o.tol_eq; c.tol_ineq = o.tol_ineq; c.tol_comp = o.tol_comp;
  c.mu0 = o.mu0; c.mu_min = o.mu_min; c.reg = o.reg;
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
}

static int free_all(rmpc_solver* s) {
  cudaFree(s->xinit_soa); cudaFree(s->z_soa); cudaFree(s->p_soa); cudaFree(s->res_soa); cudaFree(s->flag_soa);
  cudaFree(s->d_idx); cudaFree(s->h_xinit); cudaFree(s->h_x0); cudaFree(s->h_params); cudaFree(s->h_z);
  cudaFree(s->h_action); cudaFree(s->h_res); cudaFree(s->h_flag);
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
  s->bpad = ((long)max_batch + 31) / 32 * 32;
  rmpc_default_options(&s->opt);
  auto fail = [&](const char* what, cudaError_t err) -> rmpc_solver* {
    snprintf(g_err, sizeof g_err, "%s: %s", what, cudaGetErrorString(err));
    free_all(s);
    delete s;
    return nullptr;
  };
  if ((e = cudaSetDevice(device)) != cudaSuccess) return fail("cudaSetDevice", e);
  const size_t bp = (size_t)s->bpad;
  if ((e = cudaMalloc(&s->xinit_soa, sizeof(double) * M::NX * bp)) != cudaSuccess) return fail("cudaMalloc", e);
  if ((e = cudaMalloc(&s->z_soa, sizeof(double) * M::N * M::NVAR * bp)) != cudaSuccess) return fail("cudaMalloc", e);
  if ((e = cudaMalloc(&s->p_soa, sizeof(double) * M::N * M::NPAR * bp)) != cudaSuccess) return fail("cudaMalloc", e);
  if ((e = cudaMalloc(&s->res_soa, sizeof(double) * 5 * bp)) != cudaSuccess) return fail("cudaMalloc", e);
  if ((e = cudaMalloc(&s->flag_soa, sizeof(int) * 2 * bp)) != cudaSuccess) return fail("cudaMalloc", e);
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
  if (o->precision < 0 || o->precision > 1 || o->hessian < 0 || o->hessian > 1 || o->control_mode < 0 || o->control_mode > 1) {
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

static int launch_pack(int B, long bpad, int E, const double* src, double* dst, cudaStream_t st) {
  dim3 grid((B + 31) / 32, (E + 31) / 32), block(32, 8);
  pack_kernel<<<grid, block, 0, st>>>(B, bpad, E, src, dst);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

int rmpc_solve_batch(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params, double* z_out,
                     double* action_out, const rmpc_info* info, void* stream) {
  if (!s || !xinit || !x0 || !params) { snprintf(g_err, sizeof g_err, "null argument"); return -1; }
  if (B <= 0 || B > s->max_batch) { snprintf(g_err, sizeof g_err, "B=%d outside (0, max_batch=%d]", B, s->max_batch); return -1; }
  CK(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  const long bpad = s->bpad;
  if (launch_pack(B, bpad, M::NX, xinit, s->xinit_soa, st)) return -1;
  if (launch_pack(B, bpad, M::N * M::NVAR, x0, s->z_soa, st)) return -1;
  if (launch_pack(B, bpad, M::N * M::NPAR, params, s->p_soa, st)) return -1;
  const Options c = to_core(s->opt);
  const int grid = (B + RMPC_BLOCK - 1) / RMPC_BLOCK;
  if (s->opt.precision == 0)
    solve_kernel<double><<<grid, RMPC_BLOCK, 0, st>>>(B, bpad, s->xinit_soa, s->z_soa, s->p_soa, s->res_soa, s->flag_soa, c);
  else
    solve_kernel<float><<<grid, RMPC_BLOCK, 0, st>>>(B, bpad, s->xinit_soa, s->z_soa, s->p_soa, s->res_soa, s->flag_soa, c);
  g_launches++;
  CK(cudaGetLastError());
  if (z_out) {
    const int E = M::N * M::NVAR;
    dim3 g2((B + 31) / 32, (E + 31) / 32), blk(32, 8);
    unpack_kernel<<<g2, blk, 0, st>>>(B, bpad, E, s->z_soa, z_out);
    g_launches++;
    CK(cudaGetLastError());
  }
  rmpc_info inf;
  memset(&inf, 0, sizeof inf);
  if (info) inf = *info;
  if (action_out || info) {
    finish_kernel<<<(B + 127) / 128, 128, 0, st>>>(B, bpad, s->opt.control_mode, s->z_soa, s->res_soa, s->flag_soa, action_out, inf);
    g_launches++;
    CK(cudaGetLastError());
  }
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
  return 0;
}

int rmpc_solve_batch_host(rmpc_solver* s, int B, const double* xinit, const double* x0, const double* params,
                          double* z_out, double* action_out, const rmpc_info* info) {
  if (!s || !xinit || !x0 || !params) { snprintf(g_err, sizeof g_err, "null argument"); return -1; }
  if (B <= 0 || B > s->max_batch) { snprintf(g_err, sizeof g_err, "B=%d outside (0, max_batch=%d]", B, s->max_batch); return -1; }
  CK(cudaSetDevice(s->device));
  if (ensure_host_staging(s)) return -1;
  const size_t b = (size_t)B, mb = (size_t)s->max_batch;
  cudaStream_t st = 0;
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
  if (z_out) CK(cudaMemcpyAsync(z_out, s->h_z, sizeof(double) * b * M::N * M::NVAR, cudaMemcpyDeviceToHost, st));
  if (action_out) CK(cudaMemcpyAsync(action_out, s->h_action, sizeof(double) * b * M::NU, cudaMemcpyDeviceToHost, st));
  if (info) {
    if (info->exitflag) CK(cudaMemcpyAsync(info->exitflag, dinf.exitflag, sizeof(int) * b, cudaMemcpyDeviceToHost, st));
    if (info->iters) CK(cudaMemcpyAsync(info->iters, dinf.iters, sizeof(int) * b, cudaMemcpyDeviceToHost, st));
    if (info->obj) CK(cudaMemcpyAsync(info->obj, dinf.obj, sizeof(double) * b, cudaMemcpyDeviceToHost, st));
    if (info->res_stat) CK(cudaMemcpyAsync(info->res_stat, dinf.res_stat, sizeof(double) * b, cudaMemcpyDeviceToHost, st));
    if (info->res_eq) CK(cudaMemcpyAsync(info->res_eq, dinf.res_eq, sizeof(double) * b, cudaMemcpyDeviceToHost, st));
    if (info->res_ineq) CK(cudaMemcpyAsync(info->res_ineq, dinf.res_ineq, sizeof(double) * b, cudaMemcpyDeviceToHost, st));
    if (info->res_comp) CK(cudaMemcpyAsync(info->res_comp, dinf.res_comp, sizeof(double) * b, cudaMemcpyDeviceToHost, st));
  }
  CK(cudaStreamSynchronize(st));
  return 0;
}

int rmpc_shift_horizon(rmpc_solver* s, int B, const double* z, double* x0, void* stream) {
  if (!s || !z || !x0 || B <= 0) { snprintf(g_err, sizeof g_err, "bad argument"); return -1; }
  CK(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  const long total = (long)B * M::N * M::NVAR;
  const double* src = z;
  if (z == x0) {  // in-place: stage the plan in the (free) z workspace first
    if (B > s->max_batch) { snprintf(g_err, sizeof g_err, "B exceeds max_batch for in-place shift"); return -1; }
    CK(cudaMemcpyAsync(s->z_soa, z, sizeof(double) * total, cudaMemcpyDeviceToDevice, st));
    src = s->z_soa;
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

}  // extern "C"
