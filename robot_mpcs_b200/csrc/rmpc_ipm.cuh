// rmpc_ipm.cuh — one nonlinear-MPC solve, cooperatively by a group of G lanes.
//
//   min  sum_k J(z_k,p_k)   s.t.  x_1 = xinit,  x_{k+1} = Phi(x_k,u_k),  h(z_k,p_k) >= 0
//
// This is the component that replaces the FORCESPRO-generated solver the reference calls at
// planner/mpcPlanner.py:262.
//
// Algorithm: primal-dual interior point (log barrier) that stays strictly inside h + shift > 0, where the shift is
// zero for every row that is strictly feasible at the initial guess and is driven to zero otherwise; the dynamics
// are treated as equalities (infeasible start, l1 merit).  Per iteration
//   A  stage-parallel   lane k evaluates stage k: cost gradient, Hessian (exact Lagrangian or Gauss-Newton),
//                       inequality rows + Jacobians, dynamics value + sensitivities + defect
//   B  backward sweep   Riccati recursion over the horizon; the G lanes share the small dense products of every
//                       stage through shared memory (4 group barriers per stage); the adjoint recursion that gives
//                       the stationarity residual rides along as an extra right-hand side.  The barrier parameter
//                       enters only the gradient, so the vector recursion is carried twice (mu-independent part,
//                       coefficient of mu) and mu is chosen after the sweep from the residuals it measured.
//   C  roll-out         every lane runs the (cheap) forward recursion and keeps the step of the stage it owns
//   D  line search      stage-parallel trial evaluations, group reductions, fraction-to-boundary on the true
//                       nonlinear rows and Armijo on the l1 merit
// With G = 1 the same code is the sequential host twin (oracle/cpu_twin.cpp); with G = 32 a warp solves one
// problem (rmpc_solver.cu).
#pragma once
#include "rmpc_model.cuh"
#if defined(RMPC_TRACE)
#include <stdio.h>
#endif

// Loops over the stages a lane owns (SLOTS = ceil(N / G)) stay rolled on the device.  With one stage per lane the loop has
// one trip and disappears; with two (N > 32) the code of every stage-parallel phase exists once instead of twice, at the
// price of indexing the per-lane state dynamically (local memory instead of registers that spilled anyway).  Measured on
// the N = 40 model (B200, r2m): kernel 167 -> 110 KB, 366 k -> 424 k solves/s.  -DRMPC_UNROLL_SLOTS restores the unrolled form.
#if defined(__CUDA_ARCH__) && !defined(RMPC_UNROLL_SLOTS)
#define RMPC_UNROLL_SL _Pragma("unroll 1")
#else
#define RMPC_UNROLL_SL RMPC_UNROLL
#endif

#ifndef RMPC_SOC_DEFAULT
#define RMPC_SOC_DEFAULT 1   // second-order correction on (0: A/B builds)
#endif

namespace rmpc {

struct Options {
  int max_iter;        // iteration cap (FORCESPRO default 200)
  int hessian;         // 0 Gauss-Newton, 1 exact Lagrangian Hessian with GN fallback when not positive definite
  double tol_stat;     // scaled stationarity
  double tol_eq;       // dynamics defect
  double tol_ineq;     // remaining shift of rows that started infeasible
  double tol_comp;     // complementarity s*lambda
  double mu0;          // initial barrier parameter
  double mu_min;
  double reg;          // Levenberg term added to the diagonal of every stage Hessian
  double delta;        // rows with h < delta at the initial guess start shifted to s = delta
  double kappa;        // multipliers are kept above mu/(kappa s)
  double tau_min;      // fraction-to-boundary parameter tau = max(tau_min, 1 - mu)
  double lm0, lm_up, lm_down, lm_trigger;  // adaptive damping: start value, growth, decay, backtracks that trigger growth
  double gn_hold_max;                       // longest run of iterations that skip the exact-Hessian attempt after it failed (0: always try)
  double stall_window;                      // give up (-7) when the optimality error has not dropped 10x within this many iterations (0 = never)
  double creep_trigger, creep_factor;      // consecutive boundary-cut full steps that trigger a barrier increase, and its factor
  double soc;                               // 1: second-order correction of the step when two iterations in a row leave the interior of a curved row
  double ls_optimistic;                     // 1: the first trial step is capped by the (linear) limit rows only; the nonlinear rows are checked on their true values
};

RMPC_HD Options default_options() {
  Options o;
  o.max_iter = 200; o.hessian = 1;
  o.tol_stat = 1e-6; o.tol_eq = 1e-8; o.tol_ineq = 1e-8; o.tol_comp = 1e-8;
  o.mu0 = 1.0; o.mu_min = 1e-9; o.reg = 0.0; o.delta = 1e-2; o.kappa = 1e10; o.tau_min = 0.99; o.lm0 = 1e-2; o.lm_up = 4.0; o.lm_down = 0.25; o.lm_trigger = 3; o.creep_trigger = 3; o.creep_factor = 10; o.stall_window = 0; o.gn_hold_max = 4; o.ls_optimistic = 0; o.soc = RMPC_SOC_DEFAULT;
  return o;
}

struct Info {
  int exitflag;   // 1 optimal, 0 iteration limit, -5 factorisation, -6 NaN/Inf, -7 no progress / infeasible, -8 bad initial guess
  int iters;
  double obj, res_stat, res_eq, res_ineq, res_comp, mu;
};

// ------------------------------------------------------------------------------------------------
// Lane group: G lanes that own interleaved stages (lane l owns stages l, l+G, ...) and share `ProblemShared`.
template <int G> struct Group;
template <> struct Group<1> {
  int lane;
  RMPC_HD void sync() const {}
  template <class T> RMPC_HD T sum(T v) const { return v; }
  template <class T> RMPC_HD T max(T v) const { return v; }
  template <class T> RMPC_HD T min(T v) const { return v; }
  RMPC_HD bool any(bool p) const { return p; }
  template <class T> RMPC_HD T bcast(T v, int) const { return v; }
  RMPC_HD void align() const {}
};
#if defined(__CUDACC__)
// double reductions out of line: ~20 call sites x 15 instructions otherwise
__device__ __noinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __noinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { const double w = __shfl_xor_sync(0xffffffffu, v, o); v = w > v ? w : v; }
  return v;
}
__device__ __noinline__ float warp_sumf(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __noinline__ float warp_maxf(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
template <> struct Group<32> {
  int lane;
  __device__ __forceinline__ void sync() const { __syncwarp(); }
  __device__ __forceinline__ double sum(double v) const { return warp_sum(v); }
  __device__ __forceinline__ float sum(float v) const { return warp_sumf(v); }
  __device__ __forceinline__ float bcast(float v, int src) const { return __shfl_sync(0xffffffffu, v, src); }
  __device__ __forceinline__ double max(double v) const { return warp_max(v); }
  __device__ __forceinline__ double min(double v) const { return -warp_max(-v); }
  __device__ __forceinline__ float max(float v) const { return warp_maxf(v); }
  __device__ __forceinline__ float min(float v) const { return -max(-v); }
  __device__ __forceinline__ bool any(bool p) const { return __any_sync(0xffffffffu, p) != 0; }
  __device__ __forceinline__ double bcast(double v, int src) const { return __shfl_sync(0xffffffffu, v, src); }
  // Block-wide rendezvous at the top of every solver iteration: the warps of a block (one problem each) then walk
  // through the same code at about the same time, so one instruction-cache fill serves all of them (the kernel's
  // L1.5 instruction-request rate is at 89 % of peak).  Off by default: see RMPC_WARPS_FIXED in rmpc_solver.cu.  A warp
  // without work keeps answering with 1 until every warp does.
  __device__ __forceinline__ void align() const {
#if RMPC_WARPS > 1
    __syncthreads_and(0);
#endif
  }
};
#endif

// Parameter access: p(k, j) of one problem.  TS is the type the parameters are stored in (the caller's float64 arrays), T
// the type the solve computes in.
template <class T, class TS = T>
struct ParamView {
  const TS* base;
  long stage_stride;  // elements between consecutive stages
  long elem_stride;   // elements between consecutive parameters
  RMPC_HD T get(int k, int j) const {
#if defined(__CUDA_ARCH__) && defined(RMPC_PARAM_LOAD)
    return T(RMPC_PARAM_LOAD(base + k * stage_stride + j * elem_stride));   // experiment: __ldcg / __ldcs / __ldg
#else
    return T(base[k * stage_stride + j * elem_stride]);
#endif
  }
};

template <class M, class T, class PV>
RMPC_HD void load_params(const PV& pv, int k, T* p) {
  RMPC_UNROLL
  for (int j = 0; j < M::NPAR; ++j) p[j] = pv.get(k, j);
}

// ------------------------------------------------------------------------------------------------
// Shared (per problem) storage.  GPU: shared memory of the warp that owns the problem; host: heap.
template <class M, class = void> struct AsksForSharedState { static constexpr bool value = false; };
template <class M> struct AsksForSharedState<M, decltype(void(M::STATE_IN_SHARED))> { static constexpr bool value = M::STATE_IN_SHARED; };
template <class M> struct Dims {
  // The Riccati algebra runs on the variables the Newton system couples (DynS::NXC states + NU inputs, "compact" indices)
  static constexpr int RX = DynS<M>::NXC, RV = RX + M::NU;
  static constexpr int NDS = (M::BASE == 0) ? M::NQ : RV;          // variables coupled by the dense part of the stage Hessian
  static constexpr int NQD = NDS * NDS;                            // stored in full, column by column: a lane reads its column with compile-time offsets
  static constexpr int NHD = RV - NDS > 0 ? RV - NDS : 1;          // variables with a diagonal-only Hessian
  static constexpr int NVA = RV + 3;                               // [A B]-sized blocks carry 3 extra columns: g0, g1, adjoint
  static constexpr int NKK = M::NU * RX + 2 * M::NU;               // feedback gain K and the two feed-forward vectors
  static constexpr int NPRE = NQD + NHD + 2 * RV;                  // Hessian + g1 + gl: dead once phase 1 of the sweep has read them
  static constexpr int OFF0 = NPRE > NKK ? NPRE : NKK;             // offset of g0 (live until the line search) in StageShared::Qs
  static constexpr int OFF1 = OFF0 - 2 * RV, OFFL = OFF0 - RV;
  static constexpr int LDF = (NVA % 2) ? NVA : NVA + 1;            // odd row stride of F / PAB: row- and column-wise accesses are conflict-free
  static constexpr int NQS0 = OFF0 + RV > 2 * M::NH ? OFF0 + RV : 2 * M::NH;
  // In the stage-parallel phases lane k works on stage k, so the lanes of a warp access shared memory with the stage
  // stride.  The stride is kept = 4 (mod 8) 32-bit words (float64: 2 (mod 4) doubles): 16-byte aligned, conflict-free for
  // 128-bit accesses, two-way for 64-bit ones.  A multiple of 32 words puts every lane on the same bank (measured: 45 % of
  // all shared wavefronts were conflicts with a 112-double stage, 6 % with 114).
  // The multipliers and slacks of a stage live in shared memory (2 NH values per stage) instead of in the registers of the
  // lane that owns the stage, unless the horizon needs two stages per lane (N > 32), where the extra shared memory costs
  // more in resident problems than it saves.  B200, r2n/r2o A/B (registers -> shared): panda 436 k -> 489 k solves/s and
  // B = 1 latency 1.27 -> 0.88 ms; pointRobot 3.33 M -> 3.58 M; boxer N = 30 743 k -> 758 k, latency 1.11 -> 0.90 ms;
  // boxer N = 40 424 k -> 399 k (kept in registers / local memory).  -DRMPC_STATE_SH=0/1 forces either form.
  // A traits class can ask for the shared-memory form with `static constexpr bool STATE_IN_SHARED = true` (the single-launch
  // kernel for batches of <= 32 problems does: occupancy is no concern there, and B = 1 of the N = 40 model runs 1.28 -> 0.99 ms).
#if defined(RMPC_STATE_SH)
  static constexpr bool STATE_SH = RMPC_STATE_SH != 0;
#else
  static constexpr bool STATE_SH = M::N <= 32 || M::NH >= 24 || AsksForSharedState<M>::value;
#endif
  static constexpr int NLS = STATE_SH ? 2 * M::NH : 2;
  template <class T> static constexpr int nqs() {
    constexpr int per16 = 16 / int(sizeof(T));                       // elements per 16 bytes
    constexpr int q1 = (NQS0 + per16 - 1) / per16 * per16;            // what follows Qs stays 16-byte aligned
    constexpr int stage = 2 * M::NVAR + q1 + DynS<M>::NABV + M::NX + 2 + NLS;
    constexpr int words = (stage + per16 - 1) / per16 * per16 * int(sizeof(T)) / 4;   // alignas(16) rounds the struct up
    return q1 + ((words % 8 == 4) ? 0 : per16);
  }
};

template <class M, class T>
struct alignas(16) StageShared {
  T z[M::NVAR];
  T dz[M::NVAR];             // Newton step of this stage (written by the roll-out)
  // [Qd | Hd | .. | g1 | gl | g0] over the compact variables: dense Hessian part (NDS x NDS, symmetric, both triangles),
  // diagonal entries of the variables NDS..RV-1, g1 = -G'(1/s) (coefficient of mu in the barrier gradient),
  // gl = cost gradient - G' lambda, g0 = cost gradient.
  // Everything before g0 is dead once phase 1 of the sweep has consumed it; K | kff0 | kff1 then reuse that space.
  // After the roll-out the same array is scratch for the line search: row values of the trial point in [0, NH), multiplier
  // steps in [NH, 2 NH).
  T Qs[Dims<M>::template nqs<T>()];
  T ABv[DynS<M>::NABV];      // z-dependent entries of [A B]
  T d[M::NX];                // Phi(z_k) - x_{k+1}
  T nu01[2];                 // adjoint of (X+, Y+) from the previous sweep (diff-drive curvature)
  T ls[Dims<M>::NLS];        // STATE_SH models: lambda[NH] | s[NH] of this stage (2 unused entries otherwise)
  RMPC_HD T g0(int c) const { return Qs[Dims<M>::OFF0 + c]; }
  RMPC_HD T& K(int l, int c) { return Qs[l * Dims<M>::RX + c]; }
  RMPC_HD T K(int l, int c) const { return Qs[l * Dims<M>::RX + c]; }
  RMPC_HD T& kff0(int l) { return Qs[M::NU * Dims<M>::RX + l]; }
  RMPC_HD T kff0(int l) const { return Qs[M::NU * Dims<M>::RX + l]; }
  RMPC_HD T& kff1(int l) { return Qs[M::NU * Dims<M>::RX + M::NU + l]; }
  RMPC_HD T kff1(int l) const { return Qs[M::NU * Dims<M>::RX + M::NU + l]; }
};

template <class M, class T>
struct ProblemShared {
  static_assert((sizeof(StageShared<M, T>) / 4) % 8 == 4, "stage stride must be 4 (mod 8) words");
  StageShared<M, T> st[M::N];
  T PAB[Dims<M>::RX * Dims<M>::LDF];   // P_{k+1} [A B | d-term | p1 | nu] of the stage being processed (compact variables)
  T F[Dims<M>::RV * Dims<M>::LDF];     // [Q | g0 | g1 | gl] + [A B]' PAB, every entry valid (the matrix part is symmetric up to rounding)
};

template <class T>
struct Red {  // what phase A + the sweep measure at the current iterate
  T f, theta, logs, res_eq, res_stat, comp_min, comp_max, mult_sum, shift_max, s_min;
  bool pole, ok, finite;
};

// ------------------------------------------------------------------------------------------------
template <class M, class T, int G, class TS = T>
struct Ipm {
  static constexpr int NX = M::NX, NU = M::NU, NV = M::NVAR, NH = M::NH, N = M::N, NQ = M::NQ;
  static constexpr int NGEO1 = M::NGEO > 0 ? M::NGEO : 1;
  static constexpr int SLOTS = (N + G - 1) / G;
  static constexpr int NVA = Dims<M>::NVA, NDS = Dims<M>::NDS, RX = Dims<M>::RX, RV = Dims<M>::RV;
  using DS = DynS<M>;
  static_assert(DS::rows_avoid_inert(), "a limit row touches a state that the Newton system leaves out");

  const Group<G>& g;
  ProblemShared<M, T>& S;
  const ParamView<T, TS>& pv;
  const Options& opt;

  // per-lane state of the stages this lane owns
  static constexpr bool STATE_SH = Dims<M>::STATE_SH;
  T lam_r[SLOTS][STATE_SH ? 1 : NH], s_r[SLOTS][STATE_SH ? 1 : NH];
  RMPC_HD T& LAM(int sl, int k, int r) { if constexpr (STATE_SH) return S.st[k].ls[r]; else return lam_r[sl][r]; }
  RMPC_HD T& SLK(int sl, int k, int r) { if constexpr (STATE_SH) return S.st[k].ls[NH + r]; else return s_r[sl][r]; }
  float sh[SLOTS][NH];  // shifts of the rows that started infeasible; referenced only by the <SH = true> instantiations below
  T Gg[SLOTS][NGEO1][NQ];
  // Compiled in only for the diff-drive base, whose geometric rows depend on the heading (the end-effector point swings with
  // theta): B200 A/B r2t, correction on/off: supermarket 407 k -> 461 k solves/s and 99.79 -> 99.98 % converged, boxer N = 30 flat;
  // panda -3 %, pointRobot -2 % (code size, never used: prismatic joints / rows that are rarely cut).
  static constexpr bool SOCC = M::BASE == 1 && M::NGEO > 0;
  float soc[SLOTS][NGEO1];   // second-order correction of the geometric rows: h(z + dz) - h - G dz of the trial step that failed (iterate)

  RMPC_HD Ipm(const Group<G>& g_, ProblemShared<M, T>& S_, const ParamView<T, TS>& pv_, const Options& o_) : g(g_), S(S_), pv(pv_), opt(o_) {}


  // ---- A: stage-parallel derivatives ----------------------------------------------------------
  template <bool SH>
  RMPC_HD void derivs(const bool exact, const T reg, Red<T>& o, const bool soc_on = false) {
    LogAcc<T> lacc;
    T f = T(0), theta = T(0), res_eq = T(0), msum = T(0), smin = T(1e300);
    float cmin = 3e38f, cmax = 0.0f, shmax = 0.0f;  // only compared with tolerances: single precision is enough
    bool pole = false;
    RMPC_UNROLL_SL
    for (int sl = 0; sl < SLOTS; ++sl) {
      const int k = g.lane + G * sl;
      if (k < N) {
        StageShared<M, T>& st = S.st[k];
        T p[M::NPAR];
        load_params<M, T>(pv, k, p);
        T zk[NV];
        RMPC_UNROLL
        for (int i = 0; i < NV; ++i) zk[i] = st.z[i];
        T sig[NH], y1[NH];
        T rho[NGEO1], kappa[NGEO1];
        RMPC_UNROLL
        for (int r = 0; r < NH; ++r) {
          const T sr = SLK(sl, k, r), lr = LAM(sl, k, r);
          y1[r] = trcp(sr);
          sig[r] = lr * y1[r];
          lacc.mul(sr);
          if ((r % 12 == 11 && r + 1 < NH) || (SLOTS > 1 && r == NH - 1)) lacc.flush();
          const float cr = float(sr * lr);
          cmax = fmaxf(cmax, cr);
          cmin = fminf(cmin, cr);
          msum += lr;
          if constexpr (SH) shmax = fmaxf(shmax, sh[sl][r]);
        }
        if constexpr (SH) {  // a slack this small next to its shift cannot recover
          RMPC_UNROLL
          for (int r = 0; r < NH; ++r)
            if (sh[sl][r] > 0.0f) { smin = tmin(smin, SLK(sl, k, r) - T(1e-6f * sh[sl][r])); theta += T(sh[sl][r]); }
        }
        RMPC_UNROLL
        for (int r = 0; r < M::NGEO; ++r) { rho[r] = sig[M::geo_row(r)]; kappa[r] = -LAM(sl, k, M::geo_row(r)); }
        StageDer<M, T> sd;
        stage_derivatives<M, T>(zk, p, rho, kappa, exact, sd);
        f += sd.f;
        pole = pole || sd.pole;
        RMPC_UNROLL
        for (int r = 0; r < M::NGEO; ++r)
          RMPC_UNROLL
          for (int i = 0; i < NQ; ++i) Gg[sl][r][i] = sd.Gg[r][i];
        // gradients: qt0 = g, qt1 = -G'(1/s), gl = g - G'lambda
        T q1[NV], glv[NV];
        RMPC_UNROLL
        for (int i = 0; i < NV; ++i) { q1[i] = T(0); glv[i] = sd.g[i]; }
        RMPC_UNROLL
        for (int r = 0; r < M::NGEO; ++r) {
          const T a1 = y1[M::geo_row(r)], al = LAM(sl, k, M::geo_row(r));
          RMPC_UNROLL
          for (int i = 0; i < NQ; ++i) { q1[i] -= sd.Gg[r][i] * a1; glv[i] -= sd.Gg[r][i] * al; }
        }
        RMPC_UNROLL
        for (int r = 0; r < M::NLIM; ++r) {
          q1[M::lim_var(r)] -= T(M::lim_sign(r)) * y1[M::lim_row(r)];
          glv[M::lim_var(r)] -= T(M::lim_sign(r)) * LAM(sl, k, M::lim_row(r));
          sd.Hd[M::lim_var(r)] += sig[M::lim_row(r)];
        }
        constexpr int NQD = Dims<M>::NQD;
        RMPC_UNROLL
        for (int c = 0; c < RV; ++c) {
          const int i = DS::full(c);
          st.Qs[Dims<M>::OFF0 + c] = sd.g[i]; st.Qs[Dims<M>::OFF1 + c] = q1[i]; st.Qs[Dims<M>::OFFL + c] = glv[i];
        }
        if (!SH && SOCC && soc_on) {
          // second-order correction: the step is asked for ds = g'dz + c, c = curvature defect of the trial step that
          // failed; like a shift of -c it adds (lambda/s) c g to the mu-free right-hand side
          RMPC_UNROLL
          for (int r = 0; r < M::NGEO; ++r) {
            const T push = sig[M::geo_row(r)] * T(soc[sl][r]);
            RMPC_UNROLL
            for (int i = 0; i < NQ; ++i) st.Qs[Dims<M>::OFF0 + i] += sd.Gg[r][i] * push;
          }
        }
        if constexpr (SH) {
          // A shifted row is treated like a violated equality h + shift(1 - alpha) >= 0: the Newton step is asked to
          // remove the whole shift, ds = g'dz - shift, which adds (lambda/s) shift g to the mu-free right-hand side.
          T q0[NV];
          RMPC_UNROLL
          for (int i = 0; i < NV; ++i) q0[i] = sd.g[i];
          RMPC_UNROLL
          for (int r = 0; r < M::NGEO; ++r) {
            const T push = sig[M::geo_row(r)] * T(sh[sl][M::geo_row(r)]);
            RMPC_UNROLL
            for (int i = 0; i < NQ; ++i) q0[i] -= sd.Gg[r][i] * push;
          }
          RMPC_UNROLL
          for (int r = 0; r < M::NLIM; ++r) q0[M::lim_var(r)] -= T(M::lim_sign(r)) * sig[M::lim_row(r)] * T(sh[sl][M::lim_row(r)]);
          RMPC_UNROLL
          for (int c = 0; c < RV; ++c) st.Qs[Dims<M>::OFF0 + c] = q0[DS::full(c)];
        }
        // dense part of the Hessian (NDS x NDS, both triangles, column by column) with the diagonal folded in; q comes
        // first in both variable orders, so the q-block keeps its indices
        T Qd[NQD];
        RMPC_UNROLL
        for (int i = 0; i < NQD; ++i) Qd[i] = T(0);
        RMPC_UNROLL
        for (int a = 0; a < NQ; ++a)
          RMPC_UNROLL
          for (int b = 0; b < NQ; ++b) Qd[b * NDS + a] = sd.Hqq[a * NQ + b];
        RMPC_UNROLL
        for (int c = 0; c < RV; ++c) {
          const int i = DS::full(c);
          if (c < NDS) Qd[(c < NDS ? c : 0) * NDS + (c < NDS ? c : 0)] += sd.Hd[i] + reg;
          else st.Qs[NQD + (c < NDS ? 0 : c - NDS)] = sd.Hd[i] + reg;
        }
        if (k < N - 1) {
          T xn[NX], ABv[DS::NABV], Hc[5][5];
          dyn_eval<M, T>(zk, xn, ABv, exact, st.nu01[0], st.nu01[1], Hc);
          RMPC_UNROLL
          for (int i = 0; i < NX; ++i) {
            const T di = xn[i] - S.st[k + 1].z[i];
            st.d[i] = di;
            theta += tabs(di);
            res_eq = tmax(res_eq, tabs(di));
          }
          RMPC_UNROLL
          for (int i = 0; i < DS::NVR * DS::NVC; ++i) st.ABv[i] = ABv[i];
          if (M::BASE == 1 && exact) {
            RMPC_UNROLL
            for (int a = 0; a < 5; ++a)
              RMPC_UNROLL
              for (int b = 0; b < 5; ++b) {
                constexpr int dummy = 0;
                const int ca = DS::comp(DS::vc(a)) < 0 ? dummy : DS::comp(DS::vc(a)), cb = DS::comp(DS::vc(b)) < 0 ? dummy : DS::comp(DS::vc(b));
                Qd[(cb < NDS ? cb : 0) * NDS + (ca < NDS ? ca : 0)] += Hc[a][b];
              }
          }
        }
        RMPC_UNROLL
        for (int i = 0; i < NQD; ++i) st.Qs[i] = Qd[i];
      }
    }
    o.f = g.sum(f); o.theta = g.sum(theta); o.logs = g.sum(lacc.value()); o.res_eq = g.max(res_eq);
    o.comp_min = T(g.min(cmin)); o.comp_max = T(g.max(cmax)); o.mult_sum = g.sum(msum); 
    if constexpr (SH) { o.shift_max = T(g.max(shmax)); o.s_min = g.min(smin); }
    else { o.shift_max = T(0); o.s_min = T(1); }
    o.pole = g.any(pole);
    g.sync();
  }

  // ---- B: backward Riccati sweep, shared by the lanes -------------------------------------------
  // Three barriers per stage, over the RX coupled states and NU inputs (compact indices, DynS::full/comp).  P_{k+1}
  // itself is never stored: the lanes that own its rows keep them in registers just long enough to form PAB for the next
  // (earlier) stage.
  RMPC_HD void sweep(Red<T>& o) {
    constexpr int LDF = Dims<M>::LDF, NQD = Dims<M>::NQD;
    T res_stat = T(0), nsum = T(0);
    bool ok = true;
    for (int k = N - 1; k >= 0; --k) {
      StageShared<M, T>& st = S.st[k];
      const bool dyn = k < N - 1;
      // phase 1: lane j owns column j of F = [Q | g0 | g1 | gl] + [A B]' PAB, every row of it
      for (int j = g.lane; j < NVA; j += G) {
        const bool rhs = j >= RV;
        // column j of [Q | g0 | g1 | gl] starts at Qs[qoff]; the loads are unconditional, a column of a variable with a
        // diagonal-only Hessian selects zeros instead of what it fetched
        const int qoff = rhs ? (j == RV ? Dims<M>::OFF0 : Dims<M>::OFF0 + (j - RV - 3) * RV) : (j < NDS ? j * NDS : 0);
        const bool take = rhs || j < NDS;
        T a[RV];
        RMPC_UNROLL
        for (int i = 0; i < RV; ++i) {
          const T q = st.Qs[qoff + i];
          if (i < NDS) a[i] = (NDS == RV || take) ? q : T(0);
          else { const T h = st.Qs[NQD + (i < NDS ? 0 : i - NDS)]; a[i] = rhs ? q : (j == i ? h : T(0)); }
        }
        if (dyn) {
          T abv[DS::NABV];
          RMPC_UNROLL
          for (int i = 0; i < DS::NVR * DS::NVC; ++i) abv[i] = st.ABv[i];
          T w[RX];
          RMPC_UNROLL
          for (int l = 0; l < RX; ++l) w[l] = S.PAB[l * LDF + j];
          RMPC_UNROLL
          for (int i = 0; i < RV; ++i)
            RMPC_UNROLL
            for (int e = 0; e < DS::cn(i); ++e) a[i] += DS::template cv<T>(i, e, abv) * w[DS::crow(i, e)];
        }
        RMPC_UNROLL
        for (int i = 0; i < RV; ++i) S.F[i * LDF + j] = a[i];
      }
      g.sync();
      // phase 2: Cholesky of Fuu (every lane, redundantly) and the RX+2 solves (one column per lane)
      {
        T L[NU * NU];
        RMPC_UNROLL
        for (int a = 0; a < NU; ++a)
          RMPC_UNROLL
          for (int b = 0; b <= a; ++b) L[a * NU + b] = S.F[(RX + a) * LDF + RX + b];
        RMPC_UNROLL
        for (int j = 0; j < NU; ++j) {
          T dd = L[j * NU + j];
          RMPC_UNROLL
          for (int c = 0; c < j; ++c) dd -= L[j * NU + c] * L[j * NU + c];
          if (!(dd > T(0))) { ok = false; dd = T(1); }
          const T il = trsqrt(dd);
          L[j * NU + j] = il;  // reciprocal pivot
          RMPC_UNROLL
          for (int i = j + 1; i < NU; ++i) {
            T v = L[i * NU + j];
            RMPC_UNROLL
            for (int c = 0; c < j; ++c) v -= L[i * NU + c] * L[j * NU + c];
            L[i * NU + j] = v * il;
          }
        }
        RMPC_UNROLL
        for (int j = 0; j < NU; ++j) res_stat = tmax(res_stat, tabs(S.F[(RX + j) * LDF + RV + 2]));
        for (int c = g.lane; c < RX + 2; c += G) {
          const int col = c < RX ? c : RV + (c - RX);
          T y[NU];
          RMPC_UNROLL
          for (int i = 0; i < NU; ++i) {
            T v = S.F[(RX + i) * LDF + col];
            RMPC_UNROLL
            for (int l = 0; l < i; ++l) v -= L[i * NU + l] * y[l];
            y[i] = v * L[i * NU + i];
          }
          RMPC_UNROLL
          for (int i = NU - 1; i >= 0; --i) {
            T v = y[i];
            RMPC_UNROLL
            for (int l = i + 1; l < NU; ++l) v -= L[l * NU + i] * y[l];
            y[i] = v * L[i * NU + i];
          }
          RMPC_UNROLL
          for (int i = 0; i < NU; ++i) {
            st.Qs[c < RX ? i * RX + c : NU * RX + (c - RX) * NU + i] = -y[i];  // K(i, c) | kff0(i) | kff1(i)
          }
        }
      }
      g.sync();
      // phase 3: lane i owns row i of P = Fxx + Fux' K, p = fx + Fux' kff and the adjoint nu (x-part of the Lagrangian
      // gradient), and turns them at once into row i of PAB = P [A B | d-term | p1 | nu] of the stage before
      for (int i = g.lane; i < RX; i += G) {
        T fu[NU];
        RMPC_UNROLL
        for (int l = 0; l < NU; ++l) fu[l] = S.F[(RX + l) * LDF + i];
        T Pi[RX];
        RMPC_UNROLL
        for (int j = 0; j < RX; ++j) {
          T v = S.F[i * LDF + j];
          RMPC_UNROLL
          for (int l = 0; l < NU; ++l) v += fu[l] * st.K(l, j);
          Pi[j] = v;
        }
        T p0 = S.F[i * LDF + RV], p1 = S.F[i * LDF + RV + 1];
        const T nui = S.F[i * LDF + RV + 2];
        RMPC_UNROLL
        for (int l = 0; l < NU; ++l) { p0 += fu[l] * st.kff0(l); p1 += fu[l] * st.kff1(l); }
        nsum += tabs(nui);
        if (k > 0) {
          StageShared<M, T>& sp = S.st[k - 1];
          if (i < 2) sp.nu01[i] = nui;
          T abv[DS::NABV];
          RMPC_UNROLL
          for (int j = 0; j < DS::NVR * DS::NVC; ++j) abv[j] = sp.ABv[j];
          RMPC_UNROLL
          for (int j = 0; j < RV; ++j) {
            T a = T(0);
            RMPC_UNROLL
            for (int e = 0; e < DS::cn(j); ++e) a += DS::template cv<T>(j, e, abv) * Pi[DS::crow(j, e)];
            S.PAB[i * LDF + j] = a;
          }
          T a = p0;
          RMPC_UNROLL
          for (int l = 0; l < RX; ++l) a += Pi[l] * sp.d[DS::full(l)];
          S.PAB[i * LDF + RV] = a;
          S.PAB[i * LDF + RV + 1] = p1;
          S.PAB[i * LDF + RV + 2] = nui;
        }
      }
      g.sync();
    }
    o.res_stat = res_stat;  // identical in every lane
    o.mult_sum += g.sum(nsum);
    o.ok = ok;
    o.finite = (o.f == o.f) && (o.res_stat == o.res_stat) && (o.theta == o.theta) && (o.mult_sum == o.mult_sum);
  }

  // ---- C: forward roll-out: dz_k = (dx_k, K_k dx_k + kff_k), dx_{k+1} = [A B] dz_k + d_k, into shared memory.
  // Lane j < NU evaluates row j of the feedback law, the rows are exchanged by shuffles, and every lane carries dx (the
  // coupled states) and di (the states the Newton system leaves out: di+ = di + d).
  RMPC_HD void rollout(const T mu) {
    constexpr int NI = NX - RX > 0 ? NX - RX : 1;
    T dx[RX], di[NI];
    RMPC_UNROLL
    for (int i = 0; i < RX; ++i) dx[i] = T(0);  // x_1 is fixed
    RMPC_UNROLL
    for (int i = 0; i < NI; ++i) di[i] = T(0);
    for (int k = 0; k < N; ++k) {
      StageShared<M, T>& st = S.st[k];
      T u[NU];
      if (G == 1) {
        RMPC_UNROLL
        for (int j = 0; j < NU; ++j) {
          T a = st.kff0(j) + mu * st.kff1(j);
          RMPC_UNROLL
          for (int i = 0; i < RX; ++i) a += st.K(j, i) * dx[i];
          u[j] = a;
        }
      } else {
        const int j = g.lane < NU ? g.lane : 0;
        T a = st.kff0(j) + mu * st.kff1(j), b = T(0);
        RMPC_UNROLL
        for (int i = 0; i + 1 < RX; i += 2) { a += st.K(j, i) * dx[i]; b += st.K(j, i + 1) * dx[i + 1]; }
        if (RX % 2) a += st.K(j, RX - 1) * dx[RX - 1];
        a += b;
        RMPC_UNROLL
        for (int l = 0; l < NU; ++l) u[l] = g.bcast(a, l);
      }
      RMPC_UNROLL
      for (int i = 0; i < RX; ++i) st.dz[DS::full(i)] = dx[i];      // same value from every lane
      RMPC_UNROLL
      for (int j = 0; j < NU; ++j) st.dz[NX + j] = u[j];
      if (NX > RX) {
        int c = 0;
        RMPC_UNROLL
        for (int f = 0; f < NX; ++f)
          if (DS::comp(f) < 0) { st.dz[f] = di[c < NI ? c : 0]; if (k < N - 1) di[c < NI ? c : 0] += st.d[f]; ++c; }
      }
      if (k < N - 1) {
        T dn[RX];
        RMPC_UNROLL
        for (int i = 0; i < RX; ++i) dn[i] = st.d[DS::full(i)];
        RMPC_UNROLL
        for (int j = 0; j < RV; ++j)
          RMPC_UNROLL
          for (int e = 0; e < DS::cn(j); ++e) dn[DS::crow(j, e)] += DS::template cv<T>(j, e, st.ABv) * (j < RX ? dx[j < RX ? j : 0] : u[j < RX ? 0 : j - RX]);
        RMPC_UNROLL
        for (int i = 0; i < RX; ++i) dx[i] = dn[i];
      }
    }
    g.sync();
  }

  static constexpr RMPC_HD bool row_is_geo(int r) {
    for (int i = 0; i < M::NGEO; ++i)
      if (M::geo_row(i) == r) return true;
    return false;
  }

  RMPC_HD void row_dirs(int sl, int k, T* gd) const {
    const T* sdz = S.st[k].dz;
    RMPC_UNROLL
    for (int r = 0; r < M::NGEO; ++r) {
      T a = T(0);
      RMPC_UNROLL
      for (int i = 0; i < NQ; ++i) a += Gg[sl][r][i] * sdz[i];
      gd[M::geo_row(r)] = a;
    }
    RMPC_UNROLL
    for (int r = 0; r < M::NLIM; ++r) gd[M::lim_row(r)] = T(M::lim_sign(r)) * sdz[M::lim_var(r)];
  }

  struct Loop {  // what one iteration hands to the next
    T mu, rho_pen = T(1), err_ref = T(1e300);
    T lm = T(0);  // adaptive Levenberg-Marquardt damping: grows when the line search has to cut the step, decays on full steps
    int it, ncreep = 0, it_ref = 0, gn_hold = 0, gn_len = 1;
    bool skipped = false;   // refused before the first iteration (bad or non-finite initial point)
    Red<T> so;
  };

  // ---- the solve ----------------------------------------------------------------------------------
  // The iteration exists twice: iterate<true> knows about rows that start infeasible (shifted rows), iterate<false>
  // does not and is smaller and cheaper in registers.  MODE picks what a build of solve() contains:
  //   kAll      every problem; decides by itself which copy a problem needs (the CPU twin, and the single-launch
  //             kernel for small batches)
  //   kFeasible problems that starts_shifted() found feasible: no shift code at all
  //   kShifted  the others: iterate<true> until the last shift is gone, then iterate<false>
  // The CUDA host code classifies a batch ONCE (classify_kernel -> two index lists) and runs the last two as two
  // kernels side by side, each on its own list, so that the common case runs the smaller code.
  static constexpr int kAll = 0, kFeasible = 1, kShifted = 2, kResume = 3;   // kResume: solve_resume (mixed precision, below)

  RMPC_HD void load_guess(const TS* xinit, const TS* z_in) {
    for (int e = g.lane; e < N * NV; e += G) S.st[e / NV].z[e % NV] = T(z_in[e]);
    g.sync();
    for (int e = g.lane; e < NX; e += G) S.st[0].z[e] = T(xinit[e]);
    for (int e = g.lane; e < N; e += G) { S.st[e].nu01[0] = T(0); S.st[e].nu01[1] = T(0); }
    g.sync();
  }

  // Does a row of the loaded initial guess start below delta?  (the one place a problem is classified on the GPU path)
  RMPC_HD bool starts_shifted() {
    bool shifted = false;
    RMPC_UNROLL_SL
    for (int sl = 0; sl < SLOTS; ++sl) {
      const int k = g.lane + G * sl;
      if (k < N) {
        T p[M::NPAR];
        load_params<M, T>(pv, k, p);
        StageVal<M, T> sv;
        stage_values<M, T>(S.st[k].z, p, sv);
        RMPC_UNROLL
        for (int r = 0; r < NH; ++r) shifted = shifted || !(sv.h[r] >= T(opt.delta));
      }
    }
    return g.any(shifted);
  }

  template <int MODE = 0>
  RMPC_HD void solve(const TS* xinit, const TS* z_in, TS* z_out, Info& info) {
    // initial point: z from the guess, x_1 = xinit
    load_guess(xinit, z_in);
    T mu = T(opt.mu0);
    bool pole0 = false, shifted = false, finite0 = true;
    if constexpr (MODE != kFeasible) {  // lanes without a stage take part in the "any shift left?" votes: they must hold zeros, not what the last problem left
      RMPC_UNROLL_SL
      for (int sl = 0; sl < SLOTS; ++sl)
        RMPC_UNROLL
        for (int r = 0; r < NH; ++r) sh[sl][r] = 0.0f;
    }
    RMPC_UNROLL_SL
    for (int sl = 0; sl < SLOTS; ++sl) {
      const int k = g.lane + G * sl;
      if (k < N) {
        T p[M::NPAR];
        load_params<M, T>(pv, k, p);
        StageVal<M, T> sv;
        stage_values<M, T>(S.st[k].z, p, sv);
        pole0 = pole0 || sv.pole;
        finite0 = finite0 && tabs(sv.f) < T(1e30);   // false for NaN and Inf: a state, guess or parameter that is not a number
        RMPC_UNROLL
        for (int r = 0; r < NH; ++r) {
          const T hv = sv.h[r];
          finite0 = finite0 && tabs(hv) < T(1e30);
          // a violated row starts with a slack as large as its violation: the fraction-to-boundary rule then leaves room
          // for steps that remove only part of the shift.  kFeasible: the classification found every row >= delta, and
          // the row is taken as it is (no shift storage in that build).
          float shift = 0.0f;
          if constexpr (MODE != kFeasible) {
            shift = hv >= T(opt.delta) ? 0.0f : float(tmax(T(opt.delta), -hv) - hv);
            sh[sl][r] = shift;
            shifted = shifted || shift > 0.0f;
          }
          SLK(sl, k, r) = hv + T(shift);
          LAM(sl, k, r) = mu * trcp(hv + T(shift));
        }
      }
    }
    info.exitflag = 0;
    Loop L;
    L.mu = mu; L.it = 0;
    L.so.f = L.so.res_stat = L.so.res_eq = L.so.comp_max = L.so.shift_max = T(0);
    if (g.any(pole0)) { info.exitflag = -8; L.it = opt.max_iter + 1; L.skipped = true; }
    if (g.any(!finite0)) { info.exitflag = -6; L.it = opt.max_iter + 1; L.skipped = true; }   // NaN / Inf in the inputs: no iteration is run
    // Rows that start infeasible are rare (a warm start whose obstacles moved onto the plan), and their bookkeeping
    // costs registers and instructions in every unrolled row loop: the iteration exists twice, and the copy that
    // knows about shifts runs only until the last shift of the problem has reached zero.
    bool done = false;
    if constexpr (MODE != kFeasible) {
      if (MODE == kShifted || g.any(shifted)) done = iterate<true>(L, info);
    }
    if (!done) iterate<false>(L, info);
    finish(L, z_out, info);
  }


  // ---- mixed precision (rmpc_options.precision = 1): the iteration runs in float32 while the barrier parameter is large
  // and hands its state (z, lambda, mu, iteration count) to the float64 build of the same iteration, which finishes the
  // solve to the caller's tolerances.  Late iterations need float64: active rows reach lambda/s ~ 1e9 and slacks fall below
  // the rounding of a float32 row value; early iterations do not.  The reference's generated code is all-double
  // (models/mpcModel.py:110-126 sets no precision option), so precision 0 stays the parity mode.
  //
  // solve_lowp: the float32 part (T = float).  hit = iterations done, or -1: start over in float64 (the float32 iteration
  // broke down, or the guess sits on a pole).  The caller passes options with the tolerances float32 can reach (lowp_options).
  static RMPC_HD Options lowp_options(Options o) {
    o.tol_stat = 1e-2; o.tol_eq = 1e-4; o.tol_ineq = 1e-4; o.tol_comp = 1e-3; o.mu_min = 1e-4;
    return o;
  }

  RMPC_HD void solve_lowp(const TS* xinit, const TS* z_in, float* hz, float* hlam, float* hmu, int* hit) {
    load_guess(xinit, z_in);
    T mu = T(opt.mu0);
    bool pole0 = false;
    RMPC_UNROLL_SL
    for (int sl = 0; sl < SLOTS; ++sl) {
      const int k = g.lane + G * sl;
      if (k < N) {
        T p[M::NPAR];
        load_params<M, T>(pv, k, p);
        StageVal<M, T> sv;
        stage_values<M, T>(S.st[k].z, p, sv);
        pole0 = pole0 || sv.pole;
        RMPC_UNROLL
        for (int r = 0; r < NH; ++r) {
          SLK(sl, k, r) = sv.h[r];
          LAM(sl, k, r) = mu * trcp(sv.h[r]);
          pole0 = pole0 || !(sv.h[r] > T(0));   // float32 rounding of a row the classification (float64) found >= delta: leave it to float64
        }
      }
    }
    Info info;
    info.exitflag = 0;
    Loop L;
    L.mu = mu; L.it = 0;
    L.so.f = L.so.res_stat = L.so.res_eq = L.so.comp_max = L.so.shift_max = T(0);
    if (g.any(pole0)) info.exitflag = -8;
    else iterate<false>(L, info);
    const bool go_on = info.exitflag >= 0;
    if (go_on) {
      for (int e = g.lane; e < N * NV; e += G) hz[e] = float(S.st[e / NV].z[e % NV]);
      RMPC_UNROLL_SL
      for (int sl = 0; sl < SLOTS; ++sl) {
        const int k = g.lane + G * sl;
        if (k < N) {
          RMPC_UNROLL
          for (int r = 0; r < NH; ++r) hlam[k * NH + r] = float(LAM(sl, k, r));
        }
      }
    }
    if (g.lane == 0) { *hmu = float(L.mu); *hit = go_on ? (L.it > opt.max_iter ? opt.max_iter : L.it) : -1; }
  }

  // solve_resume: the float64 part (T = double, no shift code).  At most two passes through the one copy of the float64
  // iteration: from the handed-over state, and — if there is none, or it led nowhere — from the caller's guess, as
  // precision 0 would.
  RMPC_HD void solve_resume(const TS* xinit, const TS* z_in, TS* z_out, const float* hz, const float* hlam, float hmu, int hit, Info& info) {
    Loop L;
    for (int pass = hit >= 0 ? 0 : 1; pass < 2; ++pass) {
      const bool resume = pass == 0;
      if (resume) {
        for (int e = g.lane; e < N * NV; e += G) S.st[e / NV].z[e % NV] = T(hz[e]);
        g.sync();
        for (int e = g.lane; e < NX; e += G) S.st[0].z[e] = T(xinit[e]);   // x_1 = xinit exactly
        for (int e = g.lane; e < N; e += G) { S.st[e].nu01[0] = T(0); S.st[e].nu01[1] = T(0); }
        g.sync();
      } else {
        load_guess(xinit, z_in);
      }
      const T mu = resume ? T(hmu) : T(opt.mu0);
      bool pole0 = false, finite0 = true;
      RMPC_UNROLL_SL
      for (int sl = 0; sl < SLOTS; ++sl) {
        const int k = g.lane + G * sl;
        if (k < N) {
          T p[M::NPAR];
          load_params<M, T>(pv, k, p);
          StageVal<M, T> sv;
          stage_values<M, T>(S.st[k].z, p, sv);
          pole0 = pole0 || sv.pole;
          finite0 = finite0 && tabs(sv.f) < T(1e30);
          RMPC_UNROLL
          for (int r = 0; r < NH; ++r) {
            finite0 = finite0 && tabs(sv.h[r]) < T(1e30);
            if (resume) {
              const T lr = tmax(T(hlam[k * NH + r]), T(1e-30));
              // the slack is the row's own value; where float32 rounding of z has eaten a tiny slack, a tenth of the centred
              // value mu / lambda stands in until the first accepted step re-evaluates the row
              SLK(sl, k, r) = tmax(sv.h[r], T(0.1) * mu * trcp(lr));
              LAM(sl, k, r) = lr;
            } else {
              SLK(sl, k, r) = sv.h[r];
              LAM(sl, k, r) = mu * trcp(sv.h[r]);
            }
          }
        }
      }
      info.exitflag = 0;
      L = Loop();
      L.mu = mu; L.it = resume ? hit : 0;
      L.so.f = L.so.res_stat = L.so.res_eq = L.so.comp_max = L.so.shift_max = T(0);
      if (g.any(pole0)) { info.exitflag = -8; L.it = opt.max_iter + 1; L.skipped = !resume; }
      if (g.any(!finite0)) { info.exitflag = -6; L.it = opt.max_iter + 1; L.skipped = !resume; }
      iterate<false>(L, info);
      if (info.exitflag >= 0) break;
    }
    finish(L, z_out, info);
  }

  // Runs iterations until the solve ends (returns true) or, for SH, until no row of the problem is shifted any more.
  template <bool SH>
  RMPC_HD bool iterate(Loop& L, Info& info) {
    T mu = L.mu, rho_pen = L.rho_pen, err_ref = L.err_ref, lm = L.lm;
    int it = L.it, ncreep = L.ncreep, it_ref = L.it_ref, gn_hold = L.gn_hold, gn_len = L.gn_len;
    Red<T> so = L.so;
    bool handover = false;
    bool soc_on = false;   // this pass through the loop body repeats the iteration with the second-order correction in the right-hand side
    bool soc_block = false;   // the correction was tried for this iteration and dropped
    for (; it <= opt.max_iter; ++it) {
      if constexpr (SH) {
        bool left = false;
        RMPC_UNROLL_SL
        for (int sl = 0; sl < SLOTS; ++sl)
          RMPC_UNROLL
          for (int r = 0; r < NH; ++r) left = left || sh[sl][r] > 0.0f;
        if (!g.any(left)) { handover = true; break; }
      }
      g.align();
      // An indefinite exact Hessian costs a whole derivative pass + sweep before the Gauss-Newton retry, and far from
      // the solution it is indefinite iteration after iteration: after a failure the exact attempt is skipped for
      // 1, 2, 4, 4, ... iterations (gn_hold), until it succeeds again.
      bool exact = opt.hessian == 1 && gn_hold == 0;
      if (gn_hold > 0) --gn_hold;
      T reg = T(opt.reg) + lm;
      for (int tries = 0; tries < 5; ++tries) {
        derivs<SH>(exact, reg, so, SOCC && soc_on);
        sweep(so);
        if (so.ok) { if (exact) gn_len = 1; break; }
        if (exact) { exact = false; gn_hold = gn_len; gn_len = gn_len < int(opt.gn_hold_max) ? 2 * gn_len : gn_len; }  // not positive definite: Gauss-Newton
        else reg = tmax(reg * T(100), T(1e-6));         // rounding broke it (huge barrier weights): Levenberg
      }
      if (!so.finite || !so.ok) { info.exitflag = so.ok ? -6 : -5; break; }
      if (so.comp_max > T(1e15) || so.s_min < T(1e-13)) { info.exitflag = -7; break; }  // multipliers diverge: locally infeasible
      // ---- convergence test (IPOPT scaling of the dual residual)
      const T sd = tmax(T(100), so.mult_sum * T(1.0 / (N * (NH + NX)))) * T(0.01);
      const T isd = trcp(sd);
      if (so.res_stat * isd <= T(opt.tol_stat) && so.res_eq <= T(opt.tol_eq) && so.shift_max <= T(opt.tol_ineq) &&
          so.comp_max * isd <= T(opt.tol_comp)) {
        info.exitflag = 1;
        break;
      }
      if (it == opt.max_iter) break;
      {  // stall detection on the optimality error of the original problem
        const T err = tmax(tmax(so.res_stat, so.comp_max) * isd, tmax(so.res_eq, so.shift_max));
        if (err < T(0.1) * err_ref) { err_ref = err; it_ref = it; }
        if (opt.stall_window > 0 && it - it_ref >= int(opt.stall_window)) { info.exitflag = -7; break; }
      }
      // ---- monotone barrier update (Fiacco-McCormick as in IPOPT): shrink mu while the current barrier
      //      problem is solved to within kappa_eps * mu
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
      for (int rep = 0; rep < 8; ++rep) {
        const T cmu = tmax(so.comp_max - mu, mu - so.comp_min);
        const T emu = tmax(tmax(so.res_stat, cmu) * isd, tmax(so.res_eq, so.shift_max));
        if (emu <= T(10) * mu && mu > T(opt.mu_min)) {
          const T m15 = mu * mu * trsqrt(mu);  // mu^1.5
          mu = tmax(T(opt.mu_min), tmin(T(0.2) * mu, m15));
        } else break;
      }
      // ---- Newton step, linear step-length estimates, merit slope
      rollout(mu);
      const T tau = tmax(T(opt.tau_min), T(1) - mu);
      const T mu_est = mu;
      float rp = 0.0f, rpl = 0.0f, rd = 0.0f;  // step-length caps: single precision is enough (the true rows are checked in the line search)
      T dphi = T(0);
      RMPC_UNROLL_SL
      for (int sl = 0; sl < SLOTS; ++sl) {
        const int k = g.lane + G * sl;
        if (k < N) {
          T gd[NH];
          row_dirs(sl, k, gd);
          RMPC_UNROLL
          for (int c = 0; c < RV; ++c) dphi += S.st[k].g0(c) * S.st[k].dz[DS::full(c)];
          T csoc[NH];
          RMPC_UNROLL
          for (int r = 0; r < NH; ++r) csoc[r] = T(0);
          if (!SH && SOCC && soc_on) {
            RMPC_UNROLL
            for (int r = 0; r < M::NGEO; ++r) csoc[M::geo_row(r)] = T(soc[sl][r]);
          }
          RMPC_UNROLL
          for (int r = 0; r < NH; ++r) {
            const T sr = SLK(sl, k, r), lr = LAM(sl, k, r);
            const T shr = SH ? T(sh[sl][r]) : -csoc[r];  // a second-order correction acts like a shift of -c
            const T dsr = gd[r] - shr;                   // the step is asked to remove the whole shift
            const T rs = trcp(sr);
            if (SH || (SOCC && soc_on)) dphi += (lr * rs) * shr * gd[r];   // g0 holds the cost gradient minus the shift push: undo it here
            const T dlr = mu * rs - lr - (lr * rs) * dsr;
            rp = fmaxf(rp, float(-dsr * rs));             // largest relative decrease of a slack ...
            if (!row_is_geo(r)) rpl = fmaxf(rpl, float(-dsr * rs));   // ... and of the slack of a limit row (linear: its estimate is exact)
            S.st[k].Qs[NH + r] = dlr;                     // kept for the multiplier update (K, kff and g0 are dead from here on)
            rd = fmaxf(rd, fdiv_approx(-float(dlr), float(lr)));      // ... and of a multiplier, for the fraction-to-boundary rule
            dphi -= mu * (SH ? dsr : gd[r]) * rs;
          }
        }
      }
      rp = g.max(rp); rd = g.max(rd); dphi = g.sum(dphi);
      const T ap = T(rp) > tau ? tau / T(rp) : T(1), ad = T(rd) > tau ? tau / T(rd) : T(1);
      // The slack of a row is the row's own value here, s = h(z) + shift, so only the true rows have to stay inside.  The
      // geometric rows are convex in the point they constrain (distances), their linearisation underestimates them, and
      // a cap taken from it makes the iterate crawl along a curved obstacle: the first trial is capped by the limit rows
      // only, and falls back to the linearised cap when a true row objects.
      T ap0 = ap;
      if (opt.ls_optimistic > 0) { rpl = g.max(rpl); ap0 = T(rpl) > tau ? tau / T(rpl) : T(1); }
      bool soc_reject = false;
      if (so.theta > T(1e-14)) {
        const T need = dphi * trcp(T(0.9) * so.theta);
        if (need > rho_pen) {
          if (SOCC && soc_on) soc_reject = true;   // a corrected step must not buy its descent with a larger penalty
          else rho_pen = need * T(1.1) + T(1e-3);
        }
      }
      const T phi0 = so.f - mu * so.logs + rho_pen * so.theta;
      const T slope = dphi - rho_pen * so.theta;
      if (SOCC && soc_on && (soc_reject || !(slope < T(0)))) { soc_on = false; soc_block = true; --it; g.sync(); continue; }   // back to the plain step
      // ---- backtracking: stay inside h + shift >= (1 - tau) s on the true rows, Armijo on the l1 merit
      T alpha = ap0;
      bool accepted = false;
      int nls = 0;
      bool cut = false;  // the first trial within the linearised cap left the interior of a nonlinear row
      bool lin_tried = false;
      // candidate for a second-order correction: the iteration before was cut as well, this pass is not a corrected one yet, and
      // the trial step is long enough for its curvature defect, scaled to the full step by 1/alpha^2, to mean something
      const bool soc_cand = !SH && SOCC && opt.soc > 0 && !soc_on && !soc_block && ncreep >= 1 && alpha >= T(0.25);
      bool redo = false;
      for (int ls = 0; ls < 30; ++ls) {
        T f = T(0), theta = T(0);
        LogAcc<T> la;
        bool bad = false;
        RMPC_UNROLL_SL
        for (int sl = 0; sl < SLOTS; ++sl) {
          const int k = g.lane + G * sl;
          if (k < N) {
            T p[M::NPAR];
            load_params<M, T>(pv, k, p);
            T zt[NV];
            RMPC_UNROLL
            for (int i = 0; i < NV; ++i) zt[i] = S.st[k].z[i] + alpha * S.st[k].dz[i];
            StageVal<M, T> sv;
            stage_values<M, T>(zt, p, sv);
            f += sv.f;
            bad = bad || sv.pole;
            RMPC_UNROLL
            for (int r = 0; r < NH; ++r) {
              T stv = sv.h[r];
              if constexpr (SH) {  // same rounding as the update of sh below
                const float shn = float((T(1) - alpha) * T(sh[sl][r]));
                if (shn >= 1e-12f) { stv += T(shn); theta += T(shn); }
              }
              S.st[k].Qs[r] = sv.h[r];  // kept for the update of s once the trial is accepted (K and kff are dead by now)
              if (!(stv >= (T(1) - tau) * SLK(sl, k, r))) {
                bad = true;
#if defined(RMPC_TRACE) && !defined(__CUDA_ARCH__)
                if (RMPC_TRACE > 2) { T gd[NH]; row_dirs(sl, k, gd); printf("      bad row k %d r %d s %.3g lam %.3g st %.3g lin %.3g alpha %.3g\n", k, r, double(SLK(sl, k, r)), double(LAM(sl, k, r)), double(stv), double(SLK(sl, k, r) + alpha * gd[r]), double(alpha)); }
#endif
              }
              la.mul(tmax(stv, T(0)));
              if ((r % 12 == 11 && r + 1 < NH) || (SLOTS > 1 && r == NH - 1)) la.flush();
            }
            if (soc_cand && ls == 0) {   // curvature defect of the geometric rows along this trial step
              const T ia2 = trcp(alpha * alpha);
              RMPC_UNROLL
              for (int r = 0; r < M::NGEO; ++r) {
                T lin = SLK(sl, k, M::geo_row(r));
                RMPC_UNROLL
                for (int i = 0; i < NQ; ++i) lin += alpha * Gg[sl][r][i] * S.st[k].dz[i];
                soc[sl][r] = float((sv.h[M::geo_row(r)] - lin) * ia2);
              }
            }
            if (k < N - 1) {
              T xn[NX];
              dyn_value<M, T>(zt, xn);
              RMPC_UNROLL
              for (int i = 0; i < NX; ++i) theta += tabs(xn[i] - (S.st[k + 1].z[i] + alpha * S.st[k + 1].dz[i]));
            }
          }
        }
        f = g.sum(f); theta = g.sum(theta);
        const T logs = g.sum(la.value());
        bad = g.any(bad);
        const T phi = f - mu * logs + rho_pen * theta;
        const T tol_phi = T(10) * T(sizeof(T) == 8 ? 2.2e-16 : 1.2e-7) * tabs(phi0);
#if defined(RMPC_TRACE) && !defined(__CUDA_ARCH__)
        if (RMPC_TRACE > 1) printf("    ls %d alpha %.3g f %.9g (f0 %.9g) theta %.4g (%.4g) barrier %.6g (%.6g) bad %d phi-phi0 %.4g\n", ls, double(alpha), double(f), double(so.f), double(theta), double(so.theta), double(-mu * logs), double(-mu * so.logs), int(bad), double(phi - phi0));
#endif
        if (!bad && (phi == phi) && phi <= phi0 + T(1e-4) * alpha * slope + tol_phi) { accepted = true; break; }
        if (soc_cand && ls == 0 && bad) { redo = true; break; }   // the same iteration again, with the correction in the right-hand side
        if (alpha > ap) {   // an optimistic trial beyond the linearised cap: not counted as a backtrack
          alpha = bad ? tmin(ap, T(0.5) * alpha) : T(0.5) * alpha;
          continue;
        }
        alpha *= T(0.5);
        if (!lin_tried && bad) cut = true;
        lin_tried = true;
        ++nls;
      }
#if defined(RMPC_TRACE) && !defined(__CUDA_ARCH__)
      printf("it %d mu %.3g f %.9g theta %.3g stat %.3g comp %.3g..%.3g shift %.3g | ap %.3g ad %.3g alpha %.3g dphi %.3g rho %.3g slope %.3g acc %d ex %d\n",
             it, double(mu), double(so.f), double(so.theta), double(so.res_stat), double(so.comp_min), double(so.comp_max),
             double(so.shift_max), double(ap), double(ad), double(alpha), double(dphi), double(rho_pen), double(slope), int(accepted), int(exact));
#endif
      if (SOCC && redo) { soc_on = true; --it; g.sync(); continue; }
      if (!accepted) { info.exitflag = -7; break; }
      ncreep = cut ? ncreep + 1 : 0;
      soc_on = false; soc_block = false;
      if (opt.creep_trigger > 0 && ncreep >= int(opt.creep_trigger)) { mu = tmin(mu * T(opt.creep_factor), T(opt.mu0)); ncreep = 0; }
      if (nls >= int(opt.lm_trigger)) lm = tmax(lm * T(opt.lm_up), T(opt.lm0));
      else if (nls == 0) lm = lm * T(opt.lm_down) < T(0.1) * T(opt.lm0) ? T(0) : lm * T(opt.lm_down);
      // ---- take the step
      g.sync();  // every lane has read its neighbour's z for this line search
      const T adu = tmin(ad, T(1));
      const T ikappa = T(1.0 / opt.kappa);
      RMPC_UNROLL_SL
      for (int sl = 0; sl < SLOTS; ++sl) {
        const int k = g.lane + G * sl;
        if (k < N) {
          if (mu != mu_est) {  // the creep guard raised mu after the multiplier steps were formed: d(lambda) is linear in mu
            RMPC_UNROLL
            for (int r = 0; r < NH; ++r) S.st[k].Qs[NH + r] += (mu - mu_est) * trcp(SLK(sl, k, r));
          }
          T snv[NH];  // new slacks: the row values of the accepted trial ...
          RMPC_UNROLL
          for (int r = 0; r < NH; ++r) snv[r] = S.st[k].Qs[r];
          if constexpr (SH) {  // ... plus, for rows that started infeasible, what is left of their shift after a step of length alpha
            RMPC_UNROLL
            for (int r = 0; r < NH; ++r) {
              float shn = float((T(1) - alpha) * T(sh[sl][r]));
              if (shn < 1e-12f) shn = 0.0f;
              snv[r] += T(shn);
              sh[sl][r] = shn;
            }
          }
          RMPC_UNROLL
          for (int r = 0; r < NH; ++r) {
            const T lr = LAM(sl, k, r);
            const T dlr = S.st[k].Qs[NH + r];
            const T sn = snv[r];
            T ln = lr + adu * dlr;
            const T mus = mu * trcp(sn);
            ln = tmin(tmax(ln, mus * ikappa), T(1e10) * mus);
            SLK(sl, k, r) = sn;
            LAM(sl, k, r) = ln;
          }
          RMPC_UNROLL
          for (int i = 0; i < NV; ++i) S.st[k].z[i] += alpha * S.st[k].dz[i];
        }
      }
      g.sync();
    }
    L.mu = mu; L.rho_pen = rho_pen; L.err_ref = err_ref; L.lm = lm;
    L.it = it; L.ncreep = ncreep; L.it_ref = it_ref; L.gn_hold = gn_hold; L.gn_len = gn_len;
    L.so = so;
    return !handover;
  }

  RMPC_HD void finish(const Loop& L, TS* z_out, Info& info) {
    if (z_out)
      for (int e = g.lane; e < N * NV; e += G) z_out[e] = TS(S.st[e / NV].z[e % NV]);
    // The NLP of the reference carries -100 <= z <= 100 on every stage variable (models/mpcModel.py:23-28,98-104).  The
    // box is not part of the barrier problem solved here (it is never active in a sensible scenario, and 2 nvar more rows
    // per stage would cost every solve); a converged point that touches or leaves it is not a solution of the
    // reference's NLP and is reported as such (exit flag -9) instead of as optimal.
    if (info.exitflag == 1) {
      float zmax = 0.0f;
      for (int e = g.lane; e < N * NV; e += G) zmax = fmaxf(zmax, float(tabs(S.st[e / NV].z[e % NV])));
      if (g.max(zmax) > 100.0f - 1e-4f) info.exitflag = -9;
    }
    info.iters = L.skipped ? 0 : (L.it > opt.max_iter ? opt.max_iter : L.it);
    info.obj = double(L.so.f);
    info.res_stat = double(L.so.res_stat);
    info.res_eq = double(L.so.res_eq);
    info.res_ineq = double(L.so.shift_max);
    info.res_comp = double(L.so.comp_max);
    info.mu = double(L.mu);
  }
};

}  // namespace rmpc
