// rmpc_ipm.cuh — one nonlinear-MPC solve: primal-dual interior point on the stage-structured NLP
//
//   min  sum_k J(z_k,p_k)   s.t.  x_1 = xinit,  x_{k+1} = Phi(x_k,u_k),  h(z_k,p_k) >= 0
//
// with slacks (h - s = 0, s >= 0), the inequality multipliers condensed into the stage Hessians and the
// resulting equality-constrained QP solved by a backward Riccati recursion / forward roll-out.  This is
// the component that replaces the FORCESPRO-generated solver the reference calls at
// planner/mpcPlanner.py:262.  One *lane* runs one problem (a warp = 32 problems);
// all per-problem state lives in a `Work` struct that the compiler places in registers / lane-interleaved
// local memory, so that every access of a warp is coalesced.
//
// Precision: `T` is the iterate/residual type, `TF` the type of the Riccati factorisation (the O(n^3)
// part).  (T,TF) = (double,double) is the reference-grade path, (double,float) the mixed path.
#pragma once
#include "rmpc_model.cuh"
#if defined(RMPC_TRACE)
#include <stdio.h>
#include <stdlib.h>
#endif

namespace rmpc {

struct Options {
  int max_iter;        // iteration cap (FORCESPRO default 200)
  int hessian;         // 0 Gauss-Newton, 1 exact Lagrangian Hessian with GN fallback when not positive definite
  double tol_stat;     // scaled stationarity
  double tol_eq;       // dynamics defect
  double tol_ineq;     // |h - s|
  double tol_comp;     // complementarity s*lambda
  double mu0;          // initial barrier parameter
  double mu_min;
  double reg;          // Levenberg term added to the diagonal of every stage Hessian
  int max_inner;       // force a barrier decrease after this many iterations at one mu (0: never)
};

RMPC_HD Options default_options() {
  Options o;
  o.max_iter = 200; o.hessian = 1;
  o.tol_stat = 1e-6; o.tol_eq = 1e-8; o.tol_ineq = 1e-8; o.tol_comp = 1e-8;
  o.mu0 = 1.0; o.mu_min = 1e-9; o.reg = 0.0; o.max_inner = 0;
  return o;
}

struct Info {
  int exitflag;   // 1 optimal, 0 iteration limit, -6 NaN/Inf, -7 line search failed (FORCESPRO sign convention)
  int iters;
  double obj, res_stat, res_eq, res_ineq, res_comp, mu;
};

template <class M, class T>
struct Work {
  T z[M::N][M::NVAR];
  T s[M::N][M::NH];
  T lam[M::N][M::NH];
  T dz[M::N][M::NVAR];
  T K[M::N][M::NU * M::NX];
  T kff0[M::N][M::NU];                               // feed-forward, mu-independent part
  T kff1[M::N][M::NU];                               // feed-forward, coefficient of mu
  T gf[M::N][M::NVAR];                               // cost gradient
  T hh[M::N][M::NH];                                 // h(z)
  T Gg[M::N][M::NGEO > 0 ? M::NGEO : 1][M::NQ];      // geometric row gradients
};

// Parameter access: p(k, j) of the problem this lane owns.
template <class T>
struct ParamView {
  const T* base;
  long stage_stride;  // elements between consecutive stages (0: stage-invariant)
  long elem_stride;   // elements between consecutive parameters
  RMPC_HD T get(int k, int j) const { return base[k * stage_stride + j * elem_stride]; }
};

template <class M, class T>
RMPC_HD void load_params(const ParamView<T>& pv, int k, T* p) {
  RMPC_UNROLL
  for (int j = 0; j < M::NPAR; ++j) p[j] = pv.get(k, j);
}

// directional derivative of row r of stage k along dz_k
template <class M, class T>
RMPC_HD void row_dirs(const Work<M, T>& w, int k, T* gd) {
  RMPC_UNROLL
  for (int r = 0; r < M::NGEO; ++r) {
    T a = T(0);
    RMPC_UNROLL
    for (int i = 0; i < M::NQ; ++i) a += w.Gg[k][r][i] * w.dz[k][i];
    gd[M::geo_row(r)] = a;
  }
  RMPC_UNROLL
  for (int r = 0; r < M::NLIM; ++r) gd[M::lim_row(r)] = T(M::lim_sign(r)) * w.dz[k][M::lim_var(r)];
}

// G' y for a stage (adds -G'y into out)
template <class M, class T>
RMPC_HD void sub_Gt(const Work<M, T>& w, int k, const T* y, T* out) {
  RMPC_UNROLL
  for (int r = 0; r < M::NGEO; ++r) {
    const T yr = y[M::geo_row(r)];
    RMPC_UNROLL
    for (int i = 0; i < M::NQ; ++i) out[i] -= w.Gg[k][r][i] * yr;
  }
  RMPC_UNROLL
  for (int r = 0; r < M::NLIM; ++r) out[M::lim_var(r)] -= T(M::lim_sign(r)) * y[M::lim_row(r)];
}

// Cholesky of the NU x NU block, then K = -Fuu^{-1} Fux and the two feed-forward vectors
// kff0 = -Fuu^{-1} fu0 (mu-independent part of the gradient), kff1 = -Fuu^{-1} fu1 (coefficient of mu).
// Returns false if a pivot is not positive (Hessian not positive definite on the null space of the dynamics).
template <class M, class TF>
RMPC_HD bool factor_solve(TF* Fuu, const TF* Fux, const TF* fu0, const TF* fu1, TF* K, TF* kff0, TF* kff1) {
  constexpr int NU = M::NU, NX = M::NX;
  RMPC_UNROLL
  for (int j = 0; j < NU; ++j) {
    TF d = Fuu[j * NU + j];
    RMPC_UNROLL
    for (int k = 0; k < j; ++k) d -= Fuu[j * NU + k] * Fuu[j * NU + k];
    if (!(d > TF(0))) return false;
    const TF l = tsqrt(d), il = TF(1) / l;
    Fuu[j * NU + j] = il;  // store the reciprocal pivot
    RMPC_UNROLL
    for (int i = j + 1; i < NU; ++i) {
      TF v = Fuu[i * NU + j];
      RMPC_UNROLL
      for (int k = 0; k < j; ++k) v -= Fuu[i * NU + k] * Fuu[j * NU + k];
      Fuu[i * NU + j] = v * il;
    }
  }
  RMPC_UNROLL
  for (int c = 0; c < NX + 2; ++c) {
    TF y[NU];
    RMPC_UNROLL
    for (int i = 0; i < NU; ++i) {
      TF v = (c < NX) ? Fux[i * NX + (c < NX ? c : 0)] : (c == NX ? fu0[i] : fu1[i]);
      RMPC_UNROLL
      for (int k = 0; k < i; ++k) v -= Fuu[i * NU + k] * y[k];
      y[i] = v * Fuu[i * NU + i];
    }
    RMPC_UNROLL
    for (int i = NU - 1; i >= 0; --i) {
      TF v = y[i];
      RMPC_UNROLL
      for (int k = i + 1; k < NU; ++k) v -= Fuu[k * NU + i] * y[k];
      y[i] = v * Fuu[i * NU + i];
    }
    RMPC_UNROLL
    for (int i = 0; i < NU; ++i) {
      if (c < NX) K[i * NX + (c < NX ? c : 0)] = -y[i];
      else if (c == NX) kff0[i] = -y[i];
      else kff1[i] = -y[i];
    }
  }
  return true;
}

template <class M, class T>
struct SweepOut {
  T res_stat, res_eq, res_ineq;
  T comp_min, comp_max; // extrema of s*lambda
  T f, theta, logs;     // objective, l1 infeasibility, sum log s
  T mult_sum;           // |nu|_1 + |lam|_1
  bool ok, pole, finite;
};

// One backward sweep at the current iterate: model derivatives, adjoint multipliers + KKT residuals,
// condensed stage QP, Riccati factorisation.  EXACT selects the Hessian.  The barrier parameter enters the
// Newton system only through the gradient  -mu G'(1/s); the vector recursion is therefore carried twice
// (mu-independent part and coefficient of mu) so that mu is chosen *after* the sweep, from the residuals
// the sweep itself measured.
template <class M, class T, class TF, bool EXACT>
RMPC_HD void backward_sweep(Work<M, T>& w, const ParamView<T>& pv, T reg, SweepOut<M, T>& o) {
  constexpr int NX = M::NX, NU = M::NU, NV = M::NVAR, NH = M::NH, N = M::N;
  TF P[NX * NX], p0[NX], p1[NX];
  T nu[NX];
  RMPC_UNROLL
  for (int i = 0; i < NX * NX; ++i) P[i] = TF(0);
  RMPC_UNROLL
  for (int i = 0; i < NX; ++i) { p0[i] = TF(0); p1[i] = TF(0); nu[i] = T(0); }
  o.res_stat = o.res_eq = o.res_ineq = T(0);
  o.comp_min = T(1e30); o.comp_max = T(0);
  o.f = o.theta = o.logs = o.mult_sum = T(0);
  o.ok = true; o.pole = false; o.finite = true;

  for (int k = N - 1; k >= 0; --k) {
    T p[M::NPAR];
    load_params<M, T>(pv, k, p);
    T zk[NV];
    RMPC_UNROLL
    for (int i = 0; i < NV; ++i) zk[i] = w.z[k][i];
    // --- barrier quantities
    T sig[NH], y0[NH], y1[NH], lamk[NH];
    T rho[M::NGEO > 0 ? M::NGEO : 1], kappa[M::NGEO > 0 ? M::NGEO : 1];
    RMPC_UNROLL
    for (int r = 0; r < NH; ++r) { lamk[r] = w.lam[k][r]; y1[r] = T(1) / w.s[k][r]; sig[r] = lamk[r] * y1[r]; }
    RMPC_UNROLL
    for (int r = 0; r < M::NGEO; ++r) { rho[r] = sig[M::geo_row(r)]; kappa[r] = -lamk[M::geo_row(r)]; }
    StageDer<M, T> sd;
    stage_derivatives<M, T, EXACT>(zk, p, rho, kappa, sd);
    o.f += sd.f;
    o.pole = o.pole || sd.pole;
    RMPC_UNROLL
    for (int r = 0; r < NH; ++r) {
      const T sr = w.s[k][r];
      const T rh = sd.h[r] - sr;
      w.hh[k][r] = sd.h[r];
      y0[r] = lamk[r] - sig[r] * sd.h[r];     // y = y0 + mu * y1
      o.res_ineq = tmax(o.res_ineq, tabs(rh));
      o.theta += tabs(rh);
      o.comp_max = tmax(o.comp_max, sr * lamk[r]);
      o.comp_min = tmin(o.comp_min, sr * lamk[r]);
      o.logs += tlog(sr);
      o.mult_sum += tabs(lamk[r]);
    }
    RMPC_UNROLL
    for (int r = 0; r < M::NGEO; ++r)
      RMPC_UNROLL
      for (int i = 0; i < M::NQ; ++i) w.Gg[k][r][i] = sd.Gg[r][i];
    RMPC_UNROLL
    for (int i = 0; i < NV; ++i) w.gf[k][i] = sd.g[i];

    // --- stage Hessian Q (NV x NV); QP gradient qt0 + mu*qt1 = grad f - G'(y0 + mu y1); Lagrangian gradient gl
    T Q[NV * NV];
    RMPC_UNROLL
    for (int i = 0; i < NV * NV; ++i) Q[i] = T(0);
    RMPC_UNROLL
    for (int i = 0; i < M::NQ; ++i)
      RMPC_UNROLL
      for (int j = 0; j < M::NQ; ++j) Q[i * NV + j] = sd.Hqq[i * M::NQ + j];
    RMPC_UNROLL
    for (int i = 0; i < NV; ++i) Q[i * NV + i] += sd.Hd[i];
    RMPC_UNROLL
    for (int r = 0; r < M::NLIM; ++r) Q[M::lim_var(r) * NV + M::lim_var(r)] += sig[M::lim_row(r)];
    RMPC_UNROLL
    for (int j = 0; j < NV; ++j) Q[j * NV + j] += reg;
    T qt0[NV], qt1[NV], gl[NV];
    RMPC_UNROLL
    for (int i = 0; i < NV; ++i) { qt0[i] = sd.g[i]; gl[i] = sd.g[i]; qt1[i] = T(0); }
    sub_Gt<M, T>(w, k, y0, qt0);
    sub_Gt<M, T>(w, k, y1, qt1);
    sub_Gt<M, T>(w, k, lamk, gl);

    TF Fxx[NX * NX], Fux[NU * NX], Fuu[NU * NU], fx0[NX], fu0[NU], fx1[NX], fu1[NU];
    if (k < N - 1) {
      // --- dynamics to the next stage: defect d, Jacobian [A B], adjoint, curvature
      T xn[NX], AB[NX * NV];
      dyn_derivatives<M, T, EXACT>(zk, xn, AB, nu, Q);
      T d[NX];
      RMPC_UNROLL
      for (int i = 0; i < NX; ++i) {
        d[i] = xn[i] - w.z[k + 1][i];
        o.res_eq = tmax(o.res_eq, tabs(d[i]));
        o.theta += tabs(d[i]);
        o.mult_sum += tabs(nu[i]);
      }
      // adjoint: gl += [A B]' nu_{k+1}
      RMPC_UNROLL
      for (int j = 0; j < NV; ++j) {
        T a = T(0);
        RMPC_UNROLL
        for (int i = 0; i < NX; ++i) a += AB[i * NV + j] * nu[i];
        gl[j] += a;
      }
      // Riccati products (factorisation precision TF)
      TF Pd[NX];
      RMPC_UNROLL
      for (int i = 0; i < NX; ++i) {
        TF a = p0[i];
        RMPC_UNROLL
        for (int j = 0; j < NX; ++j) a += P[i * NX + j] * TF(d[j]);
        Pd[i] = a;
      }
      TF PAB[NX * NV];
      RMPC_UNROLL
      for (int i = 0; i < NX; ++i)
        RMPC_UNROLL
        for (int j = 0; j < NV; ++j) {
          TF a = TF(0);
          RMPC_UNROLL
          for (int l = 0; l < NX; ++l) a += P[i * NX + l] * TF(AB[l * NV + j]);
          PAB[i * NV + j] = a;
        }
      RMPC_UNROLL
      for (int i = 0; i < NV; ++i) {
        TF g0 = TF(qt0[i]), g1 = TF(qt1[i]);
        RMPC_UNROLL
        for (int l = 0; l < NX; ++l) { g0 += TF(AB[l * NV + i]) * Pd[l]; g1 += TF(AB[l * NV + i]) * p1[l]; }
        if (i < NX) { fx0[i < NX ? i : 0] = g0; fx1[i < NX ? i : 0] = g1; }
        else { fu0[i >= NX ? i - NX : 0] = g0; fu1[i >= NX ? i - NX : 0] = g1; }
        RMPC_UNROLL
        for (int j = 0; j <= i; ++j) {
          TF a = TF(Q[i * NV + j]);
          RMPC_UNROLL
          for (int l = 0; l < NX; ++l) a += TF(AB[l * NV + i]) * PAB[l * NV + j];
          if (i < NX) { Fxx[(i < NX ? i : 0) * NX + j] = a; Fxx[j * NX + (i < NX ? i : 0)] = a; }
          else if (j < NX) Fux[(i >= NX ? i - NX : 0) * NX + j] = a;
          else { Fuu[(i - NX) * NU + (j - NX)] = a; Fuu[(j - NX) * NU + (i - NX)] = a; }
        }
      }
    } else {
      RMPC_UNROLL
      for (int i = 0; i < NV; ++i) {
        if (i < NX) { fx0[i < NX ? i : 0] = TF(qt0[i]); fx1[i < NX ? i : 0] = TF(qt1[i]); }
        else { fu0[i >= NX ? i - NX : 0] = TF(qt0[i]); fu1[i >= NX ? i - NX : 0] = TF(qt1[i]); }
        RMPC_UNROLL
        for (int j = 0; j <= i; ++j) {
          const TF a = TF(Q[i * NV + j]);
          if (i < NX) { Fxx[(i < NX ? i : 0) * NX + j] = a; Fxx[j * NX + (i < NX ? i : 0)] = a; }
          else if (j < NX) Fux[(i >= NX ? i - NX : 0) * NX + j] = a;
          else { Fuu[(i - NX) * NU + (j - NX)] = a; Fuu[(j - NX) * NU + (i - NX)] = a; }
        }
      }
    }
    // stationarity residual (inputs) and adjoint for the next (earlier) stage
    RMPC_UNROLL
    for (int j = 0; j < NU; ++j) o.res_stat = tmax(o.res_stat, tabs(gl[NX + j]));
    RMPC_UNROLL
    for (int i = 0; i < NX; ++i) nu[i] = gl[i];

    TF Kk[NU * NX], kf0[NU], kf1[NU];
    if (!factor_solve<M, TF>(Fuu, Fux, fu0, fu1, Kk, kf0, kf1)) { o.ok = false; return; }
    RMPC_UNROLL
    for (int i = 0; i < NU * NX; ++i) w.K[k][i] = T(Kk[i]);
    RMPC_UNROLL
    for (int i = 0; i < NU; ++i) { w.kff0[k][i] = T(kf0[i]); w.kff1[k][i] = T(kf1[i]); }
    // P_k = Fxx + Fux' K ,  p_k = fx + Fux' kff
    RMPC_UNROLL
    for (int i = 0; i < NX; ++i) {
      TF a0 = fx0[i], a1 = fx1[i];
      RMPC_UNROLL
      for (int l = 0; l < NU; ++l) { a0 += Fux[l * NX + i] * kf0[l]; a1 += Fux[l * NX + i] * kf1[l]; }
      p0[i] = a0; p1[i] = a1;
      RMPC_UNROLL
      for (int j = 0; j <= i; ++j) {
        TF v = Fxx[i * NX + j];
        RMPC_UNROLL
        for (int l = 0; l < NU; ++l) v += Fux[l * NX + i] * Kk[l * NX + j];
        P[i * NX + j] = v;
        P[j * NX + i] = v;
      }
    }
  }
  o.finite = (o.f == o.f) && (o.res_stat == o.res_stat) && (o.theta == o.theta);
}

// -------------------------------------------------------------------------------------------------
template <class M, class T, class TF>
RMPC_HD void solve_one(Work<M, T>& w, const T* xinit, const ParamView<T>& pv, const Options& opt, Info& info) {
  constexpr int NX = M::NX, NU = M::NU, NV = M::NVAR, NH = M::NH, N = M::N;
  T mu = T(opt.mu0);
  // ---- initial point: x_1 = xinit; slacks from h(z0); lambda = mu / s
  RMPC_UNROLL
  for (int i = 0; i < NX; ++i) w.z[0][i] = xinit[i];
  for (int k = 0; k < N; ++k) {
    T p[M::NPAR];
    load_params<M, T>(pv, k, p);
    StageVal<M, T> sv;
    stage_values<M, T>(w.z[k], p, sv);
    RMPC_UNROLL
    for (int r = 0; r < NH; ++r) {
      const T sr = tmax(sv.h[r], T(1e-2));
      w.s[k][r] = sr;
      w.lam[k][r] = mu / sr;
    }
  }
  T rho_pen = T(1);
  info.exitflag = 0;
  int it = 0;
  SweepOut<M, T> so;
  T res_comp = T(0);
  int inner = 0;
  for (; it <= opt.max_iter; ++it) {
    bool ok = false;
    if (opt.hessian == 1) {
      backward_sweep<M, T, TF, true>(w, pv, T(opt.reg), so);
      ok = so.ok;
    }
    if (!ok) {
      // Gauss-Newton Hessian; if rounding still breaks positive definiteness (huge barrier weights), add a
      // growing Levenberg term
      T reg = T(opt.reg);
      for (int tries = 0; tries < 4; ++tries) {
        backward_sweep<M, T, TF, false>(w, pv, reg, so);
        if (so.ok) break;
        reg = tmax(reg * T(100), T(1e-6));
      }
    }
    if (!so.finite || !so.ok) { info.exitflag = so.ok ? -6 : -5; break; }
    if (so.comp_max > T(1e15)) { info.exitflag = -7; break; }  // multipliers diverge: locally infeasible
    // ---- convergence test (IPOPT scaling of the dual residual)
    const T sd = tmax(T(100), so.mult_sum / T(N * (NH + NX))) / T(100);
    res_comp = so.comp_max;
    if (so.res_stat / sd <= T(opt.tol_stat) && so.res_eq <= T(opt.tol_eq) && so.res_ineq <= T(opt.tol_ineq) &&
        so.comp_max / sd <= T(opt.tol_comp)) {
      info.exitflag = 1;
      break;
    }
    if (it == opt.max_iter) break;
    // ---- monotone barrier update (Fiacco-McCormick as in IPOPT): shrink mu while the current barrier
    //      problem is solved to within kappa_eps * mu
    ++inner;
    for (int rep = 0; rep < 8; ++rep) {
      const T cmu = tmax(so.comp_max - mu, mu - so.comp_min);
      const T emu = tmax(tmax(so.res_stat / sd, cmu / sd), tmax(so.res_eq, so.res_ineq));
      const bool force = opt.max_inner > 0 && inner > opt.max_inner && rep == 0;
      if ((emu <= T(10) * mu || force) && mu > T(opt.mu_min)) {
        inner = 0;
        const T m15 = mu * tsqrt(mu);
        mu = tmax(T(opt.mu_min), tmin(T(0.2) * mu, m15));
      } else break;
    }

    // ---- forward roll-out of the Newton step; fraction-to-boundary step lengths; merit slope
    const T tau = tmax(T(0.99), T(1) - mu);
    T ap = T(1), ad = T(1);
    T dphi = T(0);
    {
      T dx[NX];
      RMPC_UNROLL
      for (int i = 0; i < NX; ++i) dx[i] = T(0);  // x_1 is fixed
      for (int k = 0; k < N; ++k) {
        T du[NU];
        RMPC_UNROLL
        for (int j = 0; j < NU; ++j) {
          T a = w.kff0[k][j] + mu * w.kff1[k][j];
          RMPC_UNROLL
          for (int i = 0; i < NX; ++i) a += w.K[k][j * NX + i] * dx[i];
          du[j] = a;
        }
        RMPC_UNROLL
        for (int i = 0; i < NX; ++i) w.dz[k][i] = dx[i];
        RMPC_UNROLL
        for (int j = 0; j < NU; ++j) w.dz[k][NX + j] = du[j];
        RMPC_UNROLL
        for (int i = 0; i < NV; ++i) dphi += w.gf[k][i] * w.dz[k][i];
        T gd[NH];
        row_dirs<M, T>(w, k, gd);
        RMPC_UNROLL
        for (int r = 0; r < NH; ++r) {
          const T sr = w.s[k][r], lr = w.lam[k][r];
          const T dsr = gd[r] + w.hh[k][r] - sr;
          const T dlr = mu / sr - lr - (lr / sr) * dsr;
          if (dsr < T(0)) ap = tmin(ap, -tau * sr / dsr);
          if (dlr < T(0)) ad = tmin(ad, -tau * lr / dlr);
          dphi -= mu * dsr / sr;
        }
        if (k < N - 1) {
          // linearised dynamics: dx+ = [A B] dz + d
          T xn[NX], AB[NX * NV];
          T zk[NV];
          RMPC_UNROLL
          for (int i = 0; i < NV; ++i) zk[i] = w.z[k][i];
          dyn_derivatives<M, T, false>(zk, xn, AB, (const T*)0, (T*)0);
          RMPC_UNROLL
          for (int i = 0; i < NX; ++i) {
            T a = xn[i] - w.z[k + 1][i];
            RMPC_UNROLL
            for (int j = 0; j < NV; ++j) a += AB[i * NV + j] * w.dz[k][j];
            dx[i] = a;
          }
        }
      }
    }
    // ---- l1 merit: penalty update and backtracking line search on (z, s)
    if (so.theta > T(1e-14)) {
      const T need = dphi / (T(0.9) * so.theta);
      if (need > rho_pen) rho_pen = need * T(1.1) + T(1e-3);
    }
    const T phi0 = so.f - mu * so.logs + rho_pen * so.theta;
    const T slope = dphi - rho_pen * so.theta;
    T alpha = ap;
    bool accepted = false;
    for (int ls = 0; ls < 30; ++ls) {
      T f = T(0), theta = T(0), logs = T(0);
      bool pole = false;
      T xn[NX];
      for (int k = 0; k < N; ++k) {
        T p[M::NPAR];
        load_params<M, T>(pv, k, p);
        T zt[NV];
        RMPC_UNROLL
        for (int i = 0; i < NV; ++i) zt[i] = w.z[k][i] + alpha * w.dz[k][i];
        if (k > 0) {
          RMPC_UNROLL
          for (int i = 0; i < NX; ++i) theta += tabs(xn[i] - zt[i]);
        }
        StageVal<M, T> sv;
        stage_values<M, T>(zt, p, sv);
        f += sv.f;
        pole = pole || sv.pole;
        T gd[NH];
        row_dirs<M, T>(w, k, gd);
        RMPC_UNROLL
        for (int r = 0; r < NH; ++r) {
          const T sr = w.s[k][r];
          const T st = sr + alpha * (gd[r] + w.hh[k][r] - sr);
          theta += tabs(sv.h[r] - st);
          logs += tlog(st);
        }
        if (k < N - 1) dyn_value<M, T>(zt, xn);
      }
      const T phi = f - mu * logs + rho_pen * theta;
      const T tol_phi = T(10) * T(sizeof(T) == 8 ? 2.2e-16 : 1.2e-7) * tabs(phi0);
      if (!pole && (phi == phi) && phi <= phi0 + T(1e-4) * alpha * slope + tol_phi) { accepted = true; break; }
      alpha *= T(0.5);
    }
#if defined(RMPC_TRACE) && !defined(__CUDA_ARCH__)
    printf("it %d mu %.3g f %.9g theta %.3g stat %.3g comp %.3g..%.3g | ap %.3g ad %.3g alpha %.3g dphi %.3g rho %.3g slope %.3g acc %d\n", it,
           double(mu), double(so.f), double(so.theta), double(so.res_stat), double(so.comp_min), double(so.comp_max), double(ap), double(ad),
           double(alpha), double(dphi), double(rho_pen), double(slope), int(accepted));
#endif
    if (!accepted) { info.exitflag = -7; break; }
    // ---- take the step
    const T adu = tmin(ad, T(1));
    for (int k = 0; k < N; ++k) {
      T gd[NH];
      row_dirs<M, T>(w, k, gd);
      RMPC_UNROLL
      for (int r = 0; r < NH; ++r) {
        const T sr = w.s[k][r], lr = w.lam[k][r];
        const T dsr = gd[r] + w.hh[k][r] - sr;
        const T dlr = mu / sr - lr - (lr / sr) * dsr;
        w.s[k][r] = sr + alpha * dsr;
        w.lam[k][r] = lr + adu * dlr;
      }
      RMPC_UNROLL
      for (int i = 0; i < NV; ++i) w.z[k][i] += alpha * w.dz[k][i];
    }
  }
  info.iters = it;
  info.obj = double(so.f);
  info.res_stat = double(so.res_stat);
  info.res_eq = double(so.res_eq);
  info.res_ineq = double(so.res_ineq);
  info.res_comp = double(res_comp);
  info.mu = double(mu);
}

}  // namespace rmpc
