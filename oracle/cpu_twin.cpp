// ORACLE / CPU BASELINE — test and benchmark infrastructure only, never on the product path.
//
// "C++ CPU twin" of SURVEY.md §8d / BASELINE.md §3: the same interior-point + Riccati algorithm and the same
// model functions as the CUDA solver (it includes the product headers), compiled for the host in float64,
// one problem per thread, OpenMP over the batch.  It stands in for the reference's CPU path — the
// FORCESPRO-generated C solver called at planner/mpcPlanner.py:262, which is proprietary and absent — as the
// timed CPU baseline (`cpu_baseline.kind = "port"`), and gives a CUDA-independent execution of the algorithm
// for debugging.  The independent correctness oracle is oracle/solve.py (scipy), not this file.
//
// build: g++ -O3 -march=native -fopenmp -std=c++17 -shared -fPIC -I <solver_dir> -I robot_mpcs_b200/csrc
#include <omp.h>
#include <cstring>
#include <vector>
#include "model_traits.h"
#include "rmpc_ipm.cuh"

using M = ModelTraits;

template <class T>
static void run(int B, const double* xinit, const double* x0, const double* params, double* z_out, int* exitflag,
                int* iters, double* resid, const rmpc::Options& opt, bool mixed) {
#pragma omp parallel
  {
    auto* S = new rmpc::ProblemShared<M, T>();
    auto* Sf = new rmpc::ProblemShared<M, float>();
    std::vector<float> hz(M::N * M::NVAR), hlam(M::N * M::NH);
    rmpc::Group<1> g{0};
#pragma omp for schedule(dynamic, 8)
    for (int b = 0; b < B; ++b) {
      const double* xi = xinit + size_t(b) * M::NX;
      const double* zin = x0 + size_t(b) * M::N * M::NVAR;
      double* zo = z_out + size_t(b) * M::N * M::NVAR;
      rmpc::ParamView<T, double> pv{params + size_t(b) * M::N * M::NPAR, M::NPAR, 1};
      rmpc::Info info;
      auto* ipm = new rmpc::Ipm<M, T, 1, double>(g, *S, pv, opt);
      if (mixed) {   // precision 1: float32 iterations while the barrier parameter is large, float64 to finish (rmpc_ipm.cuh)
        using If = rmpc::Ipm<M, float, 1, double>;
        rmpc::ParamView<float, double> pvf{pv.base, M::NPAR, 1};
        const rmpc::Options lo = If::lowp_options(opt);
        auto* f = new If(g, *Sf, pvf, lo);
        float hmu = 0.0f;
        int hit = -1;
        f->solve_lowp(xi, zin, hz.data(), hlam.data(), &hmu, &hit);
        delete f;
        if constexpr (sizeof(T) == 8) ipm->solve_resume(xi, zin, zo, hz.data(), hlam.data(), hmu, hit, info);
      } else {
        ipm->solve(xi, zin, zo, info);
      }
      delete ipm;
      exitflag[b] = info.exitflag;
      iters[b] = info.iters;
      if (resid) {
        double* r = resid + size_t(b) * 6;
        r[0] = info.obj; r[1] = info.res_stat; r[2] = info.res_eq; r[3] = info.res_ineq; r[4] = info.res_comp; r[5] = info.mu;
      }
    }
    delete S; delete Sf;
  }
}

extern "C" {
const char* twin_name() { return M::NAME; }
void twin_dims(int* d) { d[0] = M::N; d[1] = M::NX; d[2] = M::NU; d[3] = M::NVAR; d[4] = M::NPAR; d[5] = M::NH; }
int twin_max_threads() { return omp_get_max_threads(); }

// opts: [max_iter, hessian, tol_stat, tol_eq, tol_ineq, tol_comp, mu0, mu_min, reg, delta, kappa, tau_min]; precision: 0 f64, 1 f32 then f64 (rmpc_options.precision = 1), 2 f32 throughout (experiments)
int twin_solve_batch(int B, const double* xinit, const double* x0, const double* params, double* z_out, int* exitflag,
                     int* iters, double* resid, const double* opts, int nthreads, int precision) {
  rmpc::Options o = rmpc::default_options();
  if (opts) {
    o.max_iter = int(opts[0]); o.hessian = int(opts[1]); o.tol_stat = opts[2]; o.tol_eq = opts[3]; o.tol_ineq = opts[4];
    o.tol_comp = opts[5]; o.mu0 = opts[6]; o.mu_min = opts[7]; o.reg = opts[8]; o.delta = opts[9]; o.kappa = opts[10]; o.tau_min = opts[11]; o.lm0 = opts[12]; o.lm_up = opts[13]; o.lm_down = opts[14]; o.lm_trigger = opts[15]; o.creep_trigger = opts[16]; o.creep_factor = opts[17]; o.stall_window = opts[18]; o.gn_hold_max = opts[19]; o.ls_optimistic = opts[20]; o.soc = opts[21];
  }
  if (nthreads > 0) omp_set_num_threads(nthreads);
  if (precision == 2) run<float>(B, xinit, x0, params, z_out, exitflag, iters, resid, o, false);
  else run<double>(B, xinit, x0, params, z_out, exitflag, iters, resid, o, precision == 1);
  return 0;
}

// model-function probes for the derivative / FK parity tests
void twin_sincos(int n, const double* x, double* s, double* c) { for (int i = 0; i < n; ++i) rmpc::tsincos(x[i], s + i, c + i); }
void twin_stage_values(const double* z, const double* p, double* f, double* h) {
  rmpc::StageVal<M, double> sv;
  rmpc::stage_values<M, double>(z, p, sv);
  *f = sv.f;
  for (int i = 0; i < M::NH; ++i) h[i] = sv.h[i];
}
void twin_dynamics(const double* z, double* xn, double* AB) {
  double ABv[rmpc::DynS<M>::NABV], Hc[5][5];
  rmpc::dyn_eval<M, double>(z, xn, ABv, false, 0.0, 0.0, Hc);
  rmpc::dyn_dense<M, double>(ABv, AB);
}
// exact: Hqq includes rho/kappa terms; returns cost grad g[NVAR], Hqq[NQ*NQ], Hd[NVAR], h[NH], Gg[NGEO*NQ]
void twin_stage_derivatives(const double* z, const double* p, const double* rho, const double* kappa, int exact,
                            double* f, double* g, double* Hqq, double* Hd, double* h, double* Gg) {
  rmpc::StageDer<M, double> sd;
  rmpc::stage_derivatives<M, double>(z, p, rho, kappa, exact != 0, sd);
  *f = sd.f;
  for (int i = 0; i < M::NVAR; ++i) { g[i] = sd.g[i]; Hd[i] = sd.Hd[i]; }
  for (int i = 0; i < M::NQ * M::NQ; ++i) Hqq[i] = sd.Hqq[i];
  for (int i = 0; i < M::NH; ++i) h[i] = sd.h[i];
  for (int r = 0; r < M::NGEO; ++r)
    for (int i = 0; i < M::NQ; ++i) Gg[r * M::NQ + i] = sd.Gg[r][i];
}
// Q (NVAR x NVAR, row-major) += nu-weighted second derivative of Phi (diff-drive only; a no-op otherwise)
void twin_dyn_curvature(const double* z, const double* nu, double* Q) {
  double xn[M::NX], ABv[rmpc::DynS<M>::NABV], Hc[5][5];
  rmpc::dyn_eval<M, double>(z, xn, ABv, true, nu[0], nu[1], Hc);
  if (M::BASE == 1)
    for (int a = 0; a < 5; ++a)
      for (int b = 0; b < 5; ++b) Q[rmpc::DynS<M>::vc(a) * M::NVAR + rmpc::DynS<M>::vc(b)] += Hc[a][b];
}
}
