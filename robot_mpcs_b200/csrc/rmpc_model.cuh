// rmpc_model.cuh — per-stage model functions of the robotmpcs NLP as fused device functions with analytic
// Jacobians and (Gauss-Newton or exact) Hessians.  Everything is templated on a generated `M` traits struct
// (emitted by robot_mpcs_b200/models/mpcModel.py:traits_source) so the compiler sees every dimension, every kinematic constant and
// every parameter index as an immediate.
//
// What is restated here (reference paths relative to /root/reference):
//   forward kinematics      forwardkinematics.GenericURDFFk, call sites models/mpcBase.py:46-51,89-94
//   RadialConstraints       models/mpcBase.py:82-101, inequalities/RadialConstraints.py
//   LinearConstraints       inequalities/LinearConstraints.py:25-40, utils/utils.py:48-52
//   SelfCollision...        inequalities/SelfCollisionAvoidanceConstraints.py:18-27
//   Joint/Input/VelLimit    inequalities/{JointLimit,InputLimit,VelLimit}Constraints.py
//   GoalReaching            objectives/goal_reaching.py:19-34
//   ConstraintAvoidance     objectives/constraint_avoidance.py:25-32
//   objective               objectives/ObjectiveManager.py:28-46
//   dynamics                models/mpcModel.py:65-69,118-120, models/diff_drive_mpc_model.py:24-41
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define RMPC_HD __host__ __device__ __forceinline__
#else
#define RMPC_HD inline __attribute__((always_inline))
#endif

#if defined(__CUDA_ARCH__)
#define RMPC_UNROLL _Pragma("unroll")
#else
#define RMPC_UNROLL
#endif

namespace rmpc {

template <class T> RMPC_HD T tmax(T a, T b) { return a > b ? a : b; }
template <class T> RMPC_HD T tmin(T a, T b) { return a < b ? a : b; }
RMPC_HD float tabs(float a) { return fabsf(a); }
RMPC_HD double tabs(double a) { return fabs(a); }
RMPC_HD float fdiv_approx(float a, float b) {  // 2 ulp: for ratios that only cap a step length
#if defined(__CUDA_ARCH__)
  return __fdividef(a, b);
#else
  return a / b;
#endif
}
RMPC_HD float tsqrt(float a) { return sqrtf(a); }
RMPC_HD double tsqrt(double a) { return sqrt(a); }
RMPC_HD float tlog(float a) { return logf(a); }
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ float trcp(float a) { return __frcp_rn(a); }
__device__ __forceinline__ float trsqrt(float a) { return rsqrtf(a); }
__device__ __forceinline__ void tsincos(float a, float* s, float* c) { sincosf(a, s, c); }
#else
RMPC_HD float trcp(float a) { return 1.0f / a; }
RMPC_HD float trsqrt(float a) { return 1.0f / sqrtf(a); }
RMPC_HD void tsincos(float a, float* s, float* c) { *s = sinf(a); *c = cosf(a); }
#endif
#if defined(__CUDA_ARCH__)
// Device: the libm expansions of sincos / log are ~150 / ~60 instructions each; one out-of-line copy keeps the solve
// kernel inside the instruction cache.  Reciprocal and reciprocal square root: the 64-bit MUFU seed (rcp/rsqrt
// .approx.ftz.f64, about 20 bits, one instruction, no conversions) + two Newton steps in double: a few ulp
// (tests/test_gpu_parity.py::test_device_reciprocals) instead of the IEEE division / square-root sequences with
// their slow-path branches.  Arguments are positive normal numbers here.
__device__ __noinline__ void tsincos_lib(double a, double* s, double* c) { sincos(a, s, c); }
__device__ __noinline__ double tlog(double a) { return log(a); }
__device__ __forceinline__ double trcp(double a) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a));
  r = r * (2.0 - a * r);
  r = r * (2.0 - a * r);
  return r;
}
__device__ __forceinline__ double trsqrt(double a) {
  double r;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a));
  const double ha = 0.5 * a;
  r = r * (1.5 - ha * r * r);
  r = r * (1.5 - ha * r * r);
  return r;
}
#else
RMPC_HD void tsincos_lib(double a, double* s, double* c) { *s = sin(a); *c = cos(a); }
RMPC_HD double tlog(double a) { return log(a); }
RMPC_HD double trcp(double a) { return 1.0 / a; }
RMPC_HD double trsqrt(double a) { return 1.0 / sqrt(a); }
#endif

// sin and cos together: Cody-Waite reduction by pi/2 in two pieces (exact for |x| < 1e5) and the fdlibm kernel
// polynomials on [-pi/4, pi/4]; ~35 instructions against ~90 plus a call for the library routine, which remains the
// path for huge arguments.  Accuracy ~1 ulp of the result's scale (tests/test_oracle_model.py::test_fast_sincos).
#if defined(__CUDA_ARCH__)
__device__ __noinline__   // one copy: the solve kernel is instruction-cache bound (inlined at its 12 call sites it ran 7-18 % slower)
#else
RMPC_HD
#endif
void tsincos(double x, double* sn, double* cs) {
  if (!(tabs(x) < 1e5)) { tsincos_lib(x, sn, cs); return; }
  const double magic = 6755399441055744.0;  // 1.5 * 2^52: adding it rounds to the nearest integer
  const double t = x * 0.63661977236758134308 + magic;
  const double kd = t - magic;
#if defined(__CUDA_ARCH__)
  const int q = __double2loint(t);
  double r = __fma_rn(-kd, 1.57079632673412561417e+00, x);
  r = __fma_rn(-kd, 6.07710050650619224932e-11, r);
#else
  long long tb;
  __builtin_memcpy(&tb, &t, 8);
  const int q = int(tb & 0xffffffffLL);
  double r = __builtin_fma(-kd, 1.57079632673412561417e+00, x);
  r = __builtin_fma(-kd, 6.07710050650619224932e-11, r);
#endif
  const double z = r * r;
  double ps = 1.58969099521155010221e-10;
  ps = ps * z + -2.50507602534068634195e-08;
  ps = ps * z + 2.75573137070700676789e-06;
  ps = ps * z + -1.98412698298579493134e-04;
  ps = ps * z + 8.33333333332248946124e-03;
  ps = ps * z + -1.66666666666666324348e-01;
  const double sr = r + (r * z) * ps;
  double pc = -1.13596475577881948265e-11;
  pc = pc * z + 2.08757232129817482790e-09;
  pc = pc * z + -2.75573143513906633035e-07;
  pc = pc * z + 2.48015872894767294178e-05;
  pc = pc * z + -1.38888888888741095749e-03;
  pc = pc * z + 4.16666666666666019037e-02;
  const double cr = (1.0 - 0.5 * z) + (z * z) * pc;
  const double a = (q & 1) ? cr : sr, b = (q & 1) ? sr : cr;
  *sn = (q & 2) ? -a : a;
  *cs = ((q + 1) & 2) ? -b : b;
}

// sum_i log(x_i) as log(prod mantissas) + ln2 * sum exponents: one log per stage instead of one per row.
template <class T> struct LogAcc;
template <> struct LogAcc<double> {
  double m, lg;
  RMPC_HD LogAcc() : m(1.0), lg(0.0) {}
  // x > 0.  A plain running product: the caller flushes it into a logarithm every 12 factors (flush_every), which keeps
  // 12 slacks between 1e-16 and 1e16 clear of under- and overflow; a factor of 0 (a trial point outside the domain)
  // gives -inf, which the line search rejects.
  RMPC_HD void mul(double x) { m *= x; }
  RMPC_HD void flush() { lg += tlog(m); m = 1.0; }
  RMPC_HD double value() const { return lg + tlog(m); }
};
template <> struct LogAcc<float> {
  float a;
  RMPC_HD LogAcc() : a(0.0f) {}
  RMPC_HD void mul(float x) { a += logf(x); }
  RMPC_HD void flush() {}
  RMPC_HD float value() const { return a; }
};

// ------------------------------------------------------------------------------------------------
// Kinematics: world frames of the NJ tree joints that matter, positions of the NPT points of interest
// and their position Jacobians w.r.t. q.
// ------------------------------------------------------------------------------------------------
template <class M, class T>
struct Kin {
  T R[M::NJ > 0 ? M::NJ : 1][9];   // world rotation of the child-link frame of joint j
  T t[M::NJ > 0 ? M::NJ : 1][3];   // world origin of that frame
  T a[M::NJ > 0 ? M::NJ : 1][3];   // world axis of joint j
  T p[M::NPT > 0 ? M::NPT : 1][3];            // points of interest
  T Jp[M::NPT > 0 ? M::NPT : 1][3][M::NQ];    // d p / d q
};

template <class M, class T>
RMPC_HD void kin_frames(const T* q, Kin<M, T>& kn) {
  RMPC_UNROLL
  for (int j = 0; j < M::NJ; ++j) {
    const int par = M::j_parent(j);
    T Rp[9], tp[3];
    if (par < 0) {
      RMPC_UNROLL
      for (int i = 0; i < 9; ++i) Rp[i] = T((i % 4) == 0 ? 1 : 0);
      tp[0] = tp[1] = tp[2] = T(0);
    } else {
      RMPC_UNROLL
      for (int i = 0; i < 9; ++i) Rp[i] = kn.R[par][i];
      RMPC_UNROLL
      for (int i = 0; i < 3; ++i) tp[i] = kn.t[par][i];
    }
    T Ro[9], to[3];
    RMPC_UNROLL
    for (int r = 0; r < 3; ++r) {
      to[r] = tp[r];
      RMPC_UNROLL
      for (int c = 0; c < 3; ++c) {
        if (M::j_T(j, c) != 0.0) to[r] += Rp[3 * r + c] * T(M::j_T(j, c));
        T acc = T(0);
        RMPC_UNROLL
        for (int k = 0; k < 3; ++k)
          if (M::j_R(j, 3 * k + c) != 0.0) acc += Rp[3 * r + k] * T(M::j_R(j, 3 * k + c));
        Ro[3 * r + c] = acc;
      }
    }
    T aw[3];
    RMPC_UNROLL
    for (int r = 0; r < 3; ++r) {
      T acc = T(0);
      RMPC_UNROLL
      for (int k = 0; k < 3; ++k)
        if (M::j_axis(j, k) != 0.0) acc += Ro[3 * r + k] * T(M::j_axis(j, k));
      aw[r] = acc;
      kn.a[j][r] = acc;
    }
    const int typ = M::j_type(j);
    if (typ == 2) {  // prismatic
      const T qj = q[M::j_q(j)];
      RMPC_UNROLL
      for (int r = 0; r < 3; ++r) to[r] += aw[r] * qj;
    } else if (typ == 1) {  // revolute: Ro <- Ro * Rot(axis, q)  (Rodrigues, constant local axis)
      T s, c;
      tsincos(q[M::j_q(j)], &s, &c);
      const T v = T(1) - c;
      T Rj[9];
      const double ax = M::j_axis(j, 0), ay = M::j_axis(j, 1), az = M::j_axis(j, 2);
      Rj[0] = c + T(ax * ax) * v;      Rj[1] = T(ax * ay) * v - T(az) * s; Rj[2] = T(ax * az) * v + T(ay) * s;
      Rj[3] = T(ay * ax) * v + T(az) * s; Rj[4] = c + T(ay * ay) * v;      Rj[5] = T(ay * az) * v - T(ax) * s;
      Rj[6] = T(az * ax) * v - T(ay) * s; Rj[7] = T(az * ay) * v + T(ax) * s; Rj[8] = c + T(az * az) * v;
      T Rn[9];
      RMPC_UNROLL
      for (int r = 0; r < 3; ++r)
        RMPC_UNROLL
        for (int cc = 0; cc < 3; ++cc)
          Rn[3 * r + cc] = Ro[3 * r] * Rj[cc] + Ro[3 * r + 1] * Rj[3 + cc] + Ro[3 * r + 2] * Rj[6 + cc];
      RMPC_UNROLL
      for (int i = 0; i < 9; ++i) Ro[i] = Rn[i];
    }
    RMPC_UNROLL
    for (int i = 0; i < 9; ++i) kn.R[j][i] = Ro[i];
    RMPC_UNROLL
    for (int i = 0; i < 3; ++i) kn.t[j][i] = to[i];
  }
}

template <class T> RMPC_HD void cross3(const T* a, const T* b, T* o) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}

template <class M, class T, bool JAC>
RMPC_HD void kin_points(const T* q, Kin<M, T>& kn) {
  kin_frames<M, T>(q, kn);
  RMPC_UNROLL
  for (int i = 0; i < M::NPT; ++i) {
    const int jl = M::pt_joint(i);
    RMPC_UNROLL
    for (int c = 0; c < 3; ++c) kn.p[i][c] = (jl < 0) ? T(0) : kn.t[jl][c];
    if (JAC) {
      RMPC_UNROLL
      for (int c = 0; c < 3; ++c)
        RMPC_UNROLL
        for (int k = 0; k < M::NQ; ++k) kn.Jp[i][c][k] = T(0);
      RMPC_UNROLL
      for (int j = 0; j < M::NJ; ++j) {
        if (!((M::pt_mask(i) >> j) & 1u)) continue;
        const int typ = M::j_type(j);
        if (typ == 0) continue;
        const int qi = M::j_q(j);
        if (typ == 2) {
          RMPC_UNROLL
          for (int c = 0; c < 3; ++c) kn.Jp[i][c][qi] = kn.a[j][c];
        } else {
          T d[3], cr[3];
          RMPC_UNROLL
          for (int c = 0; c < 3; ++c) d[c] = kn.p[i][c] - kn.t[j][c];
          cross3(kn.a[j], d, cr);
          RMPC_UNROLL
          for (int c = 0; c < 3; ++c) kn.Jp[i][c][qi] = cr[c];
        }
      }
    }
  }
}

// Sum_c w[c] * d^2 p_c / dq dq  added into Hqq (row-major NQ x NQ, both triangles).
template <class M, class T>
RMPC_HD void kin_point_curvature(const Kin<M, T>& kn, int i, const T* w, T* Hqq) {
  RMPC_UNROLL
  for (int k = 0; k < M::NJ; ++k) {
    if (!((M::pt_mask(i) >> k) & 1u)) continue;
    const int tk = M::j_type(k);
    if (tk == 0) continue;
    T inner[3];
    if (tk == 1) {
      T d[3];
      RMPC_UNROLL
      for (int c = 0; c < 3; ++c) d[c] = kn.p[i][c] - kn.t[k][c];
      cross3(kn.a[k], d, inner);
    } else {
      RMPC_UNROLL
      for (int c = 0; c < 3; ++c) inner[c] = kn.a[k][c];
    }
    RMPC_UNROLL
    for (int j = 0; j <= k; ++j) {  // j is an ancestor of (or equal to) k
      if (!((M::pt_mask(i) >> j) & 1u)) continue;
      if (M::j_type(j) != 1) continue;  // only an upstream revolute joint bends the downstream column
      T v[3];
      cross3(kn.a[j], inner, v);
      const T val = w[0] * v[0] + w[1] * v[1] + w[2] * v[2];
      const int qj = M::j_q(j), qk = M::j_q(k);
      Hqq[qj * M::NQ + qk] += val;
      if (qj != qk) Hqq[qk * M::NQ + qj] += val;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Stage functions
// ------------------------------------------------------------------------------------------------
constexpr int GEO_SPHERE = 0, GEO_PLANE = 1, GEO_PAIR = 2;

template <class M, class T>
struct StageVal {  // values only (line search)
  T f;
  T h[M::NH];
  bool pole;  // a ConstraintAvoidance row is on the wrong side of its pole (h <= 0)
};

template <class M, class T>
struct StageDer {
  T f;
  T g[M::NVAR];                      // cost gradient
  T Hqq[M::NQ * M::NQ];              // d2/dq2 of (cost + barrier rank-1 + optional curvature), full symmetric
  T Hd[M::NVAR];                     // diagonal part of the cost Hessian for every variable
  T h[M::NH];
  T Gg[M::NGEO > 0 ? M::NGEO : 1][M::NQ];  // gradients of the geometric rows w.r.t. q
  bool pole;
};

// geometric row r: value + unit direction(s) in point space
template <class M, class T>
RMPC_HD T geo_value(const Kin<M, T>& kn, const T* p, int r, T* nvec, T* dist_out) {
  const int typ = M::geo_type(r), A = M::geo_a(r);
  const T rb = p[M::geo_rb(r)];
  if (typ == GEO_SPHERE) {
    const int o = M::geo_p(r);
    T d[3] = {kn.p[A][0] - p[o], kn.p[A][1] - p[o + 1], kn.p[A][2] - p[o + 2]};
    const T d2 = tmax(d[0] * d[0] + d[1] * d[1] + d[2] * d[2], T(1e-24));
    const T inv = trsqrt(d2), dist = d2 * inv;
    nvec[0] = d[0] * inv; nvec[1] = d[1] * inv; nvec[2] = d[2] * inv;
    *dist_out = dist;
    return dist - p[o + 3] - rb;
  } else if (typ == GEO_PLANE) {
    const int o = M::geo_p(r);
    const T v = p[o] * kn.p[A][0] + p[o + 1] * kn.p[A][1] + p[o + 2] * kn.p[A][2] + p[o + 3];
    const T inrm = trsqrt(p[o] * p[o] + p[o + 1] * p[o + 1] + p[o + 2] * p[o + 2]);
    const T sg = v < T(0) ? -inrm : inrm;
    nvec[0] = sg * p[o]; nvec[1] = sg * p[o + 1]; nvec[2] = sg * p[o + 2];
    *dist_out = T(0);
    return tabs(v) * inrm - rb;
  } else {
    const int Bp = M::geo_b(r);
    T d[3] = {kn.p[A][0] - kn.p[Bp][0], kn.p[A][1] - kn.p[Bp][1], kn.p[A][2] - kn.p[Bp][2]};
    const T d2 = tmax(d[0] * d[0] + d[1] * d[1] + d[2] * d[2], T(1e-24));
    const T inv = trsqrt(d2), dist = d2 * inv;
    nvec[0] = d[0] * inv; nvec[1] = d[1] * inv; nvec[2] = d[2] * inv;
    *dist_out = dist;
    return dist - T(2) * rb;
  }
}

template <class M, class T>
RMPC_HD void stage_values(const T* z, const T* p, StageVal<M, T>& out) {
  Kin<M, T> kn;
  kin_points<M, T, false>(z, kn);
  RMPC_UNROLL
  for (int r = 0; r < M::NGEO; ++r) {
    T nv[3], dist;
    out.h[M::geo_row(r)] = geo_value<M, T>(kn, p, r, nv, &dist);
  }
  RMPC_UNROLL
  for (int r = 0; r < M::NLIM; ++r) {
    const T zv = z[M::lim_var(r)], pv = p[M::lim_p(r)];
    out.h[M::lim_row(r)] = M::lim_sign(r) > 0 ? zv - pv : pv - zv;
  }
  T f = T(0);
  if (M::GOAL_PT >= 0) {
    RMPC_UNROLL
    for (int c = 0; c < 3; ++c) {
      const T e = kn.p[M::GOAL_PT >= 0 ? M::GOAL_PT : 0][c] - p[M::P_GOAL + c];
      f += p[M::P_WGOAL + c] * e * e;
    }
  }
  bool pole = false;
  RMPC_UNROLL
  for (int i = 0; i < M::NCA; ++i) {
    const T w = p[M::ca_w(i)];
    const T hv = out.h[M::ca_row(i)];
    if (w != T(0)) {
      if (hv <= T(0)) pole = true;
      f += T(M::N) * w * trcp(hv);
    }
  }
  RMPC_UNROLL
  for (int j = 0; j < M::NU; ++j) f += p[M::P_WU + j] * z[M::NX + j] * z[M::NX + j];
  out.f = f;
  out.pole = pole;
}

// rho[r]   : coefficient of g_r g_r' for geometric row r (barrier lambda/s), CA rank-1 is added inside
// kappa[r] : coefficient of the row's own Hessian (-lambda_r); CA adds -N w / h^2.  Used only if EXACT.
template <class M, class T>
RMPC_HD void stage_derivatives(const T* z, const T* p, const T* rho, const T* kappa, const bool EXACT, StageDer<M, T>& out) {
  Kin<M, T> kn;
  kin_points<M, T, true>(z, kn);
  RMPC_UNROLL
  for (int i = 0; i < M::NQ * M::NQ; ++i) out.Hqq[i] = T(0);
  RMPC_UNROLL
  for (int i = 0; i < M::NVAR; ++i) { out.g[i] = T(0); out.Hd[i] = T(0); }
  T wv[M::NPT > 0 ? M::NPT : 1][3];   // per-point weight of the FK second derivative (EXACT)
  T cJJ[M::NPT > 0 ? M::NPT : 1];     // per-point coefficient of Jp'Jp (EXACT)
  RMPC_UNROLL
  for (int i = 0; i < M::NPT; ++i) { wv[i][0] = wv[i][1] = wv[i][2] = T(0); cJJ[i] = T(0); }
  T nvec[M::NGEO > 0 ? M::NGEO : 1][3], dist[M::NGEO > 0 ? M::NGEO : 1];
  T r1[M::NGEO > 0 ? M::NGEO : 1], kap[M::NGEO > 0 ? M::NGEO : 1];
  // geometric rows: value, gradient
  RMPC_UNROLL
  for (int r = 0; r < M::NGEO; ++r) {
    out.h[M::geo_row(r)] = geo_value<M, T>(kn, p, r, nvec[r], &dist[r]);
    const int A = M::geo_a(r);
    RMPC_UNROLL
    for (int k = 0; k < M::NQ; ++k) {
      T gk = nvec[r][0] * kn.Jp[A][0][k] + nvec[r][1] * kn.Jp[A][1][k] + nvec[r][2] * kn.Jp[A][2][k];
      if (M::geo_type(r) == GEO_PAIR) {
        const int Bp = M::geo_b(r);
        gk -= nvec[r][0] * kn.Jp[Bp][0][k] + nvec[r][1] * kn.Jp[Bp][1][k] + nvec[r][2] * kn.Jp[Bp][2][k];
      }
      out.Gg[r][k] = gk;
    }
    r1[r] = rho[r];
    kap[r] = EXACT ? kappa[r] : T(0);
  }
  RMPC_UNROLL
  for (int r = 0; r < M::NLIM; ++r) {
    const T zv = z[M::lim_var(r)], pv = p[M::lim_p(r)];
    out.h[M::lim_row(r)] = M::lim_sign(r) > 0 ? zv - pv : pv - zv;
  }
  T f = T(0);
  // GoalReaching
  if (M::GOAL_PT >= 0) {
    constexpr int E = M::GOAL_PT >= 0 ? M::GOAL_PT : 0;
    T we[3];
    RMPC_UNROLL
    for (int c = 0; c < 3; ++c) {
      const T e = kn.p[E][c] - p[M::P_GOAL + c];
      const T w = p[M::P_WGOAL + c];
      f += w * e * e;
      we[c] = T(2) * w * e;
      if (EXACT) wv[E][c] += we[c];
    }
    RMPC_UNROLL
    for (int k = 0; k < M::NQ; ++k) {
      out.g[k] += we[0] * kn.Jp[E][0][k] + we[1] * kn.Jp[E][1][k] + we[2] * kn.Jp[E][2][k];
      RMPC_UNROLL
      for (int l = 0; l <= k; ++l) {
        T v = T(0);
        RMPC_UNROLL
        for (int c = 0; c < 3; ++c) v += T(2) * p[M::P_WGOAL + c] * kn.Jp[E][c][k] * kn.Jp[E][c][l];
        out.Hqq[k * M::NQ + l] += v;
        if (l != k) out.Hqq[l * M::NQ + k] += v;
      }
    }
  }
  // ConstraintAvoidance: N * w / h on the first row of each module
  bool pole = false;
  RMPC_UNROLL
  for (int i = 0; i < M::NCA; ++i) {
    const T w = p[M::ca_w(i)];
    if (w == T(0)) continue;
    const int row = M::ca_row(i);
    const T hv = out.h[row];
    if (hv <= T(0)) pole = true;
    const T ih = trcp(hv);
    const T c0 = T(M::N) * w * ih;   // value
    const T c1 = -c0 * ih;           // d/dh
    const T c2 = T(2) * c0 * ih * ih;  // d2/dh2
    f += c0;
    if (M::ca_geo(i) >= 0) {
      const int r = M::ca_geo(i) >= 0 ? M::ca_geo(i) : 0;
      RMPC_UNROLL
      for (int k = 0; k < M::NQ; ++k) out.g[k] += c1 * out.Gg[r][k];
      r1[r] += c2;
      if (EXACT) kap[r] += c1;
    } else {
      const int l = M::ca_lim(i) >= 0 ? M::ca_lim(i) : 0;
      out.g[M::lim_var(l)] += c1 * T(M::lim_sign(l));
      out.Hd[M::lim_var(l)] += c2;
    }
  }
  // rank-1 and curvature of the geometric rows
  RMPC_UNROLL
  for (int r = 0; r < M::NGEO; ++r) {
    T c = r1[r];
    if (EXACT && M::geo_type(r) != GEO_PLANE) {
      const T kd = kap[r] * trcp(tmax(dist[r], T(1e-12)));
      c -= kd;  // kappa (I - n n')/dist  ->  -kappa/dist g g'  + kappa/dist J'J
      if (M::geo_type(r) == GEO_SPHERE) {
        cJJ[M::geo_a(r)] += kd;
      } else {
        const int A = M::geo_a(r), Bp = M::geo_b(r);
        RMPC_UNROLL
        for (int k = 0; k < M::NQ; ++k)
          RMPC_UNROLL
          for (int l = 0; l <= k; ++l) {
            T v = T(0);
            RMPC_UNROLL
            for (int cc = 0; cc < 3; ++cc) v += (kn.Jp[A][cc][k] - kn.Jp[Bp][cc][k]) * (kn.Jp[A][cc][l] - kn.Jp[Bp][cc][l]);
            v *= kd;
            out.Hqq[k * M::NQ + l] += v;
            if (l != k) out.Hqq[l * M::NQ + k] += v;
          }
      }
    }
    if (EXACT) {
      RMPC_UNROLL
      for (int cc = 0; cc < 3; ++cc) {
        wv[M::geo_a(r)][cc] += kap[r] * nvec[r][cc];
        if (M::geo_type(r) == GEO_PAIR) wv[M::geo_b(r) >= 0 ? M::geo_b(r) : 0][cc] -= kap[r] * nvec[r][cc];
      }
    }
    RMPC_UNROLL
    for (int k = 0; k < M::NQ; ++k)
      RMPC_UNROLL
      for (int l = 0; l <= k; ++l) {
        const T v = c * out.Gg[r][k] * out.Gg[r][l];
        out.Hqq[k * M::NQ + l] += v;
        if (l != k) out.Hqq[l * M::NQ + k] += v;
      }
  }
  if (EXACT) {
    RMPC_UNROLL
    for (int i = 0; i < M::NPT; ++i) {
      if (M::pt_mask(i) == 0u) continue;
      RMPC_UNROLL
      for (int k = 0; k < M::NQ; ++k)
        RMPC_UNROLL
        for (int l = 0; l <= k; ++l) {
          T v = T(0);
          RMPC_UNROLL
          for (int cc = 0; cc < 3; ++cc) v += kn.Jp[i][cc][k] * kn.Jp[i][cc][l];
          v *= cJJ[i];
          out.Hqq[k * M::NQ + l] += v;
          if (l != k) out.Hqq[l * M::NQ + k] += v;
        }
      kin_point_curvature<M, T>(kn, i, wv[i], out.Hqq);
    }
  }
  // u' diag(wu) u
  RMPC_UNROLL
  for (int j = 0; j < M::NU; ++j) {
    const T w = p[M::P_WU + j], u = z[M::NX + j];
    f += w * u * u;
    out.g[M::NX + j] += T(2) * w * u;
    out.Hd[M::NX + j] += T(2) * w;
  }
  out.f = f;
  out.pole = pole;
}

// ------------------------------------------------------------------------------------------------
// Discrete dynamics x+ = Phi(x,u): ERK2 (explicit midpoint), Ts = dt, 5 nodes.
//   holonomic double integrator: exact closed form (any 2-stage order-2 ERK gives it)
//   diffdrive: closed form of the 5 midpoint substeps (angles/speeds are polynomial in the substep index)
// ------------------------------------------------------------------------------------------------
template <class M, class T>
RMPC_HD void dyn_value(const T* z, T* xn) {
  const T dt = T(M::DT);
  if (M::BASE == 0) {
    RMPC_UNROLL
    for (int i = 0; i < M::NQ; ++i) {
      xn[i] = z[i] + dt * z[M::NQ + i] + T(0.5) * dt * dt * z[M::NX + i];
      xn[M::NQ + i] = z[M::NQ + i] + dt * z[M::NX + i];
    }
  } else {
    const T h = dt / T(5);
    const T th = z[2], v = z[6], om = z[7], u0 = z[8], u1 = z[9];
    T dx = T(0), dy = T(0);
    RMPC_UNROLL
    for (int j = 0; j < 5; ++j) {
      const T a = (T(j) + T(0.5)) * h, b = T(0.5) * T(j) * (T(j) + T(1)) * h * h;
      T s, c;
      tsincos(th + a * om + b * u1, &s, &c);
      const T w = v + a * u0;
      dx += h * c * w;
      dy += h * s * w;
    }
    xn[0] = z[0] + dx; xn[1] = z[1] + dy; xn[2] = th + dt * om + T(0.5) * dt * dt * u1;
    xn[3] = z[3]; xn[4] = z[4]; xn[5] = z[5];
    xn[6] = v + dt * u0; xn[7] = om + dt * u1;
  }
}

// Sparse structure of the Jacobian [A B] (NX x NVAR), column by column, resolved at compile time when the column
// index is a constant: column j holds n(j) entries (row, value); a value is either a model constant or one of the
// NVR*NVC z-dependent entries `ABv` (diff-drive: rows 0,1 = d(X+,Y+)/d(th, om, u1, v, u0), i.e. variables
// VC = {2, 7, 9, 6, 8}, ABv[r*5 + c]).
template <class M>
struct DynS {
  static constexpr int NVR = (M::BASE == 0) ? 0 : 2;
  static constexpr int NVC = (M::BASE == 0) ? 0 : 5;
  static constexpr int NABV = NVR * NVC > 0 ? NVR * NVC : 1;
  static constexpr RMPC_HD int vc(int c) { return c == 0 ? 2 : c == 1 ? 7 : c == 2 ? 9 : c == 3 ? 6 : 8; }
  // tables as arithmetic on packed literals (a constexpr local array would be rebuilt on the stack at run time)
  static constexpr RMPC_HD unsigned sel10(int j, unsigned v0, unsigned v1, unsigned v2, unsigned v3, unsigned v4, unsigned v5,
                                          unsigned v6, unsigned v7, unsigned v8, unsigned v9) {
    return j == 0 ? v0 : j == 1 ? v1 : j == 2 ? v2 : j == 3 ? v3 : j == 4 ? v4 : j == 5 ? v5 : j == 6 ? v6 : j == 7 ? v7 : j == 8 ? v8 : v9;
  }
  static constexpr RMPC_HD int n(int j) {
    if (M::BASE == 0) return j < M::NQ ? 1 : 2;
    return int(sel10(j, 1, 1, 3, 1, 1, 1, 3, 4, 3, 4));
  }
  static constexpr RMPC_HD int row(int j, int e) {  // one nibble per entry
    if (M::BASE == 0) return j < M::NQ ? j : (j < M::NX ? (e == 0 ? j - M::NQ : j) : (e == 0 ? j - M::NX : j - M::NX + M::NQ));
    return int((sel10(j, 0x0000, 0x0001, 0x0210, 0x0003, 0x0004, 0x0005, 0x0610, 0x7210, 0x0610, 0x7210) >> (4 * e)) & 15u);
  }
  static constexpr RMPC_HD int var(int j, int e) {  // index into ABv, -1 for a constant entry (nibble 15)
    if (M::BASE == 0) return -1;
    const unsigned v = (sel10(j, 0xffff, 0xffff, 0xff50, 0xffff, 0xffff, 0xffff, 0xff83, 0xff61, 0xff94, 0xff72) >> (4 * e)) & 15u;
    return v == 15u ? -1 : int(v);
  }
  static constexpr RMPC_HD double cval(int j, int e) {
    constexpr double dt = M::DT, h2 = 0.5 * M::DT * M::DT;
    if (M::BASE == 0) return j < M::NQ ? 1.0 : (j < M::NX ? (e == 0 ? dt : 1.0) : (e == 0 ? h2 : dt));
    // code per entry: 0 -> 0, 1 -> 1, 2 -> dt, 3 -> dt^2/2
    const unsigned c = (sel10(j, 0x0001, 0x0001, 0x0100, 0x0001, 0x0001, 0x0001, 0x0100, 0x1200, 0x0200, 0x2300) >> (4 * e)) & 15u;
    return c == 0u ? 0.0 : c == 1u ? 1.0 : c == 2u ? dt : h2;
  }
  template <class T>
  static RMPC_HD T val(int j, int e, const T* ABv) { return var(j, e) < 0 ? T(cval(j, e)) : ABv[var(j, e) < 0 ? 0 : var(j, e)]; }

  // Variables the Newton system couples ("compact" indices).  The diff-drive state carries three entries (x3..x5, the
  // world-frame velocities of the planar pose, diff_drive_mpc_model.py:24-31) that are constant under the dynamics,
  // enter no cost and no row, and feed no other state: their Newton step is dx+ = dx + d, their multipliers are zero, and
  // the Riccati algebra runs without them (5 states instead of 8).  Holonomic models: every variable is coupled.
  static constexpr int NXC = (M::BASE == 0) ? M::NX : 5;
  static constexpr int NVCP = NXC + M::NU;
  static constexpr RMPC_HD int full(int c) { return M::BASE == 0 ? c : (c < 3 ? c : c + 3); }
  static constexpr RMPC_HD int comp(int f) { return M::BASE == 0 ? f : (f < 3 ? f : (f < 6 ? -1 : f - 3)); }
  static constexpr RMPC_HD int cn(int jc) { return n(full(jc)); }
  static constexpr RMPC_HD int crow(int jc, int e) { return comp(row(full(jc), e)); }
  template <class T>
  static RMPC_HD T cv(int jc, int e, const T* ABv) { return val<T>(full(jc), e, ABv); }
  static constexpr RMPC_HD bool rows_avoid_inert() {
    for (int r = 0; r < M::NLIM; ++r)
      if (comp(M::lim_var(r)) < 0) return false;
    return true;
  }
};

// x+ = Phi(z) and the z-dependent Jacobian entries ABv (see DynS).  If `exact`, Hc (5x5 over VC, diff-drive only)
// receives nu0 * d2(X+) + nu1 * d2(Y+), the multiplier-weighted second derivative of the dynamics.
template <class M, class T>
RMPC_HD void dyn_eval(const T* z, T* xn, T* ABv, const bool exact, const T nu0, const T nu1, T Hc[5][5]) {
  const T dt = T(M::DT);
  if (M::BASE == 0) {
    RMPC_UNROLL
    for (int i = 0; i < M::NQ; ++i) {
      xn[i] = z[i] + dt * z[M::NQ + i] + T(0.5) * dt * dt * z[M::NX + i];
      xn[M::NQ + i] = z[M::NQ + i] + dt * z[M::NX + i];
    }
  } else {
    const T h = dt / T(5);
    const T th = z[2], v = z[6], om = z[7], u0 = z[8], u1 = z[9];
    T dx = T(0), dy = T(0);
    T gx[5] = {0, 0, 0, 0, 0}, gy[5] = {0, 0, 0, 0, 0};
    if (exact) {
      RMPC_UNROLL
      for (int i = 0; i < 5; ++i)
        RMPC_UNROLL
        for (int j = 0; j < 5; ++j) Hc[i][j] = T(0);
    }
    RMPC_UNROLL
    for (int j = 0; j < 5; ++j) {
      const T a = (T(j) + T(0.5)) * h, b = T(0.5) * T(j) * (T(j) + T(1)) * h * h;
      T s, c;
      tsincos(th + a * om + b * u1, &s, &c);
      const T w = v + a * u0;
      dx += h * c * w;
      dy += h * s * w;
      const T gp[3] = {T(1), a, b};  // d phi / d(th,om,u1)
      const T gw[2] = {T(1), a};     // d w / d(v,u0)
      RMPC_UNROLL
      for (int i = 0; i < 3; ++i) { gx[i] += -h * s * w * gp[i]; gy[i] += h * c * w * gp[i]; }
      RMPC_UNROLL
      for (int i = 0; i < 2; ++i) { gx[3 + i] += h * c * gw[i]; gy[3 + i] += h * s * gw[i]; }
      if (exact) {
        const T cpp = h * w * (-nu0 * c - nu1 * s);  // d2/dphi2
        const T cpw = h * (-nu0 * s + nu1 * c);      // d2/dphi dw
        RMPC_UNROLL
        for (int i = 0; i < 3; ++i) {
          RMPC_UNROLL
          for (int l = 0; l < 3; ++l) Hc[i][l] += cpp * gp[i] * gp[l];
          RMPC_UNROLL
          for (int l = 0; l < 2; ++l) { Hc[i][3 + l] += cpw * gp[i] * gw[l]; Hc[3 + l][i] += cpw * gp[i] * gw[l]; }
        }
      }
    }
    xn[0] = z[0] + dx; xn[1] = z[1] + dy; xn[2] = th + dt * om + T(0.5) * dt * dt * u1;
    xn[3] = z[3]; xn[4] = z[4]; xn[5] = z[5];
    xn[6] = v + dt * u0; xn[7] = om + dt * u1;
    RMPC_UNROLL
    for (int i = 0; i < 5; ++i) { ABv[i] = gx[i]; ABv[5 + i] = gy[i]; }
  }
}

// Dense [A B] (row-major NX x NVAR) from the sparse description — tests and debugging only.
template <class M, class T>
RMPC_HD void dyn_dense(const T* ABv, T* AB) {
  for (int i = 0; i < M::NX * M::NVAR; ++i) AB[i] = T(0);
  RMPC_UNROLL
  for (int j = 0; j < M::NVAR; ++j)
    RMPC_UNROLL
    for (int e = 0; e < DynS<M>::n(j); ++e) AB[DynS<M>::row(j, e) * M::NVAR + j] = DynS<M>::template val<T>(j, e, ABv);
}

}  // namespace rmpc
