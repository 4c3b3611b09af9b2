// rmpc_fsd.cuh — batched free-space decomposition: lidar point cloud -> K half-planes per (problem, stage), written
// straight into the stage-varying `lin_constrs_i` slots of the parameter array.
//
// Restates robotmpcs/utils/free_space_decomposition.py:61-116 (FreeSpaceDecomposition.compute_constraints + asdict)
// and its caller examples/boxer_example.py:132-148,191-203, which runs it N times per control step in Python:
//   seed   p = (x, y) + Rot(theta) (lidar_offset, 0), z = lidar_z           (boxer_example.py:138-144)
//   points sorted by distance to p, those with distance < max_radius kept   (free_space_decomposition.py:85-88)
//   greedy: nearest remaining point q -> half-plane normal n = p - q, constant c = -n.q; points with n.x + c <= 0 are
//           dropped; at most K planes                                       (:89-99, HalfPlane.__init__ :11-14)
//   unused slots: the dummy plane HalfPlane(p + (20,20,0), p)               (asdict, :104-116)
// One warp per (problem, stage): each lane holds up to 4 points, "nearest remaining" is a warp arg-min (ties go to the
// lower index; numpy's argsort leaves ties unspecified), the 4K coefficients are gathered in registers and stored with
// one coalesced write.  No sort, no shared memory.
#pragma once
#include <cuda_runtime.h>

namespace rmpc {

constexpr int kFsdMaxK = 16;       // planes per stage (two output registers per lane)
constexpr int kFsdMaxPoints = 128;  // points per cloud (four per lane)

struct FsdArgs {
  int B, N, P, K;
  const double* points;   // [B][P][3] absolute positions
  const double* ref;      // seed states (x, y, theta): ref[b * ref_bs + k * ref_ss + {0,1,2}]
  long ref_bs, ref_ss;
  double max_radius, lidar_offset, lidar_z;
  double* out;            // [B][N][K][4] or null
  double* params;         // [B][N][npar] or null
  int npar;
  int idx[kFsdMaxK];      // first parameter index of plane c (lin_constrs_c)
};

__global__ void __launch_bounds__(256) fsd_kernel(const FsdArgs a) {
  const int lane = threadIdx.x & 31;
  const long w = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (w >= (long)a.B * a.N) return;
  const int b = int(w / a.N), k = int(w - (long)b * a.N);
  const double* r = a.ref + b * a.ref_bs + k * a.ref_ss;
  double sn, cs;
  sincos(r[2], &sn, &cs);
  // rot_matrix @ (offset, 0) + state[0:2]  (no fused multiply-add: keeps the arithmetic of the numpy reference)
  const double px = __dadd_rn(__dmul_rn(cs, a.lidar_offset), r[0]);
  const double py = __dadd_rn(__dmul_rn(sn, a.lidar_offset), r[1]);
  const double pz = a.lidar_z;
  const double* pts = a.points + (size_t)b * a.P * 3;
  double x[4], y[4], z[4], d[4];
  unsigned alive = 0;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const int j = lane + 32 * t;
    x[t] = y[t] = z[t] = 0.0;
    d[t] = 1e300;
    if (j < a.P) {
      x[t] = pts[3 * j]; y[t] = pts[3 * j + 1]; z[t] = pts[3 * j + 2];
      const double dx = x[t] - px, dy = y[t] - py, dz = z[t] - pz;
      d[t] = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz)));
      if (d[t] < a.max_radius) alive |= 1u << t;
    }
  }
  double o0 = 0.0, o1 = 0.0;  // this lane's coefficients: entry `lane` and entry `lane + 32` of the K x 4 block
  int c = 0;
  for (; c < a.K; ++c) {
    double best = 1e300;
    int bi = 0x7fffffff;
#pragma unroll
    for (int t = 0; t < 4; ++t)
      if (((alive >> t) & 1u) && d[t] < best) { best = d[t]; bi = lane + 32 * t; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
    }
    if (bi == 0x7fffffff) break;  // no point left inside the radius
    const double qx = pts[3 * bi], qy = pts[3 * bi + 1], qz = pts[3 * bi + 2];
    const double nx = px - qx, ny = py - qy, nz = pz - qz;
    const double cst = -__dadd_rn(__dadd_rn(__dmul_rn(nx, qx), __dmul_rn(ny, qy)), __dmul_rn(nz, qz));
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const double v = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(nx, x[t]), __dmul_rn(ny, y[t])), __dmul_rn(nz, z[t])), cst);
      if (!(v > 0.0)) alive &= ~(1u << t);
    }
    const int e = lane & 3;
    const double val = e == 0 ? nx : e == 1 ? ny : e == 2 ? nz : cst;
    if ((lane >> 2) == c) o0 = val;
    if ((lane >> 2) + 8 == c) o1 = val;
  }
  if (c < a.K) {  // dummy plane through p + (20, 20, 0) with normal (-20, -20, 0) for the unused slots
    const double qx = px + 20.0, qy = py + 20.0;
    const double nx = px - qx, ny = py - qy;
    const double cst = -__dadd_rn(__dadd_rn(__dmul_rn(nx, qx), __dmul_rn(ny, qy)), __dmul_rn(0.0, pz));
    const int e = lane & 3;
    const double val = e == 0 ? nx : e == 1 ? ny : e == 2 ? 0.0 : cst;
    if ((lane >> 2) >= c) o0 = val;
    if ((lane >> 2) + 8 >= c) o1 = val;
  }
  const int n4 = 4 * a.K;
  if (a.out) {
    double* o = a.out + (size_t)w * n4;
    if (lane < n4) o[lane] = o0;
    if (lane + 32 < n4) o[lane + 32] = o1;
  }
  if (a.params) {
    double* p = a.params + (size_t)w * a.npar;
    if (lane < n4) p[a.idx[lane >> 2] + (lane & 3)] = o0;
    if (lane + 32 < n4) p[a.idx[(lane >> 2) + 8] + (lane & 3)] = o1;
  }
}

}  // namespace rmpc
