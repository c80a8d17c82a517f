// Rigid-body mortar contact boundary quadrature on the device.
//
// Reference path replaced: mimi::integrators::MortarContact (src/mimi/integrators/mortar_contact.{hpp,cpp}):
//   Prepare (cpp:19-133)                         -> contact_set_faces   (host C++: sort/unique, adjacency; setup)
//   ComputePressure + ElementGapAndArea (148-261) -> face_gap_area_kernel, nodal_pressure_kernel
//   ElementResidual / ElementResidualAndGrad (hpp:99-134, cpp:263-295) -> face_residual_kernel<DIM, GRAD>
//   AddBoundaryResidual[AndGrad] (297-421)        -> contact_residual_gather_kernel, contact_csr_gather_kernel
//   GapNorm (423-467)                             -> gap_norm_kernel
// Every per-face and per-node value follows the reference's operation order at nthreads = 1 (compiled -fmad=false):
// residual and Jacobian contributions are bit-identical. The whole-boundary scalars (area, force, pressure integral,
// gap norm) are summed per face in the reference's order and then across faces in ascending face id by one thread:
// deterministic, but not the reference's single running sum (tolerance 1e-12 relative, stated in the tests).
#include <algorithm>
#include <cmath>
#include <vector>

#include "mb2_ctx.h"

namespace mb2 {

struct RigidBody {
  int kind = -1;
  double p[16] = {0};
  // MB2_RIGID_BEZIER: degrees, sampling resolution of the initial guess, control net [deg_v+1][deg_u+1][dim] (device)
  int deg_u = 0, deg_v = 0, n_seed = 0;
  const double* cp = nullptr;
  int n_patches = 1;  // B-spline bodies: their Bezier patches one after the other in cp
  int rational = 0;   // control points are homogeneous {w x, w}: dim + 1 doubles each
};
constexpr int BEZ_MAXDEG = 5;

struct ContactState {
  int64_t NF = 0;
  int ndf = 0, nqf = 0, tdf = 0;
  int64_t n_marked = 0;            // marked nodes
  std::vector<int32_t> marked_nodes;  // sorted unique node ids
  int32_t* d_fvdofs = nullptr;     // NF x tdf
  int32_t* d_flocal = nullptr;     // NF x ndf  reduced nodal index (MortarContact::local_marked_dofs_)
  double* d_fN = nullptr;          // NF x nqf x ndf
  double* d_fdN = nullptr;         // NF x nqf x (ndf x (dim-1)) column-major
  double* d_fw = nullptr;          // NF x nqf
  uint16_t* d_frank = nullptr;     // NF x ndf x ndf: rank of node b in node a's neighbour list (CSR position)
  int32_t* d_mnode = nullptr;      // n_marked node ids
  int64_t* d_mn_ptr = nullptr;     // n_marked+1: marked node -> (face, local index), ascending face id
  int32_t* d_mn_face = nullptr;
  uint8_t* d_mn_loc = nullptr;
  double* d_la = nullptr;          // NF x ndf local_area_
  double* d_lg = nullptr;          // NF x ndf local_gap_
  double* d_fsum = nullptr;        // NF x 8 per-face scalars: area, pressure integral, force[3], gap^2
  double* d_gap = nullptr;         // n_marked average_gap_
  double* d_area = nullptr;        // n_marked area_
  double* d_press = nullptr;       // n_marked average_pressure_
  uint8_t* d_active = nullptr;     // NF: !IsPressureZero
  double* d_fR = nullptr;          // NF x tdf
  double* d_fK = nullptr;          // NF x tdf x tdf (column-major, column = perturbed dof)
  double* d_scal = nullptr;        // 8: area, pressure integral, force[3], gap norm
  double* d_cp = nullptr;          // Bezier control net
  RigidBody body;
  double penalty = 1e4;            // NearestDistanceBase::coefficient_
  std::vector<std::pair<int64_t, int64_t>> runs;  // CSR value ranges of the marked nodes' block rows (merged)
  double h_scal[8] = {0};
};

// ---- Bezier rigid bodies (curve in 2D, tensor-product surface in 3D): PARITY UNPINNED against the reference, whose
// NearestDistanceToSplines hands the search to splinepy (k-d tree of sampled points, Newton with bound clipping, SLSQP
// fallback). Restated scheme: sampled initial guess + bound-clipped Newton on |S(xi) - q|^2 / 2 with the exact Hessian
// (Gauss-Newton when it is not positive definite), in the operation order of oracle/mimi_oracle_contact.c
// (orc_bezier_nearest), which the tests compare bit for bit.
__device__ __forceinline__ void bernstein(int p, double t, double* B, double* dB, double* ddB) {
  double lvl[BEZ_MAXDEG + 1][BEZ_MAXDEG + 1];
  const double s = 1.0 - t;
  lvl[0][0] = 1.0;
  for (int k = 1; k <= p; k++)
    for (int i = 0; i <= k; i++) {
      const double a = (i < k) ? s * lvl[k - 1][i] : 0.0;
      const double b = (i > 0) ? t * lvl[k - 1][i - 1] : 0.0;
      lvl[k][i] = a + b;
    }
  for (int i = 0; i <= p; i++) {
    B[i] = lvl[p][i];
    double d1 = 0.0, d2 = 0.0;
    if (p >= 1) {
      const double lo = (i > 0) ? lvl[p - 1][i - 1] : 0.0;
      const double hi = (i < p) ? lvl[p - 1][i] : 0.0;
      d1 = (double)p * (lo - hi);
    }
    if (p >= 2) {
      const double a = (i > 1) ? lvl[p - 2][i - 2] : 0.0;
      const double b = (i > 0 && i < p) ? lvl[p - 2][i - 1] : 0.0;
      const double c = (i < p - 1) ? lvl[p - 2][i] : 0.0;
      d2 = (double)(p * (p - 1)) * (a - 2.0 * b + c);
    }
    dB[i] = d1;
    ddB[i] = d2;
  }
}

// S, S_u, S_v, S_uu, S_uv, S_vv at (u, v); rational patches: numerator and denominator sums, then the quotient rule in
// the order of the oracle's bezier_eval
template <int DIM>
__device__ __noinline__ void bezier_eval(const RigidBody& b, double u, double v, double (*out)[3]) {
  double Bu[BEZ_MAXDEG + 1], dBu[BEZ_MAXDEG + 1], ddBu[BEZ_MAXDEG + 1];
  double Bv[BEZ_MAXDEG + 1], dBv[BEZ_MAXDEG + 1], ddBv[BEZ_MAXDEG + 1];
  const int stride = DIM + (b.rational ? 1 : 0);
  double W[6] = {0, 0, 0, 0, 0, 0};
  bernstein(b.deg_u, u, Bu, dBu, ddBu);
  bernstein(b.deg_v, v, Bv, dBv, ddBv);
  for (int k = 0; k < 6; k++)
    for (int d = 0; d < 3; d++) out[k][d] = 0.0;
  for (int j = 0; j <= b.deg_v; j++)
    for (int i = 0; i <= b.deg_u; i++) {
      const double* P = b.cp + ((size_t)j * (b.deg_u + 1) + i) * stride;
      const double w[6] = {Bu[i] * Bv[j], dBu[i] * Bv[j], Bu[i] * dBv[j], ddBu[i] * Bv[j], dBu[i] * dBv[j], Bu[i] * ddBv[j]};
      for (int k = 0; k < 6; k++)
#pragma unroll
        for (int d = 0; d < DIM; d++) out[k][d] += w[k] * P[d];
      if (b.rational)
        for (int k = 0; k < 6; k++) W[k] += w[k] * P[DIM];
    }
  if (b.rational) {
#pragma unroll
    for (int d = 0; d < DIM; d++) {
      const double S = out[0][d] / W[0];
      const double Su = (out[1][d] - W[1] * S) / W[0];
      const double Sv = (out[2][d] - W[2] * S) / W[0];
      const double Suu = ((out[3][d] - (2.0 * W[1]) * Su) - W[3] * S) / W[0];
      const double Suv = (((out[4][d] - W[1] * Sv) - W[2] * Su) - W[4] * S) / W[0];
      const double Svv = ((out[5][d] - (2.0 * W[2]) * Sv) - W[5] * S) / W[0];
      out[0][d] = S; out[1][d] = Su; out[2][d] = Sv; out[3][d] = Suu; out[4][d] = Suv; out[5][d] = Svv;
    }
  }
}

template <int DIM>
__device__ __noinline__ void bezier_patch_nearest(const RigidBody& b, const double* q, double* pmq, double& distance,
                                                  double* fd) {
  constexpr int PD = DIM - 1;
  const int ns = b.n_seed;
  double E[6][3];
  double best = 1e300, u = 0.0, v = 0.0;
  for (int jv = 0; jv <= (PD == 2 ? ns : 0); jv++)
    for (int iu = 0; iu <= ns; iu++) {
      const double su = (double)iu / (double)ns, sv = (PD == 2) ? (double)jv / (double)ns : 0.0;
      bezier_eval<DIM>(b, su, sv, E);
      double d2 = 0.0;
#pragma unroll
      for (int d = 0; d < DIM; d++) d2 += (E[0][d] - q[d]) * (E[0][d] - q[d]);
      if (d2 < best) {
        best = d2;
        u = su;
        v = sv;
      }
    }
  for (int it = 0; it < 40; it++) {
    bezier_eval<DIM>(b, u, v, E);
    double r[3] = {0, 0, 0};
#pragma unroll
    for (int d = 0; d < DIM; d++) r[d] = E[0][d] - q[d];
    double g0 = 0.0, g1 = 0.0, a00 = 0.0, a01 = 0.0, a11 = 0.0, h00 = 0.0, h01 = 0.0, h11 = 0.0;
#pragma unroll
    for (int d = 0; d < DIM; d++) {
      g0 += r[d] * E[1][d];
      g1 += r[d] * E[2][d];
      a00 += E[1][d] * E[1][d];
      a01 += E[1][d] * E[2][d];
      a11 += E[2][d] * E[2][d];
      h00 += r[d] * E[3][d];
      h01 += r[d] * E[4][d];
      h11 += r[d] * E[5][d];
    }
    double du, dv = 0.0;
    if (PD == 1) {
      double H = a00 + h00;
      if (!(H > 0.0)) H = a00;
      du = -g0 / H;
    } else {
      double H00 = a00 + h00, H01 = a01 + h01, H11 = a11 + h11;
      double det = H00 * H11 - H01 * H01;
      if (!(H00 > 0.0 && det > 0.0)) {
        H00 = a00; H01 = a01; H11 = a11;
        det = H00 * H11 - H01 * H01;
      }
      du = (-g0 * H11 + g1 * H01) / det;
      dv = (-g1 * H00 + g0 * H01) / det;
    }
    double un = u + du, vn = v + dv;
    if (un < 0.0) un = 0.0;
    if (un > 1.0) un = 1.0;
    if (vn < 0.0) vn = 0.0;
    if (vn > 1.0) vn = 1.0;
    const double mu = fabs(un - u), mv = fabs(vn - v);
    u = un;
    v = vn;
    if (mu <= 1e-16 && mv <= 1e-16) break;
  }
  bezier_eval<DIM>(b, u, v, E);
  double sum = 0.0;
#pragma unroll
  for (int d = 0; d < DIM; d++) {
    pmq[d] = E[0][d] - q[d];
    sum += pmq[d] * pmq[d];
  }
  distance = sqrt(sum);
#pragma unroll
  for (int k = 0; k < PD; k++)
#pragma unroll
    for (int d = 0; d < DIM; d++) fd[k * DIM + d] = E[1 + k][d];
}

// Multi-patch bodies (B-splines with interior knots, split into Bezier patches on the host): the closest point of every
// patch, the nearest wins, the first on ties (oracle: rigid_bezier_nearest in mimi_oracle_contact.c).
template <int DIM>
__device__ __noinline__ void rigid_nearest_bezier(const RigidBody& b, const double* q, double* pmq, double& distance,
                                                  double* fd) {
  if (b.n_patches <= 1) {
    bezier_patch_nearest<DIM>(b, q, pmq, distance, fd);
    return;
  }
  const size_t stride = (size_t)(b.deg_u + 1) * (b.deg_v + 1) * (DIM + (b.rational ? 1 : 0));
  double best = 1e300;
  for (int p = 0; p < b.n_patches; p++) {
    RigidBody bp = b;
    bp.cp = b.cp + p * stride;
    double pm[3], dist, f[6];
    bezier_patch_nearest<DIM>(bp, q, pm, dist, f);
    if (dist < best) {
      best = dist;
#pragma unroll
      for (int d = 0; d < DIM; d++) pmq[d] = pm[d];
#pragma unroll
      for (int k = 0; k < (DIM - 1) * DIM; k++) fd[k] = f[k];
    }
  }
  distance = best;
}

// ---- rigid bodies: closed-form closest point, the operation order of oracle/ref_harness.cpp (RigidPlane,
// RigidSphere). Fills what MortarContact reads from NearestDistanceBase::Results (nearest_distance.hpp:139-193).
template <int DIM>
__device__ __forceinline__ void rigid_nearest(const RigidBody& b, const double* q, double* pmq, double& distance,
                                              double* fd) {
  if (b.kind == MB2_RIGID_BEZIER) {
    rigid_nearest_bezier<DIM>(b, q, pmq, distance, fd);
    return;
  }
  if (b.kind == MB2_RIGID_PLANE) {
    const double* o = b.p;
    const double lo = b.p[9], hi = b.p[10];
    double diff[DIM], s[2] = {0, 0};
#pragma unroll
    for (int d = 0; d < DIM; d++) diff[d] = q[d] - o[d];
#pragma unroll
    for (int k = 0; k < DIM - 1; k++) {
      const double* t = b.p + 3 + 3 * k;
      double num = 0.0, den = 0.0;
#pragma unroll
      for (int d = 0; d < DIM; d++) num += diff[d] * t[d];
#pragma unroll
      for (int d = 0; d < DIM; d++) den += t[d] * t[d];
      double v = num / den;
      if (v < lo) v = lo;
      if (v > hi) v = hi;
      s[k] = v;
    }
    double sum = 0.0;
#pragma unroll
    for (int d = 0; d < DIM; d++) {
      double x = o[d] + s[0] * b.p[3 + d];
      if (DIM == 3) x = x + s[1] * b.p[6 + d];
      pmq[d] = x - q[d];
      sum += pmq[d] * pmq[d];
    }
    distance = sqrt(sum);
#pragma unroll
    for (int k = 0; k < DIM - 1; k++)
#pragma unroll
      for (int d = 0; d < DIM; d++) fd[k * DIM + d] = b.p[3 + 3 * k + d];
    return;
  }
  const double* c = b.p;
  const double R = b.p[3];
  double diff[DIM], e[DIM];
  double n2 = 0.0;
#pragma unroll
  for (int d = 0; d < DIM; d++) {
    diff[d] = q[d] - c[d];
    n2 += diff[d] * diff[d];
  }
  const double dn = sqrt(n2);
  double sum = 0.0;
#pragma unroll
  for (int d = 0; d < DIM; d++) {
    e[d] = diff[d] / dn;
    const double x = c[d] + R * e[d];
    pmq[d] = x - q[d];
    sum += pmq[d] * pmq[d];
  }
  distance = sqrt(sum);
  if (DIM == 2) {
    fd[0] = -e[1];
    fd[1] = e[0];
  } else {
    if (fabs(e[2]) < 0.9) {
      fd[0] = -e[1]; fd[1] = e[0]; fd[2] = 0.0;
    } else {
      fd[0] = 0.0; fd[1] = -e[2]; fd[2] = e[1];
    }
    fd[3] = e[1] * fd[2] - e[2] * fd[1];
    fd[4] = e[2] * fd[0] - e[0] * fd[2];
    fd[5] = e[0] * fd[1] - e[1] * fd[0];
  }
}

// Results::ComputeNormal<true> (nearest_distance.hpp:139-181) == ComputeUnitNormal (integrator_utils.hpp:218-251)
template <int DIM>
__device__ __forceinline__ void unit_normal(const double* d, double* n) {
  if (DIM == 2) {
    const double inv = 1. / sqrt(d[0] * d[0] + d[1] * d[1]);
    n[0] = d[1] * inv;
    n[1] = -d[0] * inv;
  } else {
    const double n0 = d[1] * d[5] - d[2] * d[4];
    const double n1 = d[2] * d[3] - d[0] * d[5];
    const double n2 = d[0] * d[4] - d[1] * d[3];
    const double inv = 1. / sqrt(n0 * n0 + n1 * n1 + n2 * n2);
    n[0] = n0 * inv;
    n[1] = n1 * inv;
    n[2] = n2 * inv;
  }
}

// mfem::DenseMatrix::Weight, DIM x (DIM-1) (third_party/mfem/linalg/densemat.cpp:559-584)
template <int DIM>
__device__ __forceinline__ double jac_weight(const double* d) {
  if (DIM == 2) return sqrt(d[0] * d[0] + d[1] * d[1]);
  const double E = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
  const double G = d[3] * d[3] + d[4] * d[4] + d[5] * d[5];
  const double F = d[0] * d[3] + d[1] * d[4] + d[2] * d[5];
  return sqrt(E * G - F * F);
}

// MortarContactWorkData::ComputeJ = MultAtB(element_x_mat_, dN_dxi, J) (integrator_utils.cpp:99-104): k sequential.
// x: ndf x DIM (shared), dN: ndf x (DIM-1). pm >= 0 replaces x[pm] by xp (finite-difference perturbation).
template <int DIM>
__device__ __forceinline__ void face_J(int ndf, const double* x, const double* dN, int pm, double xp, double* J) {
#pragma unroll
  for (int j = 0; j < DIM - 1; j++)
#pragma unroll
    for (int i = 0; i < DIM; i++) {
      double d = 0.0;
      for (int k = 0; k < ndf; k++) {
        const int idx = k + i * ndf;
        d += ((idx == pm) ? xp : x[idx]) * dN[k + j * ndf];
      }
      J[i + j * DIM] = d;
    }
}

// ---- pass 1: MortarContact::ElementGapAndArea (mortar_contact.cpp:148-195). One CTA per face. Threads q < nqf
// evaluate one quadrature point each (query, closest point, gap filter, |J|); threads i < ndf then sum
// local_area_[i] / local_gap_[i] over q in ascending order. GAPNORM: only g^2 [g < 0] (GapNorm, cpp:423-467).
template <int DIM, bool GAPNORM>
__global__ void face_gap_area_kernel(int64_t NF, int ndf, int nqf, const int32_t* __restrict__ fvdofs,
                                     const double* __restrict__ fN, const double* __restrict__ fdN,
                                     const double* __restrict__ fw, const double* __restrict__ xref,
                                     const double* __restrict__ u, const RigidBody body, double* __restrict__ la,
                                     double* __restrict__ lg, double* __restrict__ fsum) {
  extern __shared__ double sm[];
  const int tdf = DIM * ndf;
  double* sx = sm;             // tdf: current position x = u + x_ref (CurrentElementSolutionCopy)
  double* sfac = sx + tdf;     // nqf
  double* sfacg = sfac + nqf;  // nqf
  const int t = threadIdx.x;
  for (int64_t f = blockIdx.x; f < NF; f += gridDim.x) {
    for (int k = t; k < tdf; k += blockDim.x) {
      const int32_t v = fvdofs[f * tdf + k];
      sx[k] = u[v] + xref[v];
    }
    __syncthreads();
    if (t < nqf) {
      const double* Nq = fN + (f * nqf + t) * ndf;
      double query[DIM], pmq[DIM], fd[2 * DIM], nrm[DIM], J[2 * DIM], dist;
      // ComputeNearestDistanceQuery = DenseMatrix::MultTranspose (densemat.cpp:166-179)
#pragma unroll
      for (int c = 0; c < DIM; c++) {
        double y = 0.0;
        for (int r = 0; r < ndf; r++) y += Nq[r] * sx[r + c * ndf];
        query[c] = y;
      }
      rigid_nearest<DIM>(body, query, pmq, dist, fd);
      unit_normal<DIM>(fd, nrm);
      double dot = 0.0;
#pragma unroll
      for (int c = 0; c < DIM; c++) dot += pmq[c] * nrm[c];
      const double true_g = -dot;  // Results::NormalGap
      if (GAPNORM) {
        sfac[t] = (true_g < 0.0) ? true_g * true_g : 0.0;
      } else {
        double g = (0. < true_g) ? 0. : true_g;  // std::min(true_g, 0.)
        face_J<DIM>(ndf, sx, fdN + (f * nqf + t) * ndf * (DIM - 1), -1, 0.0, J);
        const double fac = fw[f * nqf + t] * jac_weight<DIM>(J);
        const double ratio = fabs(true_g) / dist;
        const double cl = (ratio < 1.) ? ratio : 1.;  // std::min(1., ratio)
        if (acos(cl) > 1.e-5) g = 0.0;
        sfac[t] = fac;
        sfacg[t] = fac * g;
      }
    }
    __syncthreads();
    if (!GAPNORM && t < ndf) {
      double a = 0.0, g = 0.0;
      for (int q = 0; q < nqf; q++) {
        const double N = fN[(f * nqf + q) * ndf + t];
        a += sfac[q] * N;
        g += sfacg[q] * N;
      }
      la[f * ndf + t] = a;
      lg[f * ndf + t] = g;
    }
    if (t == 0) {
      double s = 0.0;
      if (GAPNORM) {
        for (int q = 0; q < nqf; q++)
          if (sfac[q] != 0.0) s += sfac[q];  // the reference only adds penetrating points
        fsum[f * 8 + 5] = s;
      } else {
        for (int q = 0; q < nqf; q++) s += sfac[q];
        fsum[f * 8 + 0] = s;
      }
    }
    __syncthreads();
  }
}

// ---- MortarContact::ComputePressure, second half (cpp:226-258): nodal area / gap = sum of the faces' local vectors
// in ascending face id (thread_local_*_ at nthreads = 1), p = gap / area * penalty.
__global__ void nodal_pressure_kernel(int64_t n_marked, int ndf, const int64_t* __restrict__ ptr,
                                      const int32_t* __restrict__ face, const uint8_t* __restrict__ loc,
                                      const double* __restrict__ la, const double* __restrict__ lg, double penalty,
                                      double* __restrict__ gap, double* __restrict__ area, double* __restrict__ press) {
  const int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (n >= n_marked) return;
  double a = 0.0, g = 0.0;
  for (int64_t p = ptr[n]; p < ptr[n + 1]; p++) {
    a += la[(int64_t)face[p] * ndf + loc[p]];
    g += lg[(int64_t)face[p] * ndf + loc[p]];
  }
  a = 0.0 + a;  // InitializeGapAreaPressure, then += thread-local
  g = 0.0 + g;
  area[n] = a;
  gap[n] = g;
  press[n] = g / a * penalty;
}

// Whole-boundary scalars: per-face partial sums reduced by ONE 256-thread CTA in a fixed order (thread t adds faces
// t, t+256, ... in ascending id, then a shared-memory tree): deterministic, run-to-run identical, 1e-12 relative to the
// reference's single running sum. mode 0: area (slot 0); mode 1: pressure integral + force (slots 1..4, active faces
// only); mode 2: gap norm (slot 5).
__global__ void __launch_bounds__(256) face_scalar_sum_kernel(int64_t NF, int mode, const double* __restrict__ fsum,
                                                              const uint8_t* __restrict__ active,
                                                              double* __restrict__ scal) {
  __shared__ double sh[4][256];
  const int t = threadIdx.x;
  const int first = (mode == 0) ? 0 : (mode == 1 ? 1 : 5), cnt = (mode == 1) ? 4 : 1;
  double s[4] = {0, 0, 0, 0};
  for (int64_t f = t; f < NF; f += 256)
    if (mode != 1 || active[f])
      for (int k = 0; k < cnt; k++) s[k] += fsum[f * 8 + first + k];
  for (int k = 0; k < cnt; k++) sh[k][t] = s[k];
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if (t < w)
      for (int k = 0; k < cnt; k++) sh[k][t] += sh[k][t + w];
    __syncthreads();
  }
  if (t < cnt) {
    const double v = sh[t][0];
    scal[first + t] = (mode == 2) ? sqrt(v) : (mode == 0 ? 0.0 + v : v);
  }
}

// ---- pass 2: MortarContact::ElementResidual<true> and ElementResidualAndGrad (mortar_contact.hpp:99-134,
// cpp:263-295). One CTA per face; faces whose nodal pressures are all zero are skipped (IsPressureZero, cpp:321-325).
//   stage A  task (m, q), m = tdf is the unperturbed evaluation: J(x + h e_m), |J|, unit normal, fac = w |J| p(q)
//            -> aw[m][q][c] = normal[c] * (-fac)                       (pressure frozen, the step is on x = u + x_ref)
//   stage B  task (m, j = i + c*ndf): R_m[j] = sum_q aw[m][q][c] * N_q[i], q ascending (Ptr_AddMult_a_VWt)
//   stage C  K[j + m*tdf] = (R_m[j] - R[j]) * (1/h_m)
template <int DIM, bool GRAD>
__global__ void face_residual_kernel(int64_t NF, int ndf, int nqf, const int32_t* __restrict__ fvdofs,
                                     const int32_t* __restrict__ flocal, const double* __restrict__ fN,
                                     const double* __restrict__ fdN, const double* __restrict__ fw,
                                     const double* __restrict__ xref, const double* __restrict__ u,
                                     const double* __restrict__ press, uint8_t* __restrict__ active,
                                     double* __restrict__ fR, double* __restrict__ fK, double* __restrict__ fsum) {
  extern __shared__ double sm[];
  const int tdf = DIM * ndf;
  const int NM = GRAD ? tdf + 1 : 1;
  double* sx = sm;               // tdf
  double* spe = sx + tdf;        // ndf  element_pressure_
  double* spq = spe + ndf;       // nqf  p(q) = N_q . pe
  double* sfac = spq + nqf;      // nqf  fac(q) of the unperturbed evaluation
  double* sR = sfac + nqf;       // tdf  base residual
  double* saw = sR + tdf;        // NM x nqf x DIM
  __shared__ int s_zero;
  const int t = threadIdx.x;
  for (int64_t f = blockIdx.x; f < NF; f += gridDim.x) {
    if (t == 0) s_zero = 1;
    __syncthreads();
    for (int k = t; k < tdf; k += blockDim.x) {
      const int32_t v = fvdofs[f * tdf + k];
      sx[k] = u[v] + xref[v];
    }
    for (int i = t; i < ndf; i += blockDim.x) {
      const double p = press[flocal[f * ndf + i]];
      spe[i] = p;
      if (p != 0.0) s_zero = 0;
    }
    __syncthreads();
    const bool zero = s_zero != 0;
    if (t == 0) active[f] = zero ? 0 : 1;
    if (zero) {
      __syncthreads();
      continue;
    }
    for (int q = t; q < nqf; q += blockDim.x) {
      const double* Nq = fN + (f * nqf + q) * ndf;
      double p = 0.0;
      for (int i = 0; i < ndf; i++) p += Nq[i] * spe[i];  // MortarContactWorkData::Pressure
      spq[q] = p;
    }
    __syncthreads();
    // stage A
    for (int task = t; task < NM * nqf; task += blockDim.x) {
      const int m = GRAD ? task / nqf : 0, q = task - m * nqf;
      const bool base = !GRAD || m == tdf;
      int pm = -1;
      double xp = 0.0;
      if (!base) {
        const double orig = sx[m];
        const double h = (orig != 0.0) ? fabs(orig) * 1.0e-8 : 1.0e-10;
        pm = m;
        xp = orig + h;
      }
      double J[2 * DIM], nrm[DIM];
      face_J<DIM>(ndf, sx, fdN + (f * nqf + q) * ndf * (DIM - 1), pm, xp, J);
      const double det_J = jac_weight<DIM>(J);
      unit_normal<DIM>(J, nrm);
      const double fac = fw[f * nqf + q] * det_J * spq[q];
#pragma unroll
      for (int c = 0; c < DIM; c++) saw[(m * nqf + q) * DIM + c] = nrm[c] * (-fac);
      if (base) sfac[q] = fac;  // record == true: pressure_integral += fac, force.Add(fac, normal) -- summed below
    }
    __syncthreads();
    // per-face scalars, q ascending from 0.0 (tl_force, local_pressure of one face)
    if (t < DIM + 1) {
      const int mb = GRAD ? tdf : 0;
      double s = 0.0;
      if (t == 0) {
        for (int q = 0; q < nqf; q++) s += sfac[q];
      } else {
        // fac * normal[c] = -(normal[c] * (-fac)): exact negation
        for (int q = 0; q < nqf; q++) s += -saw[(mb * nqf + q) * DIM + (t - 1)];
      }
      fsum[f * 8 + 1 + t] = s;
    }
    // stage B (base) -> sR
    {
      const int mb = GRAD ? tdf : 0;
      for (int j = t; j < tdf; j += blockDim.x) {
        const int i = j % ndf, c = j / ndf;
        double r = 0.0;
        for (int q = 0; q < nqf; q++) r += saw[(mb * nqf + q) * DIM + c] * fN[(f * nqf + q) * ndf + i];
        sR[j] = r;
        fR[f * tdf + j] = r;
      }
    }
    __syncthreads();
    if (GRAD) {
      for (int task = t; task < tdf * tdf; task += blockDim.x) {
        const int m = task / tdf, j = task - m * tdf;
        const int i = j % ndf, c = j / ndf;
        double r = 0.0;
        for (int q = 0; q < nqf; q++) r += saw[(m * nqf + q) * DIM + c] * fN[(f * nqf + q) * ndf + i];
        const double orig = sx[m];
        const double h = (orig != 0.0) ? fabs(orig) * 1.0e-8 : 1.0e-10;
        fK[f * (int64_t)tdf * tdf + task] = (r - sR[j]) * (1. / h);
      }
    }
    __syncthreads();
  }
}

// ---- scatter, written as ordered gathers (faces ascending, straight onto the incoming values:
// residual.AddElementVector / A[A_ids[k]] += K[k] * grad_factor, mortar_contact.cpp:336-340, 392-401).
__global__ void contact_residual_gather_kernel(int dim, int64_t n_marked, int ndf, const int32_t* __restrict__ mnode,
                                               const int64_t* __restrict__ ptr, const int32_t* __restrict__ face,
                                               const uint8_t* __restrict__ loc, const uint8_t* __restrict__ active,
                                               const double* __restrict__ fR, double* __restrict__ r,
                                               const uint8_t* __restrict__ owned) {
  const int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (idx >= n_marked * dim) return;
  const int64_t n = idx / dim;
  if (owned && !owned[mnode[n]]) return;  // partitioned run: the owner rank assembles this node's entries
  const int c = (int)(idx % dim);
  const int64_t v = (int64_t)dim * mnode[n] + c;
  double acc = r[v];
  const int tdf = dim * ndf;
  for (int64_t p = ptr[n]; p < ptr[n + 1]; p++)
    if (active[face[p]]) acc += fR[(int64_t)face[p] * tdf + loc[p] + c * ndf];
  r[v] = acc;
}

// One CTA per marked node a (block row of dim CSR rows). Thread = (column dof j = b + cj*ndf, row component ci).
template <int DIM>
__global__ void contact_csr_gather_kernel(int64_t n_marked, int ndf, const int32_t* __restrict__ mnode,
                                          const int64_t* __restrict__ ptr, const int32_t* __restrict__ face,
                                          const uint8_t* __restrict__ loc, const uint8_t* __restrict__ active,
                                          const uint16_t* __restrict__ frank, const int64_t* __restrict__ I,
                                          const double* __restrict__ fK, double grad_factor, double* __restrict__ A,
                                          const uint8_t* __restrict__ owned) {
  const int tdf = DIM * ndf;
  const int t = threadIdx.x;
  const int j = t / DIM, ci = t % DIM;
  const int b = j % ndf, cj = j / ndf;
  for (int64_t n = blockIdx.x; n < n_marked; n += gridDim.x) {
    if (owned && !owned[mnode[n]]) continue;  // uniform per CTA
    const int64_t row = (int64_t)DIM * mnode[n] + ci;
    for (int64_t p = ptr[n]; p < ptr[n + 1]; p++) {
      const int64_t f = face[p];
      if (active[f] && t < tdf * DIM) {
        const int a = loc[p];
        const int64_t pos = I[row] + DIM * (int64_t)frank[(f * ndf + a) * ndf + b] + cj;
        A[pos] += fK[f * (int64_t)tdf * tdf + (int64_t)j * tdf + a + ci * ndf] * grad_factor;
      }
      __syncthreads();  // the next face may touch the same entries
    }
  }
}

// rank of face node b in the neighbour list of face node a (the block pattern built by mb2_build_pattern)
__global__ void face_rank_kernel(int dim, int64_t NF, int ndf, const int32_t* __restrict__ fvdofs,
                                 const int32_t* __restrict__ nnbr, const int64_t* __restrict__ nodeptr,
                                 const int32_t* __restrict__ nbr, uint16_t* __restrict__ frank, int* __restrict__ flag,
                                 const uint8_t* __restrict__ owned) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= NF * ndf * ndf) return;
  const int b = (int)(t % ndf), a = (int)((t / ndf) % ndf);
  const int64_t f = t / ((int64_t)ndf * ndf);
  const int tdf = dim * ndf;
  const int32_t A = fvdofs[f * tdf + a] / dim, B = fvdofs[f * tdf + b] / dim;
  if (owned && !owned[A]) {  // rows of another rank: never gathered here
    frank[t] = 0;
    return;
  }
  const int32_t* list = nbr + nodeptr[A];
  int lo = 0, hi = nnbr[A] - 1, pos = -1;
  while (lo <= hi) {
    const int mid = (lo + hi) >> 1;
    if (list[mid] == B) { pos = mid; break; }
    if (list[mid] < B) lo = mid + 1; else hi = mid - 1;
  }
  if (pos < 0) { *flag = 1; pos = 0; }
  frank[t] = (uint16_t)pos;
}

__global__ void face_A_ids_kernel(int dim, int64_t NF, int ndf, const int32_t* __restrict__ fvdofs,
                                  const uint16_t* __restrict__ frank, const int64_t* __restrict__ I,
                                  int32_t* __restrict__ out) {
  const int tdf = dim * ndf;
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= NF * tdf * tdf) return;
  const int i = (int)(t % tdf), j = (int)((t / tdf) % tdf);
  const int64_t f = t / ((int64_t)tdf * tdf);
  const int a = i % ndf, ci = i / ndf, b = j % ndf, cj = j / ndf;
  out[t] = (int32_t)(I[fvdofs[f * tdf + a] / dim * (int64_t)dim + ci] + dim * frank[(f * ndf + a) * ndf + b] + cj);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static ContactState* CS(mb2_ctx* c) { return static_cast<ContactState*>(c->contact); }

void free_contact(mb2_ctx* c) {
  ContactState* S = CS(c);
  if (!S) return;
  for (void* p : {(void*)S->d_fvdofs, (void*)S->d_flocal, (void*)S->d_fN, (void*)S->d_fdN, (void*)S->d_fw, (void*)S->d_frank,
                  (void*)S->d_mnode, (void*)S->d_mn_ptr, (void*)S->d_mn_face, (void*)S->d_mn_loc, (void*)S->d_la,
                  (void*)S->d_lg, (void*)S->d_fsum, (void*)S->d_gap, (void*)S->d_area, (void*)S->d_press,
                  (void*)S->d_active, (void*)S->d_fR, (void*)S->d_fK, (void*)S->d_scal, (void*)S->d_cp})
    if (p) cudaFree(p);
  delete S;
  c->contact = nullptr;
}

template <typename T>
static int upload(mb2_ctx* c, T*& d, const T* h, size_t n) {
  MB2_CUDA(cudaMalloc(&d, sizeof(T) * (n ? n : 1)));
  if (h) MB2_CUDA(cudaMemcpyAsync(d, h, sizeof(T) * n, cudaMemcpyHostToDevice, c->stream));
  return MB2_OK;
}

// MortarContact::Prepare (mortar_contact.cpp:19-133): sorted unique marked vdofs -> reduced nodal numbering
// (local_marked_dofs_), plus what the ordered gathers need: marked node -> (face, local index) in ascending face id,
// the CSR rank of every face node pair, and the value ranges of the marked block rows.
int contact_set_faces(mb2_ctx* c, int64_t NF, int ndf, int nqf, const int32_t* v_dofs, const double* N,
                      const double* dN_dxi, const double* w) {
  free_contact(c);
  ContactState* S = new ContactState();
  c->contact = S;
  const int dim = c->dim, tdf = dim * ndf;
  S->NF = NF; S->ndf = ndf; S->nqf = nqf; S->tdf = tdf;
  // component-major byVDIM check + marked node list
  std::vector<int32_t> nodes((size_t)NF * ndf);
  for (int64_t f = 0; f < NF; f++)
    for (int k = 0; k < tdf; k++) {
      const int32_t v = v_dofs[f * tdf + k];
      if (v < 0 || v >= c->NV || v % dim != k / ndf || (k >= ndf && v / dim != v_dofs[f * tdf + k % ndf] / dim)) {
        c->err = "contact face v_dofs are not component-major byVDIM ids of this space";
        return MB2_ERR_ARG;
      }
      if (k < ndf) nodes[(size_t)f * ndf + k] = v / dim;
    }
  S->marked_nodes = nodes;
  std::sort(S->marked_nodes.begin(), S->marked_nodes.end());
  S->marked_nodes.erase(std::unique(S->marked_nodes.begin(), S->marked_nodes.end()), S->marked_nodes.end());
  S->n_marked = (int64_t)S->marked_nodes.size();
  std::vector<int32_t> local((size_t)NF * ndf);
  std::vector<int64_t> ptr((size_t)S->n_marked + 1, 0);
  for (size_t i = 0; i < nodes.size(); i++) {
    local[i] = (int32_t)(std::lower_bound(S->marked_nodes.begin(), S->marked_nodes.end(), nodes[i]) - S->marked_nodes.begin());
    ptr[(size_t)local[i] + 1]++;
  }
  for (int64_t n = 0; n < S->n_marked; n++) ptr[n + 1] += ptr[n];
  std::vector<int32_t> mface((size_t)ptr[S->n_marked]);
  std::vector<uint8_t> mloc((size_t)ptr[S->n_marked]);
  {
    std::vector<int64_t> fill(ptr.begin(), ptr.end() - 1);
    for (int64_t f = 0; f < NF; f++)  // ascending face id per node
      for (int i = 0; i < ndf; i++) {
        const int64_t at = fill[(size_t)local[(size_t)f * ndf + i]]++;
        mface[(size_t)at] = (int32_t)f;
        mloc[(size_t)at] = (uint8_t)i;
      }
  }
  int rc;
  if ((rc = upload(c, S->d_fvdofs, v_dofs, (size_t)NF * tdf))) return rc;
  if ((rc = upload(c, S->d_flocal, local.data(), local.size()))) return rc;
  if ((rc = upload(c, S->d_fN, N, (size_t)NF * nqf * ndf))) return rc;
  if ((rc = upload(c, S->d_fdN, dN_dxi, (size_t)NF * nqf * ndf * (dim - 1)))) return rc;
  if ((rc = upload(c, S->d_fw, w, (size_t)NF * nqf))) return rc;
  if ((rc = upload(c, S->d_mnode, S->marked_nodes.data(), S->marked_nodes.size()))) return rc;
  if ((rc = upload(c, S->d_mn_ptr, ptr.data(), ptr.size()))) return rc;
  if ((rc = upload(c, S->d_mn_face, mface.data(), mface.size()))) return rc;
  if ((rc = upload(c, S->d_mn_loc, mloc.data(), mloc.size()))) return rc;
  if ((rc = upload<double>(c, S->d_la, nullptr, (size_t)NF * ndf))) return rc;
  if ((rc = upload<double>(c, S->d_lg, nullptr, (size_t)NF * ndf))) return rc;
  if ((rc = upload<double>(c, S->d_fsum, nullptr, (size_t)NF * 8))) return rc;
  if ((rc = upload<double>(c, S->d_gap, nullptr, (size_t)S->n_marked))) return rc;
  if ((rc = upload<double>(c, S->d_area, nullptr, (size_t)S->n_marked))) return rc;
  if ((rc = upload<double>(c, S->d_press, nullptr, (size_t)S->n_marked))) return rc;
  if ((rc = upload<uint8_t>(c, S->d_active, nullptr, (size_t)NF))) return rc;
  if ((rc = upload<double>(c, S->d_fR, nullptr, (size_t)NF * tdf))) return rc;
  if ((rc = upload<double>(c, S->d_fK, nullptr, (size_t)NF * tdf * tdf))) return rc;
  if ((rc = upload<double>(c, S->d_scal, nullptr, 8))) return rc;
  if ((rc = upload<uint16_t>(c, S->d_frank, nullptr, (size_t)NF * ndf * ndf))) return rc;
  MB2_CUDA(cudaMemsetAsync(S->d_fsum, 0, sizeof(double) * NF * 8, c->stream));
  MB2_CUDA(cudaMemsetAsync(S->d_scal, 0, sizeof(double) * 8, c->stream));
  MB2_CUDA(cudaMemsetAsync(S->d_active, 0, NF, c->stream));
  // CSR ranks and the marked block rows' value ranges
  MB2_CUDA(cudaMemsetAsync(c->d_flag, 0, sizeof(int), c->stream));
  const int64_t nr = NF * ndf * ndf;
  face_rank_kernel<<<(unsigned)((nr + 255) / 256), 256, 0, c->stream>>>(dim, NF, ndf, S->d_fvdofs, c->d_nnbr, c->d_nodeptr,
                                                                       c->d_nbr, S->d_frank, c->d_flag, c->d_owned);
  MB2_CUDA(cudaGetLastError());
  int flag = 0;
  MB2_CUDA(cudaMemcpyAsync(&flag, c->d_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  std::vector<int64_t> I((size_t)c->NV + 1);
  MB2_CUDA(cudaMemcpyAsync(I.data(), c->d_I, sizeof(int64_t) * (c->NV + 1), cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  if (flag) {
    c->err = "a contact face couples two nodes that share no domain element (pattern mismatch)";
    return MB2_ERR_ARG;
  }
  std::vector<uint8_t> owned;
  if (c->d_owned) {
    owned.resize((size_t)c->Nn);
    MB2_CUDA(cudaMemcpy(owned.data(), c->d_owned, (size_t)c->Nn, cudaMemcpyDeviceToHost));
  }
  for (int32_t n : S->marked_nodes) {
    if (!owned.empty() && !owned[(size_t)n]) continue;
    const int64_t off = I[(size_t)dim * n], end = I[(size_t)dim * n + dim];
    if (!S->runs.empty() && S->runs.back().first + S->runs.back().second == off) S->runs.back().second += end - off;
    else S->runs.push_back({off, end - off});
  }
  return MB2_OK;
}

template <int DIM>
static int run_pressure(mb2_ctx* c, ContactState* S) {
  const int threads = (std::max(S->nqf, S->ndf) + 31) / 32 * 32;
  const size_t smem = sizeof(double) * (S->tdf + 2 * S->nqf);
  const unsigned grid = (unsigned)std::min<int64_t>(S->NF, (int64_t)c->sm_count * 16);
  face_gap_area_kernel<DIM, false><<<grid, threads, smem, c->stream>>>(S->NF, S->ndf, S->nqf, S->d_fvdofs, S->d_fN, S->d_fdN,
                                                                       S->d_fw, c->d_xref, c->d_u, S->body, S->d_la,
                                                                       S->d_lg, S->d_fsum);
  MB2_CUDA(cudaGetLastError());
  nodal_pressure_kernel<<<(unsigned)((S->n_marked + 127) / 128), 128, 0, c->stream>>>(
      S->n_marked, S->ndf, S->d_mn_ptr, S->d_mn_face, S->d_mn_loc, S->d_la, S->d_lg, S->penalty, S->d_gap, S->d_area,
      S->d_press);
  MB2_CUDA(cudaGetLastError());
  face_scalar_sum_kernel<<<1, 256, 0, c->stream>>>(S->NF, 0, S->d_fsum, S->d_active, S->d_scal);
  MB2_CUDA(cudaGetLastError());
  c->last_launches += 3;
  return MB2_OK;
}

template <int DIM, bool GRAD>
static int run_faces(mb2_ctx* c, ContactState* S, double gf) {
  const int NM = GRAD ? S->tdf + 1 : 1;
  const size_t smem = sizeof(double) * (2 * S->tdf + S->ndf + 2 * S->nqf + (size_t)NM * S->nqf * DIM);
  auto kern = face_residual_kernel<DIM, GRAD>;
  if (smem > 48 * 1024) MB2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const unsigned grid = (unsigned)std::min<int64_t>(S->NF, (int64_t)c->sm_count * 8);
  kern<<<grid, 256, smem, c->stream>>>(S->NF, S->ndf, S->nqf, S->d_fvdofs, S->d_flocal, S->d_fN, S->d_fdN, S->d_fw,
                                       c->d_xref, c->d_u, S->d_press, S->d_active, S->d_fR, S->d_fK, S->d_fsum);
  MB2_CUDA(cudaGetLastError());
  const int64_t nv = S->n_marked * DIM;
  contact_residual_gather_kernel<<<(unsigned)((nv + 127) / 128), 128, 0, c->stream>>>(
      DIM, S->n_marked, S->ndf, S->d_mnode, S->d_mn_ptr, S->d_mn_face, S->d_mn_loc, S->d_active, S->d_fR, c->d_r, c->d_owned);
  MB2_CUDA(cudaGetLastError());
  if (GRAD) {
    const int threads = (S->tdf * DIM + 31) / 32 * 32;
    const unsigned g2 = (unsigned)std::min<int64_t>(S->n_marked, (int64_t)c->sm_count * 16);
    contact_csr_gather_kernel<DIM><<<g2, threads, 0, c->stream>>>(S->n_marked, S->ndf, S->d_mnode, S->d_mn_ptr, S->d_mn_face,
                                                                 S->d_mn_loc, S->d_active, S->d_frank, c->d_I, S->d_fK, gf,
                                                                 c->d_A, c->d_owned);
    MB2_CUDA(cudaGetLastError());
    c->last_launches += 1;
  }
  face_scalar_sum_kernel<<<1, 256, 0, c->stream>>>(S->NF, 1, S->d_fsum, S->d_active, S->d_scal);
  MB2_CUDA(cudaGetLastError());
  c->last_launches += 3;
  return MB2_OK;
}

static int contact_ready(mb2_ctx* c) {
  ContactState* S = CS(c);
  if (!S) { c->err = "call mb2_contact_set_faces first"; return MB2_ERR_ARG; }
  if (S->body.kind < 0) { c->err = "call mb2_contact_set_rigid_body first"; return MB2_ERR_ARG; }
  if (!c->d_xref) { c->err = "call mb2_set_nodes first (reference coordinates of the face dofs)"; return MB2_ERR_ARG; }
  return MB2_OK;
}

// AddBoundaryResidual / AddBoundaryResidualAndGrad on the resident u, r (and A): ComputePressure, then the faces.
int contact_assemble(mb2_ctx* c, bool grad, double gf) {
  int rc = contact_ready(c);
  if (rc) return rc;
  ContactState* S = CS(c);
  c->last_launches = 0;
  MB2_CUDA(cudaEventRecord(c->ev[0], c->stream));
  if (c->dim == 3) {
    if ((rc = run_pressure<3>(c, S))) return rc;
    rc = grad ? run_faces<3, true>(c, S, gf) : run_faces<3, false>(c, S, gf);
  } else {
    if ((rc = run_pressure<2>(c, S))) return rc;
    rc = grad ? run_faces<2, true>(c, S, gf) : run_faces<2, false>(c, S, gf);
  }
  if (rc) return rc;
  MB2_CUDA(cudaEventRecord(c->ev[1], c->stream));
  return MB2_OK;
}

bool contact_active(mb2_ctx* c) { return CS(c) && CS(c)->body.kind >= 0 && c->d_xref; }

int contact_gap_norm(mb2_ctx* c, double* out) {
  int rc = contact_ready(c);
  if (rc) return rc;
  ContactState* S = CS(c);
  const int threads = (std::max(S->nqf, S->ndf) + 31) / 32 * 32;
  const size_t smem = sizeof(double) * (S->tdf + 2 * S->nqf);
  const unsigned grid = (unsigned)std::min<int64_t>(S->NF, (int64_t)c->sm_count * 16);
  if (c->dim == 3)
    face_gap_area_kernel<3, true><<<grid, threads, smem, c->stream>>>(S->NF, S->ndf, S->nqf, S->d_fvdofs, S->d_fN, S->d_fdN,
                                                                      S->d_fw, c->d_xref, c->d_u, S->body, S->d_la, S->d_lg,
                                                                      S->d_fsum);
  else
    face_gap_area_kernel<2, true><<<grid, threads, smem, c->stream>>>(S->NF, S->ndf, S->nqf, S->d_fvdofs, S->d_fN, S->d_fdN,
                                                                      S->d_fw, c->d_xref, c->d_u, S->body, S->d_la, S->d_lg,
                                                                      S->d_fsum);
  MB2_CUDA(cudaGetLastError());
  face_scalar_sum_kernel<<<1, 256, 0, c->stream>>>(S->NF, 2, S->d_fsum, S->d_active, S->d_scal);
  MB2_CUDA(cudaGetLastError());
  MB2_CUDA(cudaMemcpyAsync(out, S->d_scal + 5, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  c->last_launches = 2;
  return MB2_OK;
}

}  // namespace mb2

// ---------------------------------------------------------------------------------------------
// C ABI (include/mimi_b200.h, "contact" section)
// ---------------------------------------------------------------------------------------------
using mb2::ContactState;
#define CCHECK(c) \
  if (!(c)) return MB2_ERR_ARG
#define CREQ(cond, msg) \
  if (!(cond)) {        \
    c->err = (msg);     \
    return MB2_ERR_ARG; \
  }

extern "C" {

int mb2_contact_set_faces(mb2_ctx* c, int64_t NF, int nd_f, int nq_f, const int32_t* v_dofs, const double* N,
                          const double* dN_dxi, const double* w) {
  CCHECK(c);
  CREQ(c->have_pattern, "call mb2_build_pattern first");
  CREQ(NF > 0 && nd_f > 0 && nd_f < 256 && nq_f > 0 && v_dofs && N && dN_dxi && w, "bad contact face arguments");
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::contact_set_faces(c, NF, nd_f, nq_f, v_dofs, N, dN_dxi, w);
}

namespace mb2 {
// Bezier extraction of a B-spline curve by Boehm knot insertion (every interior knot raised to multiplicity p):
// Q_i = alpha P_i + (1 - alpha) P_{i-1}, alpha = (u - t_i) / (t_{i+p} - t_i); the formula and order of
// oracle_py.bspline_to_bezier. P: n points of `dim` doubles. Returns [n_spans][p+1][dim].
static std::vector<double> bspline_curve_extract(int dim, int p, std::vector<double> U, std::vector<double> P) {
  std::vector<double> interior(U.begin() + p + 1, U.end() - p - 1);
  interior.erase(std::unique(interior.begin(), interior.end()), interior.end());
  for (double u : interior) {
    while ((int)std::count(U.begin(), U.end(), u) < p) {
      int k = -1;
      for (int i = 0; i + 1 < (int)U.size(); i++)
        if (U[i] <= u && u < U[i + 1]) k = i;
      if (k < 0) return {};
      const int n = (int)(P.size() / dim);
      std::vector<double> Q((size_t)(n + 1) * dim);
      for (int i = 0; i <= n; i++) {
        for (int d = 0; d < dim; d++) {
          double v;
          if (i <= k - p) v = P[(size_t)i * dim + d];
          else if (i > k) v = P[(size_t)(i - 1) * dim + d];
          else {
            const double a = (u - U[i]) / (U[i + p] - U[i]);
            v = a * P[(size_t)i * dim + d] + (1.0 - a) * P[(size_t)(i - 1) * dim + d];
          }
          Q[(size_t)i * dim + d] = v;
        }
      }
      P.swap(Q);
      U.insert(U.begin() + k + 1, u);
    }
  }
  const int n = (int)(P.size() / dim);
  if ((n - 1) % p != 0) return {};
  const int spans = (n - 1) / p;
  std::vector<double> out;
  out.reserve((size_t)spans * (p + 1) * dim);
  for (int s = 0; s < spans; s++)
    for (int i = 0; i <= p; i++)
      for (int d = 0; d < dim; d++) out.push_back(P[(size_t)(s * p + i) * dim + d]);
  return out;
}

// Curve (dv == 0) or tensor-product surface: rows first (u), then columns (v); patch order v span slow, u span fast;
// a patch is [dv+1][du+1][dim], u fastest (oracle_py.bspline_surface_to_bezier).
static std::vector<double> bspline_extract(int dim, int du, int dv, int nu, int nv, const double* ku, const double* kv,
                                           const double* P) {
  const std::vector<double> Uu(ku, ku + nu + du + 1);
  if (dv == 0) return bspline_curve_extract(dim, du, Uu, std::vector<double>(P, P + (size_t)nu * dim));
  const std::vector<double> Uv(kv, kv + nv + dv + 1);
  std::vector<std::vector<double>> rows(nv);
  for (int j = 0; j < nv; j++) {
    rows[j] = bspline_curve_extract(dim, du, Uu, std::vector<double>(P + (size_t)j * nu * dim, P + (size_t)(j + 1) * nu * dim));
    if (rows[j].empty()) return {};
  }
  const int su = (int)(rows[0].size() / ((du + 1) * dim));
  std::vector<std::vector<double>> cols((size_t)su * (du + 1));
  int sv = 0;
  for (int a = 0; a < su; a++)
    for (int i = 0; i <= du; i++) {
      std::vector<double> line((size_t)nv * dim);
      for (int j = 0; j < nv; j++)
        for (int d = 0; d < dim; d++) line[(size_t)j * dim + d] = rows[j][((size_t)a * (du + 1) + i) * dim + d];
      cols[(size_t)a * (du + 1) + i] = bspline_curve_extract(dim, dv, Uv, line);
      if (cols[(size_t)a * (du + 1) + i].empty()) return {};
      sv = (int)(cols[(size_t)a * (du + 1) + i].size() / ((dv + 1) * dim));
    }
  std::vector<double> out;
  for (int b = 0; b < sv; b++)
    for (int a = 0; a < su; a++)
      for (int jj = 0; jj <= dv; jj++)
        for (int i = 0; i <= du; i++)
          for (int d = 0; d < dim; d++) out.push_back(cols[(size_t)a * (du + 1) + i][((size_t)b * (dv + 1) + jj) * dim + d]);
  return out;
}
}  // namespace mb2

int mb2_contact_set_rigid_body(mb2_ctx* c, int kind, const double* params, int nparams, double penalty) {
  CCHECK(c);
  ContactState* S = static_cast<ContactState*>(c->contact);
  CREQ(S, "call mb2_contact_set_faces first");
  CREQ(kind == MB2_RIGID_PLANE || kind == MB2_RIGID_SPHERE || kind == MB2_RIGID_BEZIER || kind == MB2_RIGID_BSPLINE ||
           kind == MB2_RIGID_RATIONAL_BEZIER || kind == MB2_RIGID_NURBS,
       "unknown rigid body kind");
  const int rational = (kind == MB2_RIGID_RATIONAL_BEZIER || kind == MB2_RIGID_NURBS) ? 1 : 0;
  if (kind == MB2_RIGID_BEZIER || kind == MB2_RIGID_BSPLINE || rational) {
    CREQ(params && nparams >= 3, "spline body needs {deg_u, deg_v, n_seed, ...}");
    const int du = (int)params[0], dv = (int)params[1], ns = (int)params[2];
    CREQ(du >= 1 && du <= mb2::BEZ_MAXDEG && dv >= 0 && dv <= mb2::BEZ_MAXDEG && ns >= 1, "spline degree must be 1..5");
    CREQ((c->dim == 2) == (dv == 0), "a 2D body is a curve (deg_v = 0), a 3D body a surface");
    const int hd = c->dim + rational;  // doubles per control point
    const int ncp = (du + 1) * (dv + 1) * hd;
    std::vector<double> patches;
    // rational bodies: {x, w} per point -> homogeneous {w x, w}
    auto to_homogeneous = [&](const double* P, size_t npts, std::vector<double>& out) -> bool {
      out.resize(npts * hd);
      for (size_t i = 0; i < npts; i++) {
        const double w = P[i * hd + c->dim];
        if (!(w > 0.0)) return false;
        for (int d = 0; d < c->dim; d++) out[i * hd + d] = P[i * hd + d] * w;
        out[i * hd + c->dim] = w;
      }
      return true;
    };
    if (kind == MB2_RIGID_BEZIER || kind == MB2_RIGID_RATIONAL_BEZIER) {
      CREQ(nparams > 3 && (nparams - 3) % ncp == 0, "Bezier control point count does not match the degrees");
      if (rational) {
        CREQ(to_homogeneous(params + 3, (size_t)(nparams - 3) / hd, patches), "weights must be positive");
      } else {
        patches.assign(params + 3, params + nparams);
      }
    } else {
      CREQ(nparams >= 5, "B-spline body needs {deg_u, deg_v, n_seed, ncp_u, ncp_v, knots_u, knots_v, control points}");
      const int nu = (int)params[3], nv = (int)params[4];
      CREQ(nu >= du + 1 && nv >= dv + 1 && (dv > 0 || nv == 1), "B-spline control point counts");
      const int nku = nu + du + 1, nkv = dv > 0 ? nv + dv + 1 : 0;
      CREQ(nparams == 5 + nku + nkv + nu * nv * hd, "B-spline parameter count");
      const double* ku = params + 5;
      const double* kv = ku + nku;
      const double* P = kv + nkv;
      std::vector<double> Ph;
      if (rational) {
        CREQ(to_homogeneous(P, (size_t)nu * nv, Ph), "weights must be positive");
        P = Ph.data();
      }
      for (int i = 0; i + 1 < nku; i++) CREQ(ku[i] <= ku[i + 1], "knot vectors must be non-decreasing");
      for (int i = 0; i + 1 < nkv; i++) CREQ(kv[i] <= kv[i + 1], "knot vectors must be non-decreasing");
      patches = mb2::bspline_extract(hd, du, dv, nu, nv, ku, kv, P);
      CREQ(!patches.empty(), "B-spline body: open knot vectors (end multiplicity degree + 1) expected");
    }
    const int np = (int)(patches.size() / ncp);
    MB2_CUDA(cudaSetDevice(c->device));
    MB2_CUDA(cudaStreamSynchronize(c->stream));  // a previous assembly may still read the old control net
    if (S->d_cp) cudaFree(S->d_cp);
    S->d_cp = nullptr;
    MB2_CUDA(cudaMalloc(&S->d_cp, sizeof(double) * patches.size()));
    MB2_CUDA(cudaMemcpy(S->d_cp, patches.data(), sizeof(double) * patches.size(), cudaMemcpyHostToDevice));
    S->body = mb2::RigidBody();
    S->body.kind = MB2_RIGID_BEZIER;
    S->body.deg_u = du; S->body.deg_v = dv; S->body.n_seed = ns;
    S->body.cp = S->d_cp;
    S->body.n_patches = np;
    S->body.rational = rational;
    S->penalty = penalty;
    return MB2_OK;
  }
  CREQ(params && nparams >= (kind == MB2_RIGID_PLANE ? 11 : 4) && nparams <= 16, "rigid body parameter count");
  S->body = mb2::RigidBody();
  S->body.kind = kind;
  for (int i = 0; i < 16; i++) S->body.p[i] = i < nparams ? params[i] : 0.0;
  S->penalty = penalty;
  return MB2_OK;
}

int mb2_contact_n_marked_nodes(mb2_ctx* c, int64_t* n) {
  CCHECK(c);
  ContactState* S = static_cast<ContactState*>(c->contact);
  CREQ(S && n, "call mb2_contact_set_faces first");
  *n = S->n_marked;
  return MB2_OK;
}

int mb2_contact_get_marked_v_dofs(mb2_ctx* c, int32_t* out) {
  CCHECK(c);
  ContactState* S = static_cast<ContactState*>(c->contact);
  CREQ(S && out, "call mb2_contact_set_faces first");
  for (size_t i = 0; i < S->marked_nodes.size(); i++)
    for (int d = 0; d < c->dim; d++) out[i * c->dim + d] = c->dim * S->marked_nodes[i] + d;
  return MB2_OK;
}

int mb2_contact_get_A_ids_i32(mb2_ctx* c, int32_t* A_ids) {
  CCHECK(c);
  ContactState* S = static_cast<ContactState*>(c->contact);
  CREQ(S && A_ids, "call mb2_contact_set_faces first");
  CREQ(c->nnz <= 2147483647LL, "nnz does not fit int");
  const int64_t n = S->NF * S->tdf * S->tdf;
  int32_t* d_tmp = nullptr;
  MB2_CUDA(cudaMalloc(&d_tmp, sizeof(int32_t) * n));
  mb2::face_A_ids_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(c->dim, S->NF, S->ndf, S->d_fvdofs, S->d_frank,
                                                                            c->d_I, d_tmp);
  MB2_CUDA(cudaGetLastError());
  MB2_CUDA(cudaMemcpyAsync(A_ids, d_tmp, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  cudaFree(d_tmp);
  return MB2_OK;
}

int mb2_contact_residual_device(mb2_ctx* c) {
  CCHECK(c);
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::contact_assemble(c, false, 0.0);
}
int mb2_contact_residual_and_grad_device(mb2_ctx* c, double gf) {
  CCHECK(c);
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::contact_assemble(c, true, gf);
}

int mb2_add_boundary_residual(mb2_ctx* c, const double* u, double* r) {
  CCHECK(c);
  CREQ(u && r, "null vector");
  int rc;
  if ((rc = mb2_set_u(c, u))) return rc;
  if ((rc = mb2_set_r(c, r))) return rc;
  if ((rc = mb2_contact_residual_device(c))) return rc;
  return mb2_get_r(c, r);
}

// Only the CSR rows of the marked boundary nodes travel: they are the only values the faces touch.
int mb2_add_boundary_residual_and_grad(mb2_ctx* c, const double* u, double gf, double* r, double* A) {
  CCHECK(c);
  CREQ(u && r && A, "null vector");
  ContactState* S = static_cast<ContactState*>(c->contact);
  CREQ(S, "call mb2_contact_set_faces first");
  MB2_CUDA(cudaSetDevice(c->device));
  int rc;
  if ((rc = mb2_set_u(c, u))) return rc;
  if ((rc = mb2_set_r(c, r))) return rc;
  for (auto& run : S->runs)
    MB2_CUDA(cudaMemcpyAsync(c->d_A + run.first, A + run.first, sizeof(double) * run.second, cudaMemcpyHostToDevice, c->stream));
  if ((rc = mb2_contact_residual_and_grad_device(c, gf))) return rc;
  for (auto& run : S->runs)
    MB2_CUDA(cudaMemcpyAsync(A + run.first, c->d_A + run.first, sizeof(double) * run.second, cudaMemcpyDeviceToHost, c->stream));
  return mb2_get_r(c, r);
}

int mb2_contact_get_nodal(mb2_ctx* c, double* gap, double* area, double* pressure) {
  CCHECK(c);
  ContactState* S = static_cast<ContactState*>(c->contact);
  CREQ(S, "call mb2_contact_set_faces first");
  const size_t b = sizeof(double) * (size_t)S->n_marked;
  if (gap) MB2_CUDA(cudaMemcpyAsync(gap, S->d_gap, b, cudaMemcpyDeviceToHost, c->stream));
  if (area) MB2_CUDA(cudaMemcpyAsync(area, S->d_area, b, cudaMemcpyDeviceToHost, c->stream));
  if (pressure) MB2_CUDA(cudaMemcpyAsync(pressure, S->d_press, b, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}

int mb2_contact_last_scalars(mb2_ctx* c, double* out5) {
  CCHECK(c);
  ContactState* S = static_cast<ContactState*>(c->contact);
  CREQ(S && out5, "call mb2_contact_set_faces first");
  MB2_CUDA(cudaMemcpyAsync(out5, S->d_scal, sizeof(double) * 5, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}

int mb2_contact_gap_norm(mb2_ctx* c, const double* u, double* out) {
  CCHECK(c);
  CREQ(u && out, "null argument");
  MB2_CUDA(cudaSetDevice(c->device));
  int rc;
  if ((rc = mb2_set_u(c, u))) return rc;
  return mb2::contact_gap_norm(c, out);
}

}  // extern "C"
