// Device-side regeneration of the per-quadrature-point tables of one tensor-product NURBS patch.
//
// Replaces PrecomputedData::PrecomputeElementQuadData (src/mimi/utils/precomputed.cpp:264-330), whose
// heap-allocated per-qp objects (QuadData, precomputed.hpp:58-69) take ~1.5 KB per quadrature point and
// cannot be built on a host for the 1e6-element configurations. Every value is produced in MFEM's
// operation order (this translation unit is compiled -fmad=false), so the tables are bit-identical to
// the reference's:
//   1D Gauss-Legendre rule        third_party/mfem/fem/intrules.cpp:424-494 (host: needs libm cos)
//   tensor rule ordering/weights  third_party/mfem/fem/intrules.cpp:56-84   (x fastest; (wx*wy)*wz)
//   1D B-spline values            KnotVector::CalcShape,  third_party/mfem/mesh/nurbs.cpp:297-320
//   1D B-spline derivatives       KnotVector::CalcDShape, third_party/mfem/mesh/nurbs.cpp:322-376
//   rational 2D/3D gradients      NURBS{2,3}DFiniteElement::CalcDShape, third_party/mfem/fem/fe/fe_nurbs.cpp:130-175,269-311
//   geometry Jacobian             IsoparametricTransformation::EvalJacobian, third_party/mfem/fem/eltrans.cpp:406-420
//   det / inverse / dN_dX         DenseMatrix::Det (densemat.cpp:510-525), kernels::CalcInverse
//                                 (kernels.hpp:244-251), kernels::Mult (kernels.hpp:341-357)
#include <algorithm>
#include <cmath>
#include <utility>
#include <vector>

#include "mb2_ctx.h"
#include "mb2_math.cuh"

namespace mb2 {

constexpr int MAXP = 4;    // degree limit of the instantiated kernels
constexpr int MAXQ1 = 8;   // 1D quadrature points limit

// QuadratureFunctions1D::GaussLegendre on [0,1] (intrules.cpp:424-494), host arithmetic.
static void gauss_legendre_01(int np, std::vector<double>& x, std::vector<double>& w) {
  x.assign(np, 0.0);
  w.assign(np, 0.0);
  switch (np) {
    case 1: x[0] = 0.5; w[0] = 1.0; return;
    case 2:
      x[0] = 0.21132486540518711775; w[0] = 0.5;
      x[1] = 0.78867513459481288225; w[1] = 0.5;
      return;
    case 3:
      x[0] = 0.11270166537925831148; w[0] = 5. / 18.;
      x[1] = 0.5; w[1] = 4. / 9.;
      x[2] = 0.88729833462074168852; w[2] = 5. / 18.;
      return;
  }
  const int n = np, m = (n + 1) / 2;
  for (int i = 1; i <= m; i++) {
    double z = std::cos(M_PI * (i - 0.25) / (n + 0.5));
    double pp, p1, dz, xi = 0.;
    bool done = false;
    while (1) {
      double p2 = 1;
      p1 = z;
      for (int j = 2; j <= n; j++) {
        const double p3 = p2;
        p2 = p1;
        p1 = ((2 * j - 1) * z * p2 - (j - 1) * p3) / j;
      }
      pp = n * (z * p1 - p2) / (z * z - 1);
      if (done) break;
      dz = p1 / pp;
      if (std::fabs(dz) < 1e-16) {
        done = true;
        xi = ((1 - z) + dz) / 2;
      }
      z -= dz;
    }
    x[i - 1] = xi;
    x[n - i] = 1 - xi;
    w[i - 1] = w[n - i] = 1. / (4 * xi * (1 - xi) * pp * pp);
  }
}

// 1D tables: for direction d, knot span s (ip = s + p), 1D point q: shape[0..p], dshape[0..p].
// layout: B1[((s * nq1 + q) * (p+1)) + r]
__global__ void k_basis_1d(int p, int nspan, int nq1, const double* __restrict__ knot, const double* __restrict__ gx,
                           double* __restrict__ B1, double* __restrict__ D1) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nspan * nq1) return;
  const int s = t / nq1, q = t % nq1;
  const int ip = s + p;
  const double xi = gx[q];
  // KnotVector::getKnotLocation (nurbs.hpp:99-100)
  const double u = (xi * knot[ip + 1] + (1. - xi) * knot[ip]);
  double left[MAXP + 1], right[MAXP + 1];
  {
    // CalcShape, Algorithm A2.2
    double shape[MAXP + 1];
    shape[0] = 1.;
    for (int j = 1; j <= p; ++j) {
      left[j] = u - knot[ip + 1 - j];
      right[j] = knot[ip + j] - u;
      double saved = 0.;
      for (int r = 0; r < j; ++r) {
        const double tmp = shape[r] / (right[r + 1] + left[j - r]);
        shape[r] = saved + right[r + 1] * tmp;
        saved = left[j - r] * tmp;
      }
      shape[j] = saved;
    }
    for (int r = 0; r <= p; r++) B1[(size_t)t * (p + 1) + r] = shape[r];
  }
  {
    // CalcDShape, Algorithm A2.3
    double ndu[MAXP + 1][MAXP + 1];
    ndu[0][0] = 1.0;
    for (int j = 1; j <= p; j++) {
      left[j] = u - knot[ip - j + 1];
      right[j] = knot[ip + j] - u;
      double saved = 0.0;
      for (int r = 0; r < j; r++) {
        ndu[j][r] = right[r + 1] + left[j - r];
        const double temp = ndu[r][j - 1] / ndu[j][r];
        ndu[r][j] = saved + right[r + 1] * temp;
        saved = left[j - r] * temp;
      }
      ndu[j][j] = saved;
    }
    const double scale = p * (knot[ip + 1] - knot[ip]);
    for (int r = 0; r <= p; ++r) {
      double d = 0.0;
      const int rk = r - 1, pk = p - 1;
      if (r >= 1) d = ndu[rk][pk] / ndu[p][rk];
      if (r <= pk) d -= ndu[r][pk] / ndu[p][r];
      D1[(size_t)t * (p + 1) + r] = d * scale;  // grad *= p*(knot(ip+1) - knot(ip))
    }
  }
}

struct TabArgs {
  int64_t e0, ne;
  int nq1;
  const int32_t* __restrict__ vdofs;   // all elements, NE x tdof
  const int32_t* __restrict__ el_ijk;  // ne x DIM, relative to e0
  const double* __restrict__ B1[3];
  const double* __restrict__ D1[3];
  const double* __restrict__ gw;       // 1D weights
  const double* __restrict__ x_ref;    // NV byVDIM
  const double* __restrict__ wts;      // Nn or nullptr (all 1.0)
  double* __restrict__ dNdX;
  double* __restrict__ wdetJ;
  double* __restrict__ N;              // optional: NE x nq x nd shape values
  int* __restrict__ flag;
};

// One thread per (element, quadrature point).
template <int DIM, int P>
__global__ void __launch_bounds__(128) k_tables(const TabArgs g) {
  constexpr int P1 = P + 1;
  constexpr int ND = (DIM == 3) ? P1 * P1 * P1 : P1 * P1;
  constexpr int TD = DIM * ND;
  const int nq1 = g.nq1;
  const int nq = (DIM == 3) ? nq1 * nq1 * nq1 : nq1 * nq1;
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= g.ne * nq) return;
  const int64_t el = t / nq;
  const int q = (int)(t % nq);
  const int64_t e = g.e0 + el;
  int qd[3], sp[3];
  qd[0] = q % nq1;
  qd[1] = (q / nq1) % nq1;
  qd[2] = (DIM == 3) ? q / (nq1 * nq1) : 0;
  for (int d = 0; d < DIM; d++) sp[d] = g.el_ijk[el * DIM + d];
  const double *sx = g.B1[0] + ((size_t)sp[0] * nq1 + qd[0]) * P1, *dsx = g.D1[0] + ((size_t)sp[0] * nq1 + qd[0]) * P1;
  const double *sy = g.B1[1] + ((size_t)sp[1] * nq1 + qd[1]) * P1, *dsy = g.D1[1] + ((size_t)sp[1] * nq1 + qd[1]) * P1;
  const double *sz = nullptr, *dsz = nullptr;
  if (DIM == 3) {
    sz = g.B1[2] + ((size_t)sp[2] * nq1 + qd[2]) * P1;
    dsz = g.D1[2] + ((size_t)sp[2] * nq1 + qd[2]) * P1;
  }
  const int32_t* vd = g.vdofs + e * TD;

  double uu[ND], ds[TD];  // ds(o, d) = ds[o + d*ND]
  double sum = 0.0, dsum[3] = {0.0, 0.0, 0.0};
  if (DIM == 3) {
    int o = 0;
    for (int k = 0; k < P1; k++) {
      const double szk = sz[k], dszk = dsz[k];
      for (int j = 0; j < P1; j++) {
        const double sy_sz = sy[j] * szk;
        const double dsy_sz = dsy[j] * szk;
        const double sy_dsz = sy[j] * dszk;
        for (int i = 0; i < P1; i++, o++) {
          const double w = g.wts ? g.wts[vd[o] / DIM] : 1.0;
          sum += (uu[o] = sx[i] * sy_sz * w);
          dsum[0] += (ds[o] = dsx[i] * sy_sz * w);
          dsum[1] += (ds[o + ND] = sx[i] * dsy_sz * w);
          dsum[2] += (ds[o + 2 * ND] = sx[i] * sy_dsz * w);
        }
      }
    }
  } else {
    // NURBS2DFiniteElement::CalcDShape (fe_nurbs.cpp:130-175)
    int o = 0;
    for (int j = 0; j < P1; j++) {
      const double syj = sy[j], dsyj = dsy[j];
      for (int i = 0; i < P1; i++, o++) {
        const double w = g.wts ? g.wts[vd[o] / DIM] : 1.0;
        sum += (uu[o] = sx[i] * syj * w);
        dsum[0] += (ds[o] = dsx[i] * syj * w);
        dsum[1] += (ds[o + ND] = sx[i] * dsyj * w);
      }
    }
  }
  if (g.N) {
    // CalcShape: shape /= sum  -> m = 1.0/sum; shape[i] *= m  (vector.cpp:183-191)
    const double m = 1.0 / sum;
    double* No = g.N + (e * nq + q) * (int64_t)ND;
    for (int o = 0; o < ND; o++) No[o] = uu[o] * m;
  }
  sum = 1.0 / sum;
  for (int d = 0; d < DIM; d++) dsum[d] *= sum * sum;
  for (int o = 0; o < ND; o++)
    for (int d = 0; d < DIM; d++) ds[o + d * ND] = ds[o + d * ND] * sum - uu[o] * dsum[d];

  // dFdx = PointMat * dshape (kernels::Mult: zero, then j, k, i)
  double Jm[DIM * DIM];
  for (int i = 0; i < DIM * DIM; i++) Jm[i] = 0.0;
  for (int j = 0; j < DIM; j++)
    for (int k = 0; k < ND; k++) {
      const double c = ds[k + j * ND];
      for (int i = 0; i < DIM; i++) Jm[i + DIM * j] += g.x_ref[vd[i * ND + k]] * c;
    }
  const double detJ = det_small<DIM>(Jm);
  if (!(detJ > 0.0)) atomicExch(g.flag, 2);
  double Ji[DIM * DIM];
  inverse_small<DIM>(Jm, Ji);
  // dN_dX = dN_dxi * dxi_dX (kernels::Mult)
  double* out = g.dNdX + (e * nq + q) * (int64_t)TD;
  for (int j = 0; j < DIM; j++)
    for (int i = 0; i < ND; i++) {
      double a = 0.0;
      for (int k = 0; k < DIM; k++) a += ds[i + k * ND] * Ji[k + DIM * j];
      out[i + j * ND] = a;
    }
  // integration weight: (wx*wy)*wz (intrules.cpp:80), times det_dX_dxi at use (nonlinear_solid.hpp:79)
  double w = g.gw[qd[0]] * g.gw[qd[1]];
  if (DIM == 3) w = w * g.gw[qd[2]];
  g.wdetJ[e * nq + q] = w * detJ;
}

int set_nodes(mb2_ctx* c, const double* x_ref, const double* weights) {
  if (!c->d_xref) MB2_CUDA(cudaMalloc(&c->d_xref, sizeof(double) * c->NV));
  MB2_CUDA(cudaMemcpyAsync(c->d_xref, x_ref, sizeof(double) * c->NV, cudaMemcpyHostToDevice, c->stream));
  if (weights) {
    if (!c->d_nodew) MB2_CUDA(cudaMalloc(&c->d_nodew, sizeof(double) * c->Nn));
    MB2_CUDA(cudaMemcpyAsync(c->d_nodew, weights, sizeof(double) * c->Nn, cudaMemcpyHostToDevice, c->stream));
  } else if (c->d_nodew) {
    cudaFree(c->d_nodew);
    c->d_nodew = nullptr;
  }
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}

int generate_patch_tables(mb2_ctx* c, int64_t e0, int64_t ne, int p, const int32_t* ncp, const double* const* knots,
                          const int32_t* el_ijk, int quad_order, bool want_N) {
  const int dim = c->dim;
  if (p < 1 || p > MAXP) { c->err = "degree must be 1..4"; return MB2_ERR_UNSUPPORTED; }
  int nd = 1;
  for (int d = 0; d < dim; d++) nd *= p + 1;
  if (nd != c->nd) { c->err = "(p+1)^dim does not match nd of the connectivity"; return MB2_ERR_ARG; }
  // IntegrationRules::Get(geom, order): segment real order = order | 1, n = order/2 + 1 (intrules.cpp:1004-1060)
  const int order = quad_order < 0 ? 2 * p + 3 : quad_order;
  const int nq1 = (order | 1) / 2 + 1;
  int nq = 1;
  for (int d = 0; d < dim; d++) nq *= nq1;
  if (nq != c->nq || nq1 > MAXQ1) { c->err = "quadrature order does not match nq of the connectivity"; return MB2_ERR_ARG; }
  if (!c->d_xref) { c->err = "call mb2_set_nodes first"; return MB2_ERR_ARG; }
  if (e0 < 0 || ne <= 0 || e0 + ne > c->NE) { c->err = "element range out of bounds"; return MB2_ERR_ARG; }

  std::vector<double> gx, gw;
  gauss_legendre_01(nq1, gx, gw);
  double *d_gx = nullptr, *d_gw = nullptr, *d_knot[3] = {nullptr, nullptr, nullptr};
  double *d_B1[3] = {nullptr, nullptr, nullptr}, *d_D1[3] = {nullptr, nullptr, nullptr};
  int32_t* d_ijk = nullptr;
  DevTmp tmp;
  tmp.own(&d_gx); tmp.own(&d_gw); tmp.own(&d_ijk);
  for (int d = 0; d < 3; d++) { tmp.own(&d_knot[d]); tmp.own(&d_B1[d]); tmp.own(&d_D1[d]); }
  MB2_CUDA(cudaMalloc(&d_gx, sizeof(double) * nq1));
  MB2_CUDA(cudaMalloc(&d_gw, sizeof(double) * nq1));
  MB2_CUDA(cudaMemcpyAsync(d_gx, gx.data(), sizeof(double) * nq1, cudaMemcpyHostToDevice, c->stream));
  MB2_CUDA(cudaMemcpyAsync(d_gw, gw.data(), sizeof(double) * nq1, cudaMemcpyHostToDevice, c->stream));
  for (int d = 0; d < dim; d++) {
    const int nk = ncp[d] + p + 1;
    const int nspan = ncp[d] - p;  // knot spans i = 0 .. ncp-p-1 (ip = i + p), zero-length ones included
    MB2_CUDA(cudaMalloc(&d_knot[d], sizeof(double) * nk));
    MB2_CUDA(cudaMemcpyAsync(d_knot[d], knots[d], sizeof(double) * nk, cudaMemcpyHostToDevice, c->stream));
    MB2_CUDA(cudaMalloc(&d_B1[d], sizeof(double) * nspan * nq1 * (p + 1)));
    MB2_CUDA(cudaMalloc(&d_D1[d], sizeof(double) * nspan * nq1 * (p + 1)));
    k_basis_1d<<<(nspan * nq1 + 127) / 128, 128, 0, c->stream>>>(p, nspan, nq1, d_knot[d], d_gx, d_B1[d], d_D1[d]);
    MB2_CUDA(cudaGetLastError());
  }
  MB2_CUDA(cudaMalloc(&d_ijk, sizeof(int32_t) * ne * dim));
  MB2_CUDA(cudaMemcpyAsync(d_ijk, el_ijk, sizeof(int32_t) * ne * dim, cudaMemcpyHostToDevice, c->stream));
  const int64_t nqp = c->NE * c->nq;
  if (!c->d_dNdX) MB2_CUDA(cudaMalloc(&c->d_dNdX, sizeof(double) * nqp * c->tdof));
  if (!c->d_wdetJ) MB2_CUDA(cudaMalloc(&c->d_wdetJ, sizeof(double) * nqp));
  if (want_N && !c->d_N) MB2_CUDA(cudaMalloc(&c->d_N, sizeof(double) * nqp * c->nd));
  MB2_CUDA(cudaMemsetAsync(c->d_flag, 0, sizeof(int), c->stream));

  TabArgs a;
  a.e0 = e0; a.ne = ne; a.nq1 = nq1;
  a.vdofs = c->d_vdofs; a.el_ijk = d_ijk;
  for (int d = 0; d < 3; d++) { a.B1[d] = d_B1[d]; a.D1[d] = d_D1[d]; }
  a.gw = d_gw; a.x_ref = c->d_xref; a.wts = c->d_nodew;
  a.dNdX = c->d_dNdX; a.wdetJ = c->d_wdetJ; a.N = want_N ? c->d_N : nullptr; a.flag = c->d_flag;
  const int64_t nthr = ne * nq;
  const unsigned grid = (unsigned)((nthr + 127) / 128);
  bool ok = true;
#define MB2_TAB(D, PP) if (dim == D && p == PP) k_tables<D, PP><<<grid, 128, 0, c->stream>>>(a); else
  MB2_TAB(3, 1) MB2_TAB(3, 2) MB2_TAB(3, 3) MB2_TAB(2, 1) MB2_TAB(2, 2) MB2_TAB(2, 3) MB2_TAB(2, 4) { ok = false; }
#undef MB2_TAB
  if (!ok) { c->err = "no table kernel for this (dim, p)"; return MB2_ERR_UNSUPPORTED; }
  MB2_CUDA(cudaGetLastError());
  int flag = 0;
  MB2_CUDA(cudaMemcpyAsync(&flag, c->d_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  if (flag) { c->err = "non-positive geometry Jacobian determinant while generating tables"; return MB2_ERR_NUMERIC; }
  return MB2_OK;
}


// ---------------------------------------------------------------------------------------------
// Mass matrix on the device: mfem::VectorMassIntegrator::AssembleElementMatrix
// (third_party/mfem/fem/bilininteg.cpp:1542-1620) as the reference adds it in PyNonlinearSolid::Setup
// (src/mimi/py/py_nonlinear_solid.cpp:154-173), gathered by BilinearForm::Assemble in ascending element order.
//   rule      IntRules.Get(geom, 2 p + Trans.OrderW()), OrderW = p dim - 1 for Qk (eltrans.cpp:453-465)
//   per point norm = (ip.weight * Trans.Weight()) * rho;  partelmat = shape shape^t (MultVVt);  partelmat *= norm;
//             elmat(k nd + i, k nd + j) += partelmat(i, j)   -> m_ij = ((0 + (N_i N_j) norm_0) + (N_i N_j) norm_1) + ...
// The element blocks are written in the K_e layout of mb2_assembly.cu (zeros between different components, exactly
// what AddSubMatrix adds with skip_zeros = 0), so the ordered CSR gather, its chunk pipeline and the multi-GPU ghost
// exchange are reused unchanged with grad_factor = 1.
// ---------------------------------------------------------------------------------------------
struct MassTab {
  int p = 0, nq1 = 0;
  int mode = 0;  // 0: mass matrix, 2: viscosity matrix (Jacobian pipeline); 1: body-force linear form (residual pipeline)
  double rho = 1.0;
  double f[3] = {0.0, 0.0, 0.0};
  int32_t* d_ijk = nullptr;
  double *d_gw = nullptr, *d_B1[3] = {nullptr, nullptr, nullptr}, *d_D1[3] = {nullptr, nullptr, nullptr};
};

struct MassArgs {
  int64_t e0, e1;
  int nq1;
  double rho;
  double f[3];
  const int32_t* __restrict__ vdofs;
  const int32_t* __restrict__ el_ijk;
  const double* __restrict__ B1[3];
  const double* __restrict__ D1[3];
  const double* __restrict__ gw;
  const double* __restrict__ x_ref;
  const double* __restrict__ wts;
  double* __restrict__ KeT;
  double* __restrict__ Re;
};

constexpr int MASS_NT = 128;  // threads per element
constexpr int MASS_QB = 32;   // quadrature points staged per batch

// el.CalcShape(ip, shape), ip.weight and Trans.Weight() at quadrature point q of element e, in MFEM's operation order.
// GRAD: `shape` receives dshapedxt = dshape * adj(J) (ND x DIM, column-major) instead of the shape values.
template <int DIM, int P, bool GRAD = false>
__device__ void mass_point(const MassArgs& g, const int32_t* vd, int64_t e, int q, double* shape, double* ipw, double* detJ) {
  constexpr int P1 = P + 1;
  constexpr int ND = (DIM == 3) ? P1 * P1 * P1 : P1 * P1;
  constexpr int TD = DIM * ND;
  const int nq1 = g.nq1;
  int qd[3], sp[3];
  qd[0] = q % nq1;
  qd[1] = (q / nq1) % nq1;
  qd[2] = (DIM == 3) ? q / (nq1 * nq1) : 0;
  for (int d = 0; d < DIM; d++) sp[d] = g.el_ijk[e * DIM + d];
  const double *sx = g.B1[0] + ((size_t)sp[0] * nq1 + qd[0]) * P1, *dsx = g.D1[0] + ((size_t)sp[0] * nq1 + qd[0]) * P1;
  const double *sy = g.B1[1] + ((size_t)sp[1] * nq1 + qd[1]) * P1, *dsy = g.D1[1] + ((size_t)sp[1] * nq1 + qd[1]) * P1;
  double uu[ND], ds[TD];
  double sum = 0.0, dsum[3] = {0.0, 0.0, 0.0};
  if (DIM == 3) {
    const double* sz = g.B1[2] + ((size_t)sp[2] * nq1 + qd[2]) * P1;
    const double* dsz = g.D1[2] + ((size_t)sp[2] * nq1 + qd[2]) * P1;
    int o = 0;
    for (int k = 0; k < P1; k++)
      for (int j = 0; j < P1; j++) {
        const double sy_sz = sy[j] * sz[k], dsy_sz = dsy[j] * sz[k], sy_dsz = sy[j] * dsz[k];
        for (int i = 0; i < P1; i++, o++) {
          const double w = g.wts ? g.wts[vd[o] / DIM] : 1.0;
          sum += (uu[o] = sx[i] * sy_sz * w);
          dsum[0] += (ds[o] = dsx[i] * sy_sz * w);
          dsum[1] += (ds[o + ND] = sx[i] * dsy_sz * w);
          dsum[2] += (ds[o + 2 * ND] = sx[i] * sy_dsz * w);
        }
      }
  } else {
    int o = 0;
    for (int j = 0; j < P1; j++)
      for (int i = 0; i < P1; i++, o++) {
        const double w = g.wts ? g.wts[vd[o] / DIM] : 1.0;
        sum += (uu[o] = sx[i] * sy[j] * w);
        dsum[0] += (ds[o] = dsx[i] * sy[j] * w);
        dsum[1] += (ds[o + ND] = sx[i] * dsy[j] * w);
      }
  }
  // el.CalcShape: shape /= sum (fe_nurbs.cpp:125-158,236-267; Vector::operator/=: m = 1/sum, shape[i] *= m)
  const double m = 1.0 / sum;
  if (!GRAD)
    for (int o = 0; o < ND; o++) shape[o] = uu[o] * m;
  // Trans.Weight(): the Jacobian is built from CalcDShape (eltrans.cpp:406-420)
  for (int d = 0; d < DIM; d++) dsum[d] *= m * m;
  for (int o = 0; o < ND; o++)
    for (int d = 0; d < DIM; d++) ds[o + d * ND] = ds[o + d * ND] * m - uu[o] * dsum[d];
  double Jm[DIM * DIM];
  for (int i = 0; i < DIM * DIM; i++) Jm[i] = 0.0;
  for (int j = 0; j < DIM; j++)
    for (int k = 0; k < ND; k++) {
      const double cc = ds[k + j * ND];
      for (int i = 0; i < DIM; i++) Jm[i + DIM * j] += g.x_ref[vd[i * ND + k]] * cc;
    }
  double w = g.gw[qd[0]] * g.gw[qd[1]];
  if (DIM == 3) w = w * g.gw[qd[2]];
  *ipw = w;
  *detJ = det_small<DIM>(Jm);
  if (GRAD) {
    // Trans.AdjugateJacobian(): CalcAdjugate (densemat.cpp:2495-2563); a(i, j) = Jm[i + DIM * j]
    double ad[DIM * DIM];
#define A_(i, j) Jm[(i) + DIM * (j)]
#define AD_(i, j) ad[(i) + DIM * (j)]
    if (DIM == 2) {
      AD_(0, 0) = A_(1, 1); AD_(0, 1) = -A_(0, 1); AD_(1, 0) = -A_(1, 0); AD_(1, 1) = A_(0, 0);
    } else {
      AD_(0, 0) = A_(1, 1) * A_(2, 2) - A_(1, 2) * A_(2, 1);
      AD_(0, 1) = A_(0, 2) * A_(2, 1) - A_(0, 1) * A_(2, 2);
      AD_(0, 2) = A_(0, 1) * A_(1, 2) - A_(0, 2) * A_(1, 1);
      AD_(1, 0) = A_(1, 2) * A_(2, 0) - A_(1, 0) * A_(2, 2);
      AD_(1, 1) = A_(0, 0) * A_(2, 2) - A_(0, 2) * A_(2, 0);
      AD_(1, 2) = A_(0, 2) * A_(1, 0) - A_(0, 0) * A_(1, 2);
      AD_(2, 0) = A_(1, 0) * A_(2, 1) - A_(1, 1) * A_(2, 0);
      AD_(2, 1) = A_(0, 1) * A_(2, 0) - A_(0, 0) * A_(2, 1);
      AD_(2, 2) = A_(0, 0) * A_(1, 1) - A_(0, 1) * A_(1, 0);
    }
    // Mult(dshape, adjJ, dshapedxt) (kernels::Mult: zero, then j, k, i)
    for (int j = 0; j < DIM; j++)
      for (int i = 0; i < ND; i++) {
        double a = 0.0;
        for (int k = 0; k < DIM; k++) a += ds[i + k * ND] * AD_(k, j);
        shape[i + j * ND] = a;
      }
#undef A_
#undef AD_
  }
}

// VISC: mfem::VectorDiffusionIntegrator::AssembleElementMatrix (bilininteg.cpp:2854-2934) instead of the mass:
//   rule DiffusionIntegrator::GetRule = IntRules.Get(geom, 2p + dim - 1);  w = (ip.weight / Trans.Weight()) * coef;
//   dshapedxt = dshape adj(J);  pelmat(i, j) = w * sum_k dshapedxt(i, k) dshapedxt(j, k)  (Mult_a_AAt);  elmat += pelmat
template <int DIM, int P, bool VISC>
__global__ void __launch_bounds__(MASS_NT) k_mass_element(const MassArgs g) {
  constexpr int P1 = P + 1;
  constexpr int ND = (DIM == 3) ? P1 * P1 * P1 : P1 * P1;
  constexpr int TD = DIM * ND;
  constexpr int KP = (TD * DIM + 3) / 4 * 4;
  constexpr int NACC = (ND * ND + MASS_NT - 1) / MASS_NT;
  constexpr int QB = VISC ? MASS_QB / 2 : MASS_QB;  // dshapedxt rows are DIM times longer
  constexpr int ROW = VISC ? ND * DIM : ND;
  __shared__ double sN[QB][ROW];
  __shared__ double sNorm[QB];
  const int nq1 = g.nq1;
  const int nq = (DIM == 3) ? nq1 * nq1 * nq1 : nq1 * nq1;
  const int t = threadIdx.x;
  for (int64_t e = g.e0 + blockIdx.x; e < g.e1; e += gridDim.x) {
    const int32_t* vd = g.vdofs + e * TD;
    double acc[NACC];
#pragma unroll
    for (int k = 0; k < NACC; k++) acc[k] = 0.0;
    for (int q0 = 0; q0 < nq; q0 += QB) {
      const int nb = min(QB, nq - q0);
      if (t < nb) {
        double w, dj;
        mass_point<DIM, P, VISC>(g, vd, e, q0 + t, sN[t], &w, &dj);
        sNorm[t] = VISC ? (w / dj) * g.rho : (w * dj) * g.rho;
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < NACC; k++) {
        const int idx = t + k * MASS_NT;
        if (idx < ND * ND) {
          const int i = idx % ND, j = idx / ND;
          double a = acc[k];
          if (VISC) {
            for (int s = 0; s < nb; s++) {
              double d = 0.0;
              for (int k = 0; k < DIM; k++) d += sN[s][i + k * ND] * sN[s][j + k * ND];
              a += sNorm[s] * d;
            }
          } else {
            for (int s = 0; s < nb; s++) a += (sN[s][i] * sN[s][j]) * sNorm[s];
          }
          acc[k] = a;
        }
      }
      __syncthreads();
    }
    double* out = g.KeT + (e - g.e0) * (int64_t)(ND * KP);
#pragma unroll
    for (int k = 0; k < NACC; k++) {
      const int idx = t + k * MASS_NT;
      if (idx < ND * ND) {
        const int a = idx % ND, b = idx / ND;  // row node, column node
        for (int cj = 0; cj < DIM; cj++)
          for (int ci = 0; ci < DIM; ci++) out[(int64_t)a * KP + (b + cj * ND) * DIM + ci] = (ci == cj) ? acc[k] : 0.0;
      }
    }
    for (int i = t; i < TD; i += MASS_NT) g.Re[e * TD + i] = 0.0;  // the pipeline also gathers R_e: r += 0
  }
}

// Body-force linear form: mfem::VectorDomainLFIntegrator::AssembleRHSElementVect (third_party/mfem/fem/lininteg.cpp) with
// a VectorConstantCoefficient, as PyNonlinearSolid::Setup adds it (src/mimi/py/py_nonlinear_solid.cpp:221-241):
//   rule IntRules.Get(geom, 2 p);  per point val = Tr.Weight(), cf = val * f_k,  elvect(nd k + s) += ip.weight * cf * shape(s)
// The element vectors land in R_e; the ordered residual gather adds them in ascending element order (LinearForm::Assemble).
template <int DIM, int P>
__global__ void __launch_bounds__(MASS_NT) k_lform_element(const MassArgs g) {
  constexpr int P1 = P + 1;
  constexpr int ND = (DIM == 3) ? P1 * P1 * P1 : P1 * P1;
  constexpr int TD = DIM * ND;
  __shared__ double sN[MASS_QB][ND];
  __shared__ double sW[MASS_QB], sVal[MASS_QB];
  const int nq1 = g.nq1;
  const int nq = (DIM == 3) ? nq1 * nq1 * nq1 : nq1 * nq1;
  const int t = threadIdx.x;
  for (int64_t e = g.e0 + blockIdx.x; e < g.e1; e += gridDim.x) {
    const int32_t* vd = g.vdofs + e * TD;
    double acc0 = 0.0, acc1 = 0.0;  // TD <= 2 * MASS_NT
    for (int q0 = 0; q0 < nq; q0 += MASS_QB) {
      const int nb = min(MASS_QB, nq - q0);
      if (t < nb) mass_point<DIM, P>(g, vd, e, q0 + t, sN[t], &sW[t], &sVal[t]);
      __syncthreads();
      for (int s = 0; s < nb; s++) {
        if (t < TD) acc0 += sW[s] * (sVal[s] * g.f[t / ND]) * sN[s][t % ND];
        if (t + MASS_NT < TD) acc1 += sW[s] * (sVal[s] * g.f[(t + MASS_NT) / ND]) * sN[s][(t + MASS_NT) % ND];
      }
      __syncthreads();
    }
    if (t < TD) g.Re[e * TD + t] = acc0;
    if (t + MASS_NT < TD) g.Re[e * TD + t + MASS_NT] = acc1;
  }
}

int launch_lform_element(mb2_ctx* c, int64_t e0, int64_t e1, double* Re, cudaStream_t st) {
  const MassTab* T = static_cast<const MassTab*>(c->mass_tab);
  MassArgs a;
  a.e0 = e0; a.e1 = e1; a.nq1 = T->nq1; a.rho = 1.0;
  for (int d = 0; d < 3; d++) a.f[d] = T->f[d];
  a.vdofs = c->d_vdofs; a.el_ijk = T->d_ijk;
  for (int d = 0; d < 3; d++) { a.B1[d] = T->d_B1[d]; a.D1[d] = T->d_D1[d]; }
  a.gw = T->d_gw; a.x_ref = c->d_xref; a.wts = c->d_nodew;
  a.KeT = nullptr; a.Re = Re;
  const int64_t ne = e1 - e0;
  if (ne <= 0) return MB2_OK;
  const unsigned grid = (unsigned)std::min<int64_t>(ne, (int64_t)c->sm_count * 8);
  const int dim = c->dim, p = T->p;
  bool ok = true;
#define MB2_LF(D, PP) if (dim == D && p == PP) k_lform_element<D, PP><<<grid, MASS_NT, 0, st>>>(a); else
  MB2_LF(3, 1) MB2_LF(3, 2) MB2_LF(3, 3) MB2_LF(2, 1) MB2_LF(2, 2) MB2_LF(2, 3) MB2_LF(2, 4) { ok = false; }
#undef MB2_LF
  if (!ok) { c->err = "no linear-form kernel for this (dim, p)"; return MB2_ERR_UNSUPPORTED; }
  MB2_CUDA(cudaGetLastError());
  return MB2_OK;
}

int launch_mass_element(mb2_ctx* c, int64_t e0, int64_t e1, double* KeT, double* Re, cudaStream_t st) {
  const MassTab* T = static_cast<const MassTab*>(c->mass_tab);
  MassArgs a;
  a.e0 = e0; a.e1 = e1; a.nq1 = T->nq1; a.rho = T->rho;
  for (int d = 0; d < 3; d++) a.f[d] = 0.0;
  a.vdofs = c->d_vdofs; a.el_ijk = T->d_ijk;
  for (int d = 0; d < 3; d++) { a.B1[d] = T->d_B1[d]; a.D1[d] = T->d_D1[d]; }
  a.gw = T->d_gw; a.x_ref = c->d_xref; a.wts = c->d_nodew;
  a.KeT = KeT; a.Re = Re;
  const int64_t ne = e1 - e0;
  if (ne <= 0) return MB2_OK;
  const unsigned grid = (unsigned)std::min<int64_t>(ne, (int64_t)c->sm_count * 8);
  const int dim = c->dim, p = T->p;
  bool ok = true;
#define MB2_MASS(D, PP)                                                      \
  if (dim == D && p == PP) {                                                 \
    if (T->mode == 2) k_mass_element<D, PP, true><<<grid, MASS_NT, 0, st>>>(a); \
    else k_mass_element<D, PP, false><<<grid, MASS_NT, 0, st>>>(a);          \
  } else
  MB2_MASS(3, 1) MB2_MASS(3, 2) MB2_MASS(3, 3) MB2_MASS(2, 1) MB2_MASS(2, 2) MB2_MASS(2, 3) MB2_MASS(2, 4) { ok = false; }
#undef MB2_MASS
  if (!ok) { c->err = "no mass kernel for this (dim, p)"; return MB2_ERR_UNSUPPORTED; }
  MB2_CUDA(cudaGetLastError());
  return MB2_OK;
}

static void free_mass_tab(MassTab& T) {
  cudaFree(T.d_ijk);
  cudaFree(T.d_gw);
  for (int d = 0; d < 3; d++) { cudaFree(T.d_B1[d]); cudaFree(T.d_D1[d]); }
}

// mode 0: A += M on the resident CSR values; mode 1: r += b (body force). el_ijk: knot-span index of the n_compute
// elements this context computes.
static int assemble_form(mb2_ctx* c, int mode, int p, const int32_t* ncp, const double* const* knots, const int32_t* el_ijk,
                         int quad_order, double rho, const double* f) {
  const int dim = c->dim;
  if (p < 1 || p > MAXP) { c->err = "degree must be 1..4"; return MB2_ERR_UNSUPPORTED; }
  int nd = 1;
  for (int d = 0; d < dim; d++) nd *= p + 1;
  if (nd != c->nd) { c->err = "(p+1)^dim does not match nd of the connectivity"; return MB2_ERR_ARG; }
  if (!c->d_xref) { c->err = "call mb2_set_nodes first"; return MB2_ERR_ARG; }
  if (!c->have_pattern) { c->err = "call mb2_build_pattern first"; return MB2_ERR_ARG; }
  const int order = quad_order >= 0 ? quad_order : (mode == 0 ? 2 * p + (p * dim - 1) : mode == 2 ? 2 * p + dim - 1 : 2 * p);
  MassTab T;
  T.p = p;
  T.mode = mode;
  if (f)
    for (int d = 0; d < dim; d++) T.f[d] = f[d];
  T.nq1 = (order | 1) / 2 + 1;
  T.rho = rho;
  if (T.nq1 > MAXQ1) { c->err = "mass quadrature order above the instantiated limit"; return MB2_ERR_UNSUPPORTED; }
  std::vector<double> gx, gw;
  gauss_legendre_01(T.nq1, gx, gw);
  double* d_gx = nullptr;
  int rc = [&]() -> int {
    MB2_CUDA(cudaMalloc(&d_gx, sizeof(double) * T.nq1));
    MB2_CUDA(cudaMalloc(&T.d_gw, sizeof(double) * T.nq1));
    MB2_CUDA(cudaMemcpyAsync(d_gx, gx.data(), sizeof(double) * T.nq1, cudaMemcpyHostToDevice, c->stream));
    MB2_CUDA(cudaMemcpyAsync(T.d_gw, gw.data(), sizeof(double) * T.nq1, cudaMemcpyHostToDevice, c->stream));
    std::vector<double*> d_knot(dim, nullptr);
    DevTmp tmp;
    for (int d = 0; d < dim; d++) tmp.own(&d_knot[d]);
    for (int d = 0; d < dim; d++) {
      const int nk = ncp[d] + p + 1, nspan = ncp[d] - p;
      MB2_CUDA(cudaMalloc(&d_knot[d], sizeof(double) * nk));
      MB2_CUDA(cudaMemcpyAsync(d_knot[d], knots[d], sizeof(double) * nk, cudaMemcpyHostToDevice, c->stream));
      MB2_CUDA(cudaMalloc(&T.d_B1[d], sizeof(double) * nspan * T.nq1 * (p + 1)));
      MB2_CUDA(cudaMalloc(&T.d_D1[d], sizeof(double) * nspan * T.nq1 * (p + 1)));
      k_basis_1d<<<(nspan * T.nq1 + 127) / 128, 128, 0, c->stream>>>(p, nspan, T.nq1, d_knot[d], d_gx, T.d_B1[d], T.d_D1[d]);
      MB2_CUDA(cudaGetLastError());
    }
    MB2_CUDA(cudaMalloc(&T.d_ijk, sizeof(int32_t) * c->n_compute * dim));
    MB2_CUDA(cudaMemcpyAsync(T.d_ijk, el_ijk, sizeof(int32_t) * c->n_compute * dim, cudaMemcpyHostToDevice, c->stream));
    MB2_CUDA(cudaStreamSynchronize(c->stream));
    c->mass_tab = &T;
    int r = mode != 1 ? launch_residual_and_grad(c, 1.0, nullptr, nullptr) : launch_residual(c);
    if (r) return r;
    MB2_CUDA(cudaStreamSynchronize(c->stream));
    return MB2_OK;
  }();
  c->mass_tab = nullptr;
  cudaFree(d_gx);
  free_mass_tab(T);
  return rc;
}

int assemble_mass(mb2_ctx* c, int p, const int32_t* ncp, const double* const* knots, const int32_t* el_ijk, int quad_order,
                  double rho) {
  return assemble_form(c, 0, p, ncp, knots, el_ijk, quad_order, rho, nullptr);
}
int assemble_viscosity(mb2_ctx* c, int p, const int32_t* ncp, const double* const* knots, const int32_t* el_ijk,
                       int quad_order, double coef) {
  return assemble_form(c, 2, p, ncp, knots, el_ijk, quad_order, coef, nullptr);
}
int assemble_body_force(mb2_ctx* c, int p, const int32_t* ncp, const double* const* knots, const int32_t* el_ijk,
                        int quad_order, const double* f) {
  return assemble_form(c, 1, p, ncp, knots, el_ijk, quad_order, 1.0, f);
}
int form_mode(const mb2_ctx* c) { return static_cast<const MassTab*>(c->mass_tab)->mode; }


// ---------------------------------------------------------------------------------------------
// Traction linear form: mfem::VectorBoundaryLFIntegrator::AssembleRHSElementVect (third_party/mfem/fem/lininteg.cpp:375-408)
// as PyNonlinearSolid::Setup adds it to rhs (src/mimi/py/py_nonlinear_solid.cpp:243-276), on face tables the caller
// uploads the way the reference precomputes boundary data (CreateBoundaryElementQuadData + PrecomputeElementQuadData at the
// integrator's rule, order 2p: QuadData::N, integration_weight, det_dX_dxi = Tr.Weight()):
//   per point  vec = t;  vec *= Tr.Weight() * ip.weight;  elvect(ndf k + s) += vec(k) * shape(s)
// LinearForm::Assemble adds the boundary elements after the domain ones, face by face ONTO the vector, so the gather
// continues from the resident r:  r[v] = ((r[v] + f_1) + f_2) + ...  in ascending face order.
// ---------------------------------------------------------------------------------------------
__global__ void k_traction_faces(int64_t NF, int dim, int ndf, int nqf, const double* __restrict__ N,
                                 const double* __restrict__ w, const double* __restrict__ detW, double t0, double t1,
                                 double t2, double* __restrict__ fe) {
  const int tdf = dim * ndf;
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= NF * tdf) return;
  const int64_t f = i / tdf;
  const int l = (int)(i % tdf), k = l / ndf, s = l % ndf;
  const double tk = k == 0 ? t0 : (k == 1 ? t1 : t2);
  double acc = 0.0;
  for (int q = 0; q < nqf; q++) {
    const double vec = tk * (detW[f * nqf + q] * w[f * nqf + q]);
    acc += vec * N[(f * nqf + q) * ndf + s];
  }
  fe[i] = acc;
}

// one thread per loaded vdof: entries (face-major ascending) ptr[j] .. ptr[j+1]
__global__ void k_traction_gather(int64_t n, const int32_t* __restrict__ vdof, const int64_t* __restrict__ ptr,
                                  const int64_t* __restrict__ src, const double* __restrict__ fe, double* __restrict__ r) {
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= n) return;
  double x = r[vdof[j]];
  for (int64_t e = ptr[j]; e < ptr[j + 1]; e++) x += fe[src[e]];
  r[vdof[j]] = x;
}

int assemble_traction(mb2_ctx* c, int64_t NF, int ndf, int nqf, const int32_t* face_vdofs, const double* N, const double* w,
                      const double* detW, const double* traction) {
  const int dim = c->dim, tdf = dim * ndf;
  if (!c->d_r) { c->err = "call mb2_build_pattern first"; return MB2_ERR_ARG; }
  if (NF <= 0 || ndf <= 0 || nqf <= 0) { c->err = "bad face table sizes"; return MB2_ERR_ARG; }
  // vdof -> (face, local) entries in ascending face order (a face's vdofs are distinct)
  std::vector<std::pair<int32_t, int64_t>> ent((size_t)NF * tdf);
  for (int64_t i = 0; i < NF * tdf; i++) {
    const int32_t v = face_vdofs[i];
    if (v < 0 || v >= c->NV) { c->err = "face vdof out of range"; return MB2_ERR_ARG; }
    ent[(size_t)i] = {v, i};
  }
  std::stable_sort(ent.begin(), ent.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
  std::vector<int32_t> vd;
  std::vector<int64_t> ptr, src(ent.size());
  for (size_t i = 0; i < ent.size(); i++) {
    if (i == 0 || ent[i].first != ent[i - 1].first) { vd.push_back(ent[i].first); ptr.push_back((int64_t)i); }
    src[i] = ent[i].second;
  }
  ptr.push_back((int64_t)ent.size());
  const int64_t nv = (int64_t)vd.size();
  double *d_N = nullptr, *d_w = nullptr, *d_det = nullptr, *d_fe = nullptr;
  int32_t* d_vd = nullptr;
  int64_t *d_ptr = nullptr, *d_src = nullptr;
  int rc = [&]() -> int {
    const size_t nfq = (size_t)NF * nqf;
    MB2_CUDA(cudaMalloc(&d_N, sizeof(double) * nfq * ndf));
    MB2_CUDA(cudaMalloc(&d_w, sizeof(double) * nfq));
    MB2_CUDA(cudaMalloc(&d_det, sizeof(double) * nfq));
    MB2_CUDA(cudaMalloc(&d_fe, sizeof(double) * NF * tdf));
    MB2_CUDA(cudaMalloc(&d_vd, sizeof(int32_t) * nv));
    MB2_CUDA(cudaMalloc(&d_ptr, sizeof(int64_t) * (nv + 1)));
    MB2_CUDA(cudaMalloc(&d_src, sizeof(int64_t) * src.size()));
    MB2_CUDA(cudaMemcpyAsync(d_N, N, sizeof(double) * nfq * ndf, cudaMemcpyHostToDevice, c->stream));
    MB2_CUDA(cudaMemcpyAsync(d_w, w, sizeof(double) * nfq, cudaMemcpyHostToDevice, c->stream));
    MB2_CUDA(cudaMemcpyAsync(d_det, detW, sizeof(double) * nfq, cudaMemcpyHostToDevice, c->stream));
    MB2_CUDA(cudaMemcpyAsync(d_vd, vd.data(), sizeof(int32_t) * nv, cudaMemcpyHostToDevice, c->stream));
    MB2_CUDA(cudaMemcpyAsync(d_ptr, ptr.data(), sizeof(int64_t) * (nv + 1), cudaMemcpyHostToDevice, c->stream));
    MB2_CUDA(cudaMemcpyAsync(d_src, src.data(), sizeof(int64_t) * src.size(), cudaMemcpyHostToDevice, c->stream));
    const double t2 = dim == 3 ? traction[2] : 0.0;
    k_traction_faces<<<(unsigned)((NF * tdf + 255) / 256), 256, 0, c->stream>>>(NF, dim, ndf, nqf, d_N, d_w, d_det, traction[0],
                                                                                traction[1], t2, d_fe);
    MB2_CUDA(cudaGetLastError());
    k_traction_gather<<<(unsigned)((nv + 255) / 256), 256, 0, c->stream>>>(nv, d_vd, d_ptr, d_src, d_fe, c->d_r);
    MB2_CUDA(cudaGetLastError());
    MB2_CUDA(cudaStreamSynchronize(c->stream));
    return MB2_OK;
  }();
  for (void* p : {(void*)d_N, (void*)d_w, (void*)d_det, (void*)d_fe, (void*)d_vd, (void*)d_ptr, (void*)d_src})
    if (p) cudaFree(p);
  return rc;
}

}  // namespace mb2
