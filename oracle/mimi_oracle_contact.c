/*
 * TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 *
 * Plain-C, single-thread restatement of the reference's (j042/mimi) rigid-body mortar contact boundary
 * integrator, src/mimi/integrators/mortar_contact.{hpp,cpp}, at nthreads = 1. Used only by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg as the CHECKER of the CUDA path
 * (mimi_b200/csrc/mb2_contact.cu); the product never links, imports or executes this file.
 *
 * Parity status: PINNED for everything downstream of the closest-point query -- tests/test_oracle_contact.py
 * checks every function bit-for-bit against dumps of the unmodified reference (oracle/_ref/mimi_ref contact ...,
 * fixtures tests/golden/contact_*.npz). The rigid bodies are the analytic NearestDistanceBase subclasses of
 * oracle/ref_harness.cpp (closed-form closest point); the reference's own NearestDistanceToSplines delegates the
 * closest-point iteration to splinepy (third_party/splinepy @ f1662da), which is not linked into oracle/_ref:
 * spline rigid bodies are "parity unpinned".
 *
 * Compile without FMA contraction (-ffp-contract=off). All matrices column-major.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_RIGID_PLANE 0  /* p = o[3], t0[3], t1[3], lo, hi */
#define ORC_RIGID_SPHERE 1 /* p = c[3], R */

typedef struct {
  int kind;
  int dim;
  double p[16];
  /* kind == ORC_RIGID_BEZIER (below): degrees, sampling resolution of the initial guess, control net */
  int deg_u, deg_v, n_seed;
  const double* cp;
  /* a B-spline body = its Bezier patches one after the other in cp (0 or 1: a single patch) */
  int n_patches;
  /* 1: rational patches (NURBS bodies), control points homogeneous {w x, w} = dim + 1 doubles each */
  int rational;
} orc_rigid;

static void rigid_bezier_nearest(const orc_rigid* b, const double* q, double* pmq, double* distance, double* fd);

/* NearestDistanceBase::NearestDistance as implemented by RigidPlane / RigidSphere in oracle/ref_harness.cpp,
 * filling what MortarContact reads from Results (src/mimi/coefficients/nearest_distance.hpp:139-193):
 * pmq = physical - query, distance, fd = first_derivatives [para_dim x dim]. */
void orc_rigid_nearest(const orc_rigid* b, const double* q, double* pmq, double* distance, double* fd) {
  const int dim = b->dim;
  if (b->kind == 2) { /* ORC_RIGID_BEZIER */
    rigid_bezier_nearest(b, q, pmq, distance, fd);
    return;
  }
  if (b->kind == ORC_RIGID_PLANE) {
    const double* o = b->p;
    const double* t[2] = {b->p + 3, b->p + 6};
    const double lo = b->p[9], hi = b->p[10];
    double diff[3], s[2] = {0, 0};
    for (int d = 0; d < dim; d++) diff[d] = q[d] - o[d];
    for (int k = 0; k < dim - 1; k++) {
      double num = 0.0, den = 0.0;
      for (int d = 0; d < dim; d++) num += diff[d] * t[k][d];
      for (int d = 0; d < dim; d++) den += t[k][d] * t[k][d];
      double v = num / den;
      if (v < lo) v = lo;
      if (v > hi) v = hi;
      s[k] = v;
    }
    double sum = 0.0;
    for (int d = 0; d < dim; d++) {
      double x = o[d] + s[0] * t[0][d];
      if (dim == 3) x = x + s[1] * t[1][d];
      pmq[d] = x - q[d];
      sum += pmq[d] * pmq[d];
    }
    *distance = sqrt(sum);
    for (int k = 0; k < dim - 1; k++)
      for (int d = 0; d < dim; d++) fd[k * dim + d] = t[k][d];
    return;
  }
  const double* c = b->p;
  const double R = b->p[3];
  double diff[3] = {0, 0, 0}, e[3] = {0, 0, 0};
  double n2 = 0.0;
  for (int d = 0; d < dim; d++) {
    diff[d] = q[d] - c[d];
    n2 += diff[d] * diff[d];
  }
  const double dn = sqrt(n2);
  double sum = 0.0;
  for (int d = 0; d < dim; d++) {
    e[d] = diff[d] / dn;
    const double x = c[d] + R * e[d];
    pmq[d] = x - q[d];
    sum += pmq[d] * pmq[d];
  }
  *distance = sqrt(sum);
  if (dim == 2) {
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

/* Results::ComputeNormal<true> (nearest_distance.hpp:139-181) and ComputeUnitNormal (integrator_utils.hpp:218-251):
 * the same expressions on the rigid body's first derivatives and on the face Jacobian. */
static void unit_normal(int dim, const double* d, double* n) {
  if (dim == 2) {
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

/* mfem::DenseMatrix::Weight for dim x (dim-1), third_party/mfem/linalg/densemat.cpp:559-584 */
static double weight(int dim, const double* d) {
  if (dim == 2) return sqrt(d[0] * d[0] + d[1] * d[1]);
  const double E = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
  const double G = d[3] * d[3] + d[4] * d[4] + d[5] * d[5];
  const double F = d[0] * d[3] + d[1] * d[4] + d[2] * d[5];
  return sqrt(E * G - F * F);
}

/* MortarContactWorkData::ComputeJ = mfem::MultAtB(element_x_mat_, dN_dxi, J) (integrator_utils.cpp:99-104;
 * densemat.cpp:3051-3068): J(i,j) = sum_k x(k,i) dN(k,j), j outer, i inner, k sequential from 0.0 */
static void compute_J(int dim, int ndf, const double* x, const double* dN, double* J) {
  for (int j = 0; j < dim - 1; j++)
    for (int i = 0; i < dim; i++) {
      double d = 0.0;
      for (int k = 0; k < ndf; k++) d += x[k + i * ndf] * dN[k + j * ndf];
      J[i + j * dim] = d;
    }
}

/* MortarContactWorkData::CurrentElementSolutionCopy (integrator_utils.cpp:80-89): x = u[v_dofs] + x_ref */
static void current_x(int tdf, const int32_t* vd, const double* u, const double* xref, double* x) {
  for (int k = 0; k < tdf; k++) x[k] = u[vd[k]];
  for (int k = 0; k < tdf; k++) x[k] += xref[k];
}

/* MortarContact::Prepare (mortar_contact.cpp:19-133): sorted unique marked vdofs -> local (reduced) dof index of
 * every face dof. local_dofs: NF x ndf. Returns the number of marked nodes. */
static int cmp_i32(const void* a, const void* b) {
  const int32_t x = *(const int32_t*)a, y = *(const int32_t*)b;
  return (x > y) - (x < y);
}
int orc_contact_prepare(int dim, int64_t NF, int ndf, const int32_t* v_dofs, int32_t* local_dofs,
                        int32_t* marked_v_dofs /* NF*ndf*dim capacity */, int64_t* n_marked_v) {
  const int tdf = ndf * dim;
  const int64_t n = NF * tdf;
  int32_t* all = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
  memcpy(all, v_dofs, sizeof(int32_t) * (size_t)n);
  qsort(all, (size_t)n, sizeof(int32_t), cmp_i32);
  int64_t m = 0;
  for (int64_t i = 0; i < n; i++)
    if (i == 0 || all[i] != all[i - 1]) all[m++] = all[i];
  const int32_t maxv = all[m - 1];
  int32_t* lmd = (int32_t*)malloc(sizeof(int32_t) * (size_t)((maxv + 1) / dim + 1));
  for (int64_t i = 0; i < (maxv + 1) / dim + 1; i++) lmd[i] = -1;
  int counter = 0;
  for (int64_t vc = 0; vc < m; vc++) {
    if (vc % dim == 0) lmd[all[vc] / dim] = counter++;
    if (marked_v_dofs) marked_v_dofs[vc] = all[vc];
  }
  if (n_marked_v) *n_marked_v = m;
  for (int64_t f = 0; f < NF; f++)
    for (int j = 0; j < ndf; j++) local_dofs[f * ndf + j] = lmd[v_dofs[f * tdf + j] / dim];
  free(all);
  free(lmd);
  return counter;
}

/* MortarContact::ComputePressure + ElementGapAndArea (mortar_contact.cpp:148-261) at nthreads = 1.
 * N: NF x nqf x ndf, dN_dxi: NF x nqf x (ndf x (dim-1)), w: NF x nqf, x_ref: NF x (ndf x dim).
 * Outputs gap / area / pressure per marked node and the total area. */
int orc_contact_pressure(int dim, int64_t NF, int ndf, int nqf, int n_marked, const int32_t* v_dofs,
                         const int32_t* local_dofs, const double* N, const double* dN_dxi, const double* w,
                         const double* x_ref, const double* u, const orc_rigid* body, double penalty, double* gap,
                         double* area, double* pressure, double* area_total) {
  const int tdf = ndf * dim;
  double* x = (double*)malloc(sizeof(double) * (size_t)(tdf + 2 * ndf));
  double* la = x + tdf;
  double* lg = la + ndf;
  double* tl_area = (double*)calloc((size_t)n_marked, sizeof(double));
  double* tl_gap = (double*)calloc((size_t)n_marked, sizeof(double));
  double local_area = 0.0;
  for (int64_t f = 0; f < NF; f++) {
    current_x(tdf, v_dofs + f * tdf, u, x_ref + f * tdf, x);
    for (int i = 0; i < ndf; i++) la[i] = lg[i] = 0.0;
    for (int q = 0; q < nqf; q++) {
      const double* Nq = N + ((size_t)f * nqf + q) * ndf;
      const double* dNq = dN_dxi + ((size_t)f * nqf + q) * ndf * (dim - 1);
      /* ComputeNearestDistanceQuery = DenseMatrix::MultTranspose (densemat.cpp:166-179) */
      double query[3], pmq[3], fd[6], nrm[3], J[6], dist;
      for (int c = 0; c < dim; c++) {
        double y = 0.0;
        for (int r = 0; r < ndf; r++) y += Nq[r] * x[r + c * ndf];
        query[c] = y;
      }
      orc_rigid_nearest(body, query, pmq, &dist, fd);
      unit_normal(dim, fd, nrm);
      /* NormalGap (nearest_distance.hpp:185-192): -normal . (physical - query) */
      double dot = 0.0;
      for (int c = 0; c < dim; c++) dot += pmq[c] * nrm[c];
      const double true_g = -dot;
      double g = (0. < true_g) ? 0. : true_g; /* std::min(true_g, 0.) */
      compute_J(dim, ndf, x, dNq, J);
      const double fac = w[f * nqf + q] * weight(dim, J);
      local_area += fac;
      const double ratio = fabs(true_g) / dist;
      const double cl = (ratio < 1.) ? ratio : 1.; /* std::min(1., ratio) */
      if (acos(cl) > 1.e-5) g = 0.0;
      const double fac_g = fac * g;
      for (int i = 0; i < ndf; i++) {
        la[i] += fac * Nq[i];
        lg[i] += fac_g * Nq[i];
      }
    }
    for (int i = 0; i < ndf; i++) {
      tl_area[local_dofs[f * ndf + i]] += la[i];
      tl_gap[local_dofs[f * ndf + i]] += lg[i];
    }
  }
  *area_total += local_area;
  for (int i = 0; i < n_marked; i++) {
    area[i] = 0.0 + tl_area[i];
    gap[i] = 0.0 + tl_gap[i];
    pressure[i] = gap[i] / area[i] * penalty;
  }
  free(x);
  free(tl_area);
  free(tl_gap);
  return 0;
}

/* MortarContact::ElementResidual<record> (mortar_contact.hpp:99-134). pe: element nodal pressures. */
static void face_residual(int dim, int ndf, int nqf, const double* x, const double* pe, const double* N,
                          const double* dN_dxi, const double* w, double* R, double* force, double* pint) {
  for (int i = 0; i < ndf * dim; i++) R[i] = 0.0;
  for (int q = 0; q < nqf; q++) {
    const double* Nq = N + (size_t)q * ndf;
    double p = 0.0;
    for (int i = 0; i < ndf; i++) p += Nq[i] * pe[i]; /* MortarContactWorkData::Pressure */
    double J[6], nrm[3];
    compute_J(dim, ndf, x, dN_dxi + (size_t)q * ndf * (dim - 1), J);
    const double det_J = weight(dim, J);
    unit_normal(dim, J, nrm);
    const double fac = w[q] * det_J * p;
    /* Ptr_AddMult_a_VWt(-fac, N, normal, R) (integrator_utils.hpp:199-215) */
    double* out = R;
    for (int c = 0; c < dim; c++) {
      const double aw = nrm[c] * (-fac);
      for (int i = 0; i < ndf; i++) (*out++) += aw * Nq[i];
    }
    if (force) {
      for (int c = 0; c < dim; c++) force[c] += fac * nrm[c];
      *pint += fac;
    }
  }
}

/* MortarContact::AddBoundaryResidual / AddBoundaryResidualAndGrad (mortar_contact.cpp:297-421) after
 * ComputePressure, at nthreads = 1: faces ascending, zero-pressure faces skipped, every face pushed straight into
 * the global residual / CSR values. grad == 0: residual only. out3 = {force[3]}, *pint = pressure integral. */
int orc_contact_assemble(int dim, int64_t NF, int ndf, int nqf, const int32_t* v_dofs, const int32_t* local_dofs,
                         const int32_t* A_ids, const double* N, const double* dN_dxi, const double* w,
                         const double* x_ref, const double* u, const double* pressure, int grad, double grad_factor,
                         double* r, double* A, double* force3, double* pint) {
  const int tdf = ndf * dim;
  double* x = (double*)malloc(sizeof(double) * (size_t)(3 * tdf + ndf + (size_t)tdf * tdf));
  double* R = x + tdf;
  double* Rf = R + tdf;
  double* pe = Rf + tdf;
  double* K = pe + ndf;
  for (int c = 0; c < 3; c++) force3[c] = 0.0;
  *pint = 0.0;
  for (int64_t f = 0; f < NF; f++) {
    int zero = 1;
    for (int i = 0; i < ndf; i++) {
      pe[i] = pressure[local_dofs[f * ndf + i]];
      if (pe[i] != 0.0) zero = 0;
    }
    if (zero) continue;
    const double* Nf = N + (size_t)f * nqf * ndf;
    const double* dNf = dN_dxi + (size_t)f * nqf * ndf * (dim - 1);
    const int32_t* vd = v_dofs + f * tdf;
    current_x(tdf, vd, u, x_ref + f * tdf, x);
    face_residual(dim, ndf, nqf, x, pe, Nf, dNf, w + f * nqf, R, force3, pint);
    if (grad) {
      /* MortarContact::ElementResidualAndGrad (mortar_contact.cpp:263-295): the step is taken on the CURRENT
       * POSITION x = u + x_ref, pressure frozen */
      double* g = K;
      for (int i = 0; i < tdf; i++) {
        const double orig = x[i];
        const double h = (orig != 0.0) ? fabs(orig) * 1.0e-8 : 1.0e-10;
        const double hinv = 1. / h;
        x[i] = orig + h;
        face_residual(dim, ndf, nqf, x, pe, Nf, dNf, w + f * nqf, Rf, 0, 0);
        for (int j = 0; j < tdf; j++) *g++ = (Rf[j] - R[j]) * hinv;
        x[i] = orig;
      }
    }
    for (int k = 0; k < tdf; k++) r[vd[k]] += R[k];
    if (grad) {
      const int32_t* aid = A_ids + f * tdf * tdf;
      for (int k = 0; k < tdf * tdf; k++) A[aid[k]] += K[k] * grad_factor;
    }
  }
  free(x);
  return 0;
}

/* MortarContact::GapNorm (mortar_contact.cpp:423-467) */
double orc_contact_gap_norm(int dim, int64_t NF, int ndf, int nqf, const int32_t* v_dofs, const double* N,
                            const double* x_ref, const double* u, const orc_rigid* body) {
  const int tdf = ndf * dim;
  double* x = (double*)malloc(sizeof(double) * (size_t)tdf);
  double sum = 0.0;
  for (int64_t f = 0; f < NF; f++) {
    current_x(tdf, v_dofs + f * tdf, u, x_ref + f * tdf, x);
    for (int q = 0; q < nqf; q++) {
      const double* Nq = N + ((size_t)f * nqf + q) * ndf;
      double query[3], pmq[3], fd[6], nrm[3], dist;
      for (int c = 0; c < dim; c++) {
        double y = 0.0;
        for (int r = 0; r < ndf; r++) y += Nq[r] * x[r + c * ndf];
        query[c] = y;
      }
      orc_rigid_nearest(body, query, pmq, &dist, fd);
      unit_normal(dim, fd, nrm);
      double dot = 0.0;
      for (int c = 0; c < dim; c++) dot += pmq[c] * nrm[c];
      const double g = -dot;
      if (g < 0.0) sum += g * g;
    }
  }
  free(x);
  return sqrt(sum);
}

/* ---------------------------------------------------------------------------------------------------------------
 * Bezier rigid bodies (curve in 2D, tensor-product surface in 3D) -- PARITY UNPINNED.
 * The reference's NearestDistanceToSplines hands the closest-point search to splinepy (third_party/splinepy @ f1662da:
 * k-d tree of sampled points for the initial guess, Newton with bound clipping, SLSQP fallback; tolerance 1e-24, i.e.
 * iterate to the end), which is not linked into oracle/_ref. What is restated here is the published scheme -- sampled
 * initial guess + bound-clipped Newton on J = |S(xi) - q|^2 / 2 with the exact Hessian, Gauss-Newton when that is
 * not positive definite -- in a fixed operation order that mb2_contact.cu (rigid_nearest_bezier) repeats bit for bit.
 * A converged search returns the same closest point as splinepy to rounding (checked in tests by first-order
 * optimality and against dense sampling), not the same bits.
 * cp layout: [deg_v + 1][deg_u + 1][dim], u fastest; curves have deg_v = 0.
 * --------------------------------------------------------------------------------------------------------------- */
#define ORC_RIGID_BEZIER 2
#define ORC_BEZ_MAXDEG 5

typedef struct {
  int dim, deg_u, deg_v, n_seed;
  const double* cp;
  int rational; /* control points are homogeneous {w x, w}: dim + 1 doubles each */
} orc_bezier;

/* Bernstein basis of degree p at t with first and second derivatives (recurrence on the lower-degree bases) */
static void bernstein(int p, double t, double* B, double* dB, double* ddB) {
  double lvl[ORC_BEZ_MAXDEG + 1][ORC_BEZ_MAXDEG + 1];
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

/* S, S_u, S_v, S_uu, S_uv, S_vv at (u, v): out[6][3]. Rational patches (splinepy's RationalBezier / NURBS rigid
 * bodies, src/mimi/coefficients/nearest_distance.hpp:139-193 takes either): numerator A = sum B w P and denominator
 * W = sum B w with their derivatives first, then the quotient rule in the order written below. */
static void bezier_eval(const orc_bezier* b, double u, double v, double out[6][3]) {
  double Bu[ORC_BEZ_MAXDEG + 1], dBu[ORC_BEZ_MAXDEG + 1], ddBu[ORC_BEZ_MAXDEG + 1];
  double Bv[ORC_BEZ_MAXDEG + 1], dBv[ORC_BEZ_MAXDEG + 1], ddBv[ORC_BEZ_MAXDEG + 1];
  const int stride = b->dim + (b->rational ? 1 : 0);
  double W[6] = {0, 0, 0, 0, 0, 0};
  bernstein(b->deg_u, u, Bu, dBu, ddBu);
  bernstein(b->deg_v, v, Bv, dBv, ddBv);
  for (int k = 0; k < 6; k++)
    for (int d = 0; d < 3; d++) out[k][d] = 0.0;
  for (int j = 0; j <= b->deg_v; j++)
    for (int i = 0; i <= b->deg_u; i++) {
      const double* P = b->cp + ((size_t)j * (b->deg_u + 1) + i) * stride;
      const double w[6] = {Bu[i] * Bv[j], dBu[i] * Bv[j], Bu[i] * dBv[j], ddBu[i] * Bv[j], dBu[i] * dBv[j], Bu[i] * ddBv[j]};
      for (int k = 0; k < 6; k++)
        for (int d = 0; d < b->dim; d++) out[k][d] += w[k] * P[d];
      if (b->rational)
        for (int k = 0; k < 6; k++) W[k] += w[k] * P[b->dim];
    }
  if (b->rational)
    for (int d = 0; d < b->dim; d++) {
      const double S = out[0][d] / W[0];
      const double Su = (out[1][d] - W[1] * S) / W[0];
      const double Sv = (out[2][d] - W[2] * S) / W[0];
      const double Suu = ((out[3][d] - (2.0 * W[1]) * Su) - W[3] * S) / W[0];
      const double Suv = (((out[4][d] - W[1] * Sv) - W[2] * Su) - W[4] * S) / W[0];
      const double Svv = ((out[5][d] - (2.0 * W[2]) * Sv) - W[5] * S) / W[0];
      out[0][d] = S; out[1][d] = Su; out[2][d] = Sv; out[3][d] = Suu; out[4][d] = Suv; out[5][d] = Svv;
    }
}

void orc_bezier_nearest(const orc_bezier* b, const double* q, double* xi, double* pmq, double* distance, double* fd) {
  const int dim = b->dim, pd = dim - 1, ns = b->n_seed;
  double E[6][3];
  /* initial guess: nearest point of a uniform (ns+1)^pd sampling */
  double best = 1e300, u = 0.0, v = 0.0;
  for (int jv = 0; jv <= (pd == 2 ? ns : 0); jv++)
    for (int iu = 0; iu <= ns; iu++) {
      const double su = (double)iu / (double)ns, sv = (pd == 2) ? (double)jv / (double)ns : 0.0;
      bezier_eval(b, su, sv, E);
      double d2 = 0.0;
      for (int d = 0; d < dim; d++) d2 += (E[0][d] - q[d]) * (E[0][d] - q[d]);
      if (d2 < best) {
        best = d2;
        u = su;
        v = sv;
      }
    }
  for (int it = 0; it < 40; it++) {
    bezier_eval(b, u, v, E);
    double r[3] = {0, 0, 0};
    for (int d = 0; d < dim; d++) r[d] = E[0][d] - q[d];
    double g0 = 0.0, g1 = 0.0, a00 = 0.0, a01 = 0.0, a11 = 0.0, h00 = 0.0, h01 = 0.0, h11 = 0.0;
    for (int d = 0; d < dim; d++) {
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
    if (pd == 1) {
      double H = a00 + h00;
      if (!(H > 0.0)) H = a00; /* Gauss-Newton when the exact Hessian is not positive */
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
  bezier_eval(b, u, v, E);
  double sum = 0.0;
  for (int d = 0; d < dim; d++) {
    pmq[d] = E[0][d] - q[d];
    sum += pmq[d] * pmq[d];
  }
  *distance = sqrt(sum);
  for (int k = 0; k < pd; k++)
    for (int d = 0; d < dim; d++) fd[k * dim + d] = E[1 + k][d];
  if (xi) {
    xi[0] = u;
    if (pd == 2) xi[1] = v;
  }
}

/* Multi-patch bodies (B-splines with interior knots, split into Bezier patches by the caller): the closest point of
 * every patch, the nearest wins, the first on ties -- mb2_contact.cu (rigid_nearest_bezier) does the same. */
static void rigid_bezier_nearest(const orc_rigid* b, const double* q, double* pmq, double* distance, double* fd) {
  const int np = b->n_patches > 1 ? b->n_patches : 1;
  const size_t stride = (size_t)(b->deg_u + 1) * (b->deg_v + 1) * (b->dim + (b->rational ? 1 : 0));
  double best = 1e300;
  for (int p = 0; p < np; p++) {
    const orc_bezier z = {b->dim, b->deg_u, b->deg_v, b->n_seed, b->cp + p * stride, b->rational};
    double pm[3], dist, f[6];
    orc_bezier_nearest(&z, q, 0, pm, &dist, f);
    if (dist < best) {
      best = dist;
      for (int d = 0; d < b->dim; d++) pmq[d] = pm[d];
      for (int k = 0; k < (b->dim - 1) * b->dim; k++) fd[k] = f[k];
    }
  }
  *distance = best;
}

/* test helper: closest point only (xi out) */
void orc_rigid_query(const orc_rigid* b, const double* q, double* xi, double* pmq, double* distance, double* fd) {
  if (b->kind == ORC_RIGID_BEZIER && b->n_patches <= 1) {
    const orc_bezier z = {b->dim, b->deg_u, b->deg_v, b->n_seed, b->cp, b->rational};
    orc_bezier_nearest(&z, q, xi, pmq, distance, fd);
  } else {
    orc_rigid_nearest(b, q, pmq, distance, fd);
  }
}
