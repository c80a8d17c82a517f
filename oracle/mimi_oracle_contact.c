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
} orc_rigid;

/* NearestDistanceBase::NearestDistance as implemented by RigidPlane / RigidSphere in oracle/ref_harness.cpp,
 * filling what MortarContact reads from Results (src/mimi/coefficients/nearest_distance.hpp:139-193):
 * pmq = physical - query, distance, fd = first_derivatives [para_dim x dim]. */
void orc_rigid_nearest(const orc_rigid* b, const double* q, double* pmq, double* distance, double* fd) {
  const int dim = b->dim;
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
