/* TEST INFRASTRUCTURE ONLY -- part of the CPU oracle (see mimi_oracle.c). Never linked into the product.
 *
 * The linear-algebra forms PyNonlinearSolid::Setup assembles once with MFEM (src/mimi/py/py_nonlinear_solid.cpp:154-191,
 * 221-241), restated for ONE tensor-product NURBS patch so that every value is produced by MFEM's operation sequence:
 *   mass       mfem::VectorMassIntegrator::AssembleElementMatrix       third_party/mfem/fem/bilininteg.cpp:1542-1620
 *   viscosity  mfem::VectorDiffusionIntegrator::AssembleElementMatrix  third_party/mfem/fem/bilininteg.cpp:2854-2934
 *   body force mfem::VectorDomainLFIntegrator::AssembleRHSElementVect  third_party/mfem/fem/lininteg.cpp:268-308
 * with the pieces they call:
 *   1D Gauss-Legendre rule, tensor ordering     fem/intrules.cpp:424-494, 56-84
 *   KnotVector::CalcShape / CalcDShape          mesh/nurbs.cpp:297-376
 *   NURBS{2,3}DFiniteElement::CalcShape/CalcDShape   fem/fe/fe_nurbs.cpp:125-311
 *   IsoparametricTransformation Jacobian / Weight / AdjugateJacobian, OrderW   fem/eltrans.cpp:406-465
 *   CalcAdjugate, Mult, Mult_a_AAt, MultVVt     linalg/densemat.cpp:2495-2563, 3196-3222
 * Pinned against dumps of the unmodified reference: the `mass` array of every tests/golden fixture and
 * tests/golden/forms.npz (oracle/make_golden.py bodyforce), see tests/test_oracle_forms.py.
 *
 * The routines return ELEMENT blocks (scalar nd x nd matrices / dim*nd vectors); BilinearForm::Assemble's ordered
 * addition into the CSR is done by the caller (oracle_py.assemble_form). */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define FMAXP 4
#define FMAXND 64

static void gauss01(int np, double* x, double* w) {
  if (np == 1) { x[0] = 0.5; w[0] = 1.0; return; }
  if (np == 2) {
    x[0] = 0.21132486540518711775; w[0] = 0.5;
    x[1] = 0.78867513459481288225; w[1] = 0.5;
    return;
  }
  if (np == 3) {
    x[0] = 0.11270166537925831148; w[0] = 5. / 18.;
    x[1] = 0.5; w[1] = 4. / 9.;
    x[2] = 0.88729833462074168852; w[2] = 5. / 18.;
    return;
  }
  const int n = np, m = (n + 1) / 2;
  for (int i = 1; i <= m; i++) {
    double z = cos(M_PI * (i - 0.25) / (n + 0.5));
    double pp, p1, dz, xi = 0.;
    int done = 0;
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
      if (fabs(dz) < 1e-16) {
        done = 1;
        xi = ((1 - z) + dz) / 2;
      }
      z -= dz;
    }
    x[i - 1] = xi;
    x[n - i] = 1 - xi;
    w[i - 1] = w[n - i] = 1. / (4 * xi * (1 - xi) * pp * pp);
  }
}

/* KnotVector::CalcShape and CalcDShape at local coordinate xi of span i (ip = i + p) */
static void basis_1d(int p, const double* knot, int span, double xi, double* shape, double* grad) {
  const int ip = span + p;
  const double u = (xi * knot[ip + 1] + (1. - xi) * knot[ip]);
  double left[FMAXP + 1], right[FMAXP + 1];
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
  double ndu[FMAXP + 1][FMAXP + 1];
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
    grad[r] = d * scale;
  }
}

typedef struct {
  int dim, p, nd, nq1, nq;
  double gx[16], gw[16];
  const double* knots[3];
  const double* x_ref;   /* NV byVDIM */
  const double* wts;     /* per node or NULL */
} forms_ctx;

/* shape values, dshapedxt = dshape adj(J) (nd x dim column-major), ip.weight and Trans.Weight() at point q of an element */
static void point(const forms_ctx* f, const int32_t* vd, const int32_t* ijk, int q, double* shape, double* gxt, double* ipw,
                  double* detJ) {
  const int dim = f->dim, p1 = f->p + 1, nd = f->nd, nq1 = f->nq1;
  int qd[3] = {q % nq1, (q / nq1) % nq1, dim == 3 ? q / (nq1 * nq1) : 0};
  double s1[3][FMAXP + 1], d1[3][FMAXP + 1];
  for (int d = 0; d < dim; d++) basis_1d(f->p, f->knots[d], ijk[d], f->gx[qd[d]], s1[d], d1[d]);
  double uu[FMAXND], ds[3 * FMAXND];
  double sum = 0.0, dsum[3] = {0.0, 0.0, 0.0};
  int o = 0;
  if (dim == 3) {
    for (int k = 0; k < p1; k++)
      for (int j = 0; j < p1; j++) {
        const double sy_sz = s1[1][j] * s1[2][k], dsy_sz = d1[1][j] * s1[2][k], sy_dsz = s1[1][j] * d1[2][k];
        for (int i = 0; i < p1; i++, o++) {
          const double w = f->wts ? f->wts[vd[o] / dim] : 1.0;
          sum += (uu[o] = s1[0][i] * sy_sz * w);
          dsum[0] += (ds[o] = d1[0][i] * sy_sz * w);
          dsum[1] += (ds[o + nd] = s1[0][i] * dsy_sz * w);
          dsum[2] += (ds[o + 2 * nd] = s1[0][i] * sy_dsz * w);
        }
      }
  } else {
    for (int j = 0; j < p1; j++)
      for (int i = 0; i < p1; i++, o++) {
        const double w = f->wts ? f->wts[vd[o] / dim] : 1.0;
        sum += (uu[o] = s1[0][i] * s1[1][j] * w);
        dsum[0] += (ds[o] = d1[0][i] * s1[1][j] * w);
        dsum[1] += (ds[o + nd] = s1[0][i] * d1[1][j] * w);
      }
  }
  const double m = 1.0 / sum; /* shape /= sum */
  for (o = 0; o < nd; o++) shape[o] = uu[o] * m;
  for (int d = 0; d < dim; d++) dsum[d] *= m * m;
  for (o = 0; o < nd; o++)
    for (int d = 0; d < dim; d++) ds[o + d * nd] = ds[o + d * nd] * m - uu[o] * dsum[d];
  double J[9];
  for (int i = 0; i < dim * dim; i++) J[i] = 0.0;
  for (int j = 0; j < dim; j++)
    for (int k = 0; k < nd; k++) {
      const double c = ds[k + j * nd];
      for (int i = 0; i < dim; i++) J[i + dim * j] += f->x_ref[vd[i * nd + k]] * c;
    }
#define A_(i, j) J[(i) + dim * (j)]
  double ad[9];
#define AD_(i, j) ad[(i) + dim * (j)]
  if (dim == 2) {
    *detJ = A_(0, 0) * A_(1, 1) - A_(1, 0) * A_(0, 1);
    AD_(0, 0) = A_(1, 1); AD_(0, 1) = -A_(0, 1); AD_(1, 0) = -A_(1, 0); AD_(1, 1) = A_(0, 0);
  } else {
    /* DenseMatrix::Det, 3 x 3 (densemat.cpp:510-525) */
    const double* d = J;
    *detJ = d[0] * (d[4] * d[8] - d[5] * d[7]) + d[3] * (d[2] * d[7] - d[1] * d[8]) + d[6] * (d[1] * d[5] - d[2] * d[4]);
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
  for (int j = 0; j < dim; j++)
    for (int i = 0; i < nd; i++) {
      double a = 0.0;
      for (int k = 0; k < dim; k++) a += ds[i + k * nd] * AD_(k, j);
      gxt[i + j * nd] = a;
    }
#undef A_
#undef AD_
  double w = f->gw[qd[0]] * f->gw[qd[1]];
  if (dim == 3) w = w * f->gw[qd[2]];
  *ipw = w;
}

static int forms_init(forms_ctx* f, int dim, int p, int order, const double* const* knots, const double* x_ref,
                      const double* wts) {
  f->dim = dim; f->p = p;
  f->nd = 1;
  for (int d = 0; d < dim; d++) f->nd *= p + 1;
  f->nq1 = (order | 1) / 2 + 1; /* IntegrationRules::Get: real order = order | 1, n = order / 2 + 1 */
  if (p > FMAXP || f->nd > FMAXND || f->nq1 > 16) return 1;
  f->nq = 1;
  for (int d = 0; d < dim; d++) f->nq *= f->nq1;
  gauss01(f->nq1, f->gx, f->gw);
  for (int d = 0; d < dim; d++) f->knots[d] = knots[d];
  f->x_ref = x_ref; f->wts = wts;
  return 0;
}

/* kind 0: mass, coefficient rho, rule 2p + (p dim - 1); kind 2: viscosity, coefficient nu, rule 2p + dim - 1.
 * out: NE x nd x nd scalar blocks (row-major; the blocks are symmetric); quad_order < 0 = MFEM's default. */
int orc_form_element_matrices(int kind, int dim, int p, int64_t NE, const int32_t* v_dofs, const int32_t* el_ijk,
                              const double* const* knots, const double* x_ref, const double* wts, int quad_order,
                              double coef, double* out) {
  forms_ctx f;
  const int order = quad_order >= 0 ? quad_order : (kind == 0 ? 2 * p + (p * dim - 1) : 2 * p + dim - 1);
  if (forms_init(&f, dim, p, order, knots, x_ref, wts)) return 1;
  const int nd = f.nd, td = dim * nd;
  double shape[FMAXND], gxt[3 * FMAXND];
  for (int64_t e = 0; e < NE; e++) {
    double* M = out + e * nd * nd;
    memset(M, 0, sizeof(double) * nd * nd);
    for (int q = 0; q < f.nq; q++) {
      double w, dj;
      point(&f, v_dofs + e * td, el_ijk + e * dim, q, shape, gxt, &w, &dj);
      if (kind == 0) {
        const double norm = (w * dj) * coef; /* norm = ip.weight * Trans.Weight(); norm *= Q->Eval */
        for (int i = 0; i < nd; i++)
          for (int j = 0; j < nd; j++) M[i * nd + j] += (shape[i] * shape[j]) * norm; /* MultVVt; *= norm; AddMatrix */
      } else {
        const double ww = (w / dj) * coef; /* w = ip.weight / Trans.Weight(); w *= Q->Eval */
        for (int i = 0; i < nd; i++)
          for (int j = 0; j < nd; j++) {
            double d = 0.;
            for (int k = 0; k < dim; k++) d += gxt[i + k * nd] * gxt[j + k * nd];
            M[i * nd + j] += ww * d; /* Mult_a_AAt; AddMatrix */
          }
      }
    }
  }
  return 0;
}

/* body force: out NE x (dim*nd), component-major; rule 2p */
int orc_form_element_vectors(int dim, int p, int64_t NE, const int32_t* v_dofs, const int32_t* el_ijk,
                             const double* const* knots, const double* x_ref, const double* wts, int quad_order,
                             const double* force, double* out) {
  forms_ctx f;
  if (forms_init(&f, dim, p, quad_order >= 0 ? quad_order : 2 * p, knots, x_ref, wts)) return 1;
  const int nd = f.nd, td = dim * nd;
  double shape[FMAXND], gxt[3 * FMAXND];
  for (int64_t e = 0; e < NE; e++) {
    double* b = out + e * td;
    memset(b, 0, sizeof(double) * td);
    for (int q = 0; q < f.nq; q++) {
      double w, val;
      point(&f, v_dofs + e * td, el_ijk + e * dim, q, shape, gxt, &w, &val);
      for (int k = 0; k < dim; k++) {
        const double cf = val * force[k];
        for (int s = 0; s < nd; s++) b[nd * k + s] += w * cf * shape[s];
      }
    }
  }
  return 0;
}
