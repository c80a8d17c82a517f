/* TEST INFRASTRUCTURE ONLY -- part of the CPU oracle (see mimi_oracle.c). Never linked into the product.
 *
 * Spectrum of a symmetric 2x2 / 3x3 matrix exactly as the reference's J2Log material obtains it:
 * LogarithmicStrain (src/mimi/materials/material_utils.hpp:92-114) and the exponential map of J2Log::PlasticStress
 * (src/mimi/materials/materials.hpp:636-750) call mfem::DenseMatrix::CalcEigenvalues
 * (third_party/mfem/linalg/densemat.cpp:1309-1329), i.e. kernels::CalcEigenvalues<2|3>
 * (third_party/mfem/linalg/kernels.hpp:1110-1340) with its helpers (kernels.hpp:441-1065). MFEM is vendored in
 * the reference at 0952809 (v4.7.1).
 *
 * The algorithm (restated, operation order preserved so that the result is bit-identical on the same libm):
 *   3x3: scale by 2^e, closed-form extreme root of the characteristic cubic (trigonometric formula), a unit vector z
 *        in the near-kernel of A - root*I by a column/row-pivoted Householder QR, deflation Q P A P Q with the reflection
 *        that maps P z onto e_1, a Jacobi rotation of the remaining 2x2 block, back-transformation of its two
 *        eigenvectors, ascending sort.
 *   2x2: one Jacobi rotation.
 * Only the reference's `mode == 0` pivoting (smallest entry, kernels.hpp:1226) is reachable and restated.
 * Pinned through the J2Log fixtures dumped from the unmodified reference (tests/golden/j2_*.npz, keys *_log_*). */
#include <float.h>
#include <math.h>

#define ORC_PI 3.14159265358979323846 /* M_PI */

static void swapd(double* a, double* b) {
  const double t = *a;
  *a = *b;
  *b = t;
}

/* Jacobi rotation of [[a, b], [b, d]] (kernels.hpp:463-494): a, d become the eigenvalues; (c, s) the rotation */
static void jacobi2(double b, double* a, double* d, double* c, double* s) {
  const double big = sqrt(1. / DBL_EPSILON);
  if (b == 0.0) {
    *c = 1.;
    *s = 0.;
    return;
  }
  double t;
  const double zeta = (*d - *a) / (2 * b);
  const double az = fabs(zeta);
  if (az < big) t = copysign(1. / (az + sqrt(1. + zeta * zeta)), zeta);
  else t = copysign(0.5 / az, zeta);
  *c = sqrt(1. / (1. + t * t));
  *s = *c * t;
  t *= b;
  *a -= t;
  *d += t;
}

/* 2^e with vmax / 2^e in [0.5, 1) (kernels.hpp:497-517) */
static double pow2_scale(double vmax) {
  if (!(vmax > 0.)) return 1.;
  int e;
  double m = frexp(vmax, &e);
  if (e == DBL_MAX_EXP) m *= FLT_RADIX;
  return vmax / m;
}

/* near-kernel vector of the general 2x2 [[g[0], g[1]], [g[2], g[3]]] (row-major), |z1| + |z2| = 1, returned in
 * (g[0], g[3]); 1 = zero matrix (kernels.hpp:519-659, mode 0) */
static int null2(double* g) {
  double* p11 = &g[0];
  double* p12 = &g[1];
  double* p21 = &g[2];
  double* p22 = &g[3];
  double n1 = fabs(*p11) + fabs(*p21);
  double n2 = fabs(*p22) + fabs(*p12);
  const int swap_cols = n2 > n1;
  double mu;
  if (!swap_cols) {
    if (n1 == 0.) return 1;
    if (fabs(*p11) > fabs(*p21)) {
      swapd(p11, p21);
      swapd(p12, p22);
    }
  } else if (fabs(*p12) > fabs(*p22)) {
    swapd(p11, p22);
    swapd(p12, p21);
  } else {
    swapd(p11, p12);
    swapd(p21, p22);
  }
  n1 = hypot(*p11, *p21);
  if (*p21 != 0.) {
    mu = copysign(n1, *p11);
    n1 = -*p21 * (*p21 / (*p11 + mu));
    *p11 = mu;
    if (fabs(n1) <= fabs(*p21)) {
      n1 = n1 / *p21;
      mu = (2. / (1. + n1 * n1)) * (n1 * *p12 + *p22);
      *p22 = *p22 - mu;
      *p12 = *p12 - mu * n1;
    } else {
      n2 = *p21 / n1;
      mu = (2. / (1. + n2 * n2)) * (*p12 + n2 * *p22);
      *p22 = *p22 - mu * n2;
      *p12 = *p12 - mu;
    }
  }
  mu = -*p12 / *p11;
  n2 = 1. / (1. + fabs(mu));
  if (fabs(*p11) <= n2 * fabs(*p22)) {
    *p22 = 0.;
    *p11 = 1.;
  } else {
    *p22 = n2;
    *p11 = mu * n2;
  }
  if (swap_cols) swapd(p11, p22);
  return 0;
}

/* (kernels.hpp:662-678): x1 is the entry of largest magnitude */
static void unit3_lead(double x1, double x2, double x3, double* n1, double* n2, double* n3) {
  const double m = fabs(x1);
  double r = x2 / m;
  double t = 1. + r * r;
  r = x3 / m;
  t = sqrt(1. / (t + r * r));
  *n1 = copysign(t, x1);
  t /= m;
  *n2 = x2 * t;
  *n3 = x3 * t;
}

/* v <- v / |v| without overflow (kernels.hpp:681-707) */
static void unit3(double* v) {
  const double x1 = v[0], x2 = v[1], x3 = v[2];
  if (fabs(x1) >= fabs(x2)) {
    if (fabs(x1) >= fabs(x3)) {
      if (x1 != 0.) unit3_lead(x1, x2, x3, &v[0], &v[1], &v[2]);
      else v[0] = v[1] = v[2] = 0.;
      return;
    }
  } else if (fabs(x2) >= fabs(x3)) {
    unit3_lead(x2, x1, x3, &v[1], &v[0], &v[2]);
    return;
  }
  unit3_lead(x3, x1, x2, &v[2], &v[0], &v[1]);
}

/* unit near-kernel vector of a general 3x3 m (row-major m[3*r + c]) whose first column is already the pivot column:
 * Householder on column 1, then null2 on the trailing block (kernels.hpp:709-814). z out; returns the kernel
 * dimension (1, or 2: z is then orthogonal to the kernel). */
static int null3_general(double* m, double* z) {
  double *d1 = &m[0], *c12 = &m[1], *c13 = &m[2];
  double *c21 = &m[3], *d2 = &m[4], *c23 = &m[5];
  double *c31 = &m[6], *c32 = &m[7], *d3 = &m[8];
  int kdim;
  double mu, n1, n2, n3, s1, s2, s3;
  s1 = hypot(*c21, *c31);
  n1 = hypot(*d1, s1);
  if (s1 != 0.) {
    mu = copysign(n1, *d1);
    n1 = -s1 * (s1 / (*d1 + mu));
    *d1 = mu;
    const int n1_ge_c21 = fabs(n1) >= fabs(*c21);
    if (n1_ge_c21 && fabs(n1) >= fabs(*c31)) {
      s2 = *c21 / n1;
      s3 = *c31 / n1;
      mu = 2. / (1. + s2 * s2 + s3 * s3);
      n2 = mu * (*c12 + s2 * *d2 + s3 * *c32);
      n3 = mu * (*c13 + s2 * *c23 + s3 * *d3);
      *c12 = *c12 - n2;
      *d2 = *d2 - s2 * n2;
      *c32 = *c32 - s3 * n2;
      *c13 = *c13 - n3;
      *c23 = *c23 - s2 * n3;
      *d3 = *d3 - s3 * n3;
    } else if (!n1_ge_c21 && fabs(*c21) >= fabs(*c31)) {
      s1 = n1 / *c21;
      s3 = *c31 / *c21;
      mu = 2. / (1. + s1 * s1 + s3 * s3);
      n2 = mu * (s1 * *c12 + *d2 + s3 * *c32);
      n3 = mu * (s1 * *c13 + *c23 + s3 * *d3);
      *c12 = *c12 - s1 * n2;
      *d2 = *d2 - n2;
      *c32 = *c32 - s3 * n2;
      *c13 = *c13 - s1 * n3;
      *c23 = *c23 - n3;
      *d3 = *d3 - s3 * n3;
    } else {
      s1 = n1 / *c31;
      s2 = *c21 / *c31;
      mu = 2. / (1. + s1 * s1 + s2 * s2);
      n2 = mu * (s1 * *c12 + s2 * *d2 + *c32);
      n3 = mu * (s1 * *c13 + s2 * *c23 + *d3);
      *c12 = *c12 - s1 * n2;
      *d2 = *d2 - s2 * n2;
      *c32 = *c32 - n2;
      *c13 = *c13 - s1 * n3;
      *c23 = *c23 - s2 * n3;
      *d3 = *d3 - n3;
    }
  }
  double g[4] = {*d2, *c23, *c32, *d3};
  if (null2(g)) {
    z[1] = *c12 / *d1;
    z[2] = *c13 / *d1;
    z[0] = 1.;
    kdim = 2;
  } else {
    z[1] = g[0];
    z[2] = g[3];
    z[0] = -(*c12 * z[1] + *c13 * z[2]) / *d1;
    kdim = 1;
  }
  unit3(z);
  return kdim;
}

/* unit near-kernel vector of the symmetric [[d[0], o12, o13], [o12, d[1], o23], [o13, o23, d[2]]]: pivot the column of
 * largest l1 norm to the front, the row with the smallest pivot entry to the top (mode 0), then null3_general
 * (kernels.hpp:816-942). z overwrites d. Returns 3 for the zero matrix. */
static int null3_sym(double o12, double o13, double o23, double* d) {
  double c12 = o12, c13 = o13, c23 = o23;
  double c21, c31, c32;
  int col, row;
  c32 = fabs(d[0]) + fabs(c12) + fabs(c13);
  c31 = fabs(d[1]) + fabs(c12) + fabs(c23);
  c21 = fabs(d[2]) + fabs(c13) + fabs(c23);
  if (c32 >= c21) col = (c32 >= c31) ? 1 : 2;
  else col = (c31 >= c21) ? 2 : 3;
  if (col == 1) {
    if (c32 == 0.) return 3;
  } else if (col == 2) {
    if (c31 == 0.) return 3;
    swapd(&c13, &c23);
    swapd(&d[0], &d[1]);
  } else {
    if (c21 == 0.) return 3;
    swapd(&c12, &c23);
    swapd(&d[0], &d[2]);
  }
  if (fabs(d[0]) <= fabs(c13)) row = (fabs(d[0]) <= fabs(c12)) ? 1 : 2;
  else row = (fabs(c12) <= fabs(c13)) ? 2 : 3;
  double m[9];
  if (row == 1) {
    m[0] = d[0]; m[1] = c12;  m[2] = c13;
    m[3] = c12;  m[4] = d[1]; m[5] = c23;
    m[6] = c13;  m[7] = c23;  m[8] = d[2];
  } else if (row == 2) { /* rows 1 <-> 2 */
    m[0] = c12;  m[1] = d[1]; m[2] = c23;
    m[3] = d[0]; m[4] = c12;  m[5] = c13;
    m[6] = c13;  m[7] = c23;  m[8] = d[2];
  } else { /* rows 1 <-> 3 */
    m[0] = c13;  m[1] = c23;  m[2] = d[2];
    m[3] = c12;  m[4] = d[1]; m[5] = c23;
    m[6] = d[0]; m[7] = c12;  m[8] = c13;
  }
  const int kdim = null3_general(m, d);
  if (col == 2) swapd(&d[0], &d[1]);
  else if (col == 3) swapd(&d[0], &d[2]);
  return kdim;
}

/* deflation (kernels.hpp:944-1065, mode 0): with the unit eigenvector z, B = Q P A P Q = diag(b1, [[b2, b23], [b23, b3]]),
 * P swapping entries 1 and k (k: |z_k| smallest), Q = I - g v v^T mapping P z to e_1. d = diagonal, o = {a12, a13, a23};
 * b1, b2, b3 -> d, b23 -> o[2]. Returns k (1-based). */
static int deflate3(double* d, double* o, double* z, double* v, double* g) {
  int k;
  double s, w1, w2, w3;
  if (fabs(z[0]) <= fabs(z[2])) k = (fabs(z[0]) <= fabs(z[1])) ? 1 : 2;
  else k = (fabs(z[1]) <= fabs(z[2])) ? 2 : 3;
  if (k == 2) {
    swapd(&o[1], &o[2]);
    swapd(&d[0], &d[1]);
    swapd(&z[0], &z[1]);
  } else if (k == 3) {
    swapd(&o[0], &o[2]);
    swapd(&d[0], &d[2]);
    swapd(&z[0], &z[2]);
  }
  s = hypot(z[1], z[2]);
  if (s == 0.) {
    v[0] = v[1] = v[2] = 0.;
    *g = 1.;
  } else {
    double gg = copysign(1., z[0]);
    v[0] = -s * (s / (z[0] + gg));
    gg = fabs(v[0]);
    if (fabs(z[1]) > gg) gg = fabs(z[1]);
    if (fabs(z[2]) > gg) gg = fabs(z[2]);
    v[0] = v[0] / gg;
    v[1] = z[1] / gg;
    v[2] = z[2] / gg;
    gg = 2. / (v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    w1 = gg * (d[0] * v[0] + o[0] * v[1] + o[1] * v[2]);
    w2 = gg * (o[0] * v[0] + d[1] * v[1] + o[2] * v[2]);
    w3 = gg * (o[1] * v[0] + o[2] * v[1] + d[2] * v[2]);
    s = (gg / 2) * (v[0] * w1 + v[1] * w2 + v[2] * w3);
    w1 -= s * v[0];
    w2 -= s * v[1];
    w3 -= s * v[2];
    d[0] -= 2 * v[0] * w1;
    d[1] -= 2 * v[1] * w2;
    o[2] -= v[1] * w3 + v[2] * w2;
    d[2] -= 2 * v[2] * w3;
    *g = gg;
  }
  if (k == 2) swapd(&z[0], &z[1]);
  else if (k == 3) swapd(&z[0], &z[2]);
  return k;
}

static void identity3(double* vec) {
  for (int i = 0; i < 9; i++) vec[i] = (i % 4 == 0) ? 1. : 0.;
}

/* kernels::CalcEigenvalues<3> (kernels.hpp:1141-1340): A column-major 3x3 (upper triangle read), lambda ascending,
 * eigenvector i in vec[3 i .. 3 i + 2] */
static void eig_sym3(const double* A, double* lambda, double* vec) {
  double d[3] = {A[0], A[4], A[8]};
  double o[3] = {A[3], A[6], A[7]}; /* a12, a13, a23 */
  double vmax = fabs(d[0]);
  if (vmax < fabs(d[1])) vmax = fabs(d[1]);
  if (vmax < fabs(d[2])) vmax = fabs(d[2]);
  if (vmax < fabs(o[0])) vmax = fabs(o[0]);
  if (vmax < fabs(o[1])) vmax = fabs(o[1]);
  if (vmax < fabs(o[2])) vmax = fabs(o[2]);
  const double mult = pow2_scale(vmax);
  d[0] /= mult; d[1] /= mult; d[2] /= mult;
  o[0] /= mult; o[1] /= mult; o[2] /= mult;
  double aa = (d[0] + d[1] + d[2]) / 3;
  double c[3] = {d[0] - aa, d[1] - aa, d[2] - aa};
  const double Q = (2 * (o[0] * o[0] + o[1] * o[1] + o[2] * o[2]) + c[0] * c[0] + c[1] * c[1] + c[2] * c[2]) / 6;
  double R = (c[0] * (o[2] * o[2] - c[1] * c[2]) + o[0] * (o[0] * c[2] - 2 * o[1] * o[2]) + o[1] * o[1] * c[1]) / 2;
  if (Q <= 0.) {
    lambda[0] = lambda[1] = lambda[2] = aa;
    identity3(vec);
  } else {
    const double sqrtQ = sqrt(Q);
    const double sqrtQ3 = Q * sqrtQ;
    double r;
    if (fabs(R) >= sqrtQ3) {
      r = (R < 0.) ? 2 * sqrtQ : -2 * sqrtQ;
    } else {
      R = R / sqrtQ3;
      if (R < 0.) r = -2 * sqrtQ * cos((acos(R) + 2.0 * ORC_PI) / 3);
      else r = -2 * sqrtQ * cos(acos(R) / 3);
    }
    aa += r;
    c[0] = d[0] - aa;
    c[1] = d[1] - aa;
    c[2] = d[2] - aa;
    if (null3_sym(o[0], o[1], o[2], c) == 3) {
      lambda[0] = lambda[1] = lambda[2] = aa;
      identity3(vec);
    } else {
      double v[3], g, cs, sn;
      const int k = deflate3(d, o, c, v, &g);
      jacobi2(o[2], &d[1], &d[2], &cs, &sn);
      double *e1, *e2, *e3; /* eigenvectors of d[0], d[1], d[2] */
      if (d[0] <= d[1]) {
        if (d[1] <= d[2]) { lambda[0] = d[0]; e1 = vec; lambda[1] = d[1]; e2 = vec + 3; lambda[2] = d[2]; e3 = vec + 6; }
        else if (d[0] <= d[2]) { lambda[0] = d[0]; e1 = vec; lambda[1] = d[2]; e3 = vec + 3; lambda[2] = d[1]; e2 = vec + 6; }
        else { lambda[0] = d[2]; e3 = vec; lambda[1] = d[0]; e1 = vec + 3; lambda[2] = d[1]; e2 = vec + 6; }
      } else {
        if (d[0] <= d[2]) { lambda[0] = d[1]; e2 = vec; lambda[1] = d[0]; e1 = vec + 3; lambda[2] = d[2]; e3 = vec + 6; }
        else if (d[1] <= d[2]) { lambda[0] = d[1]; e2 = vec; lambda[1] = d[2]; e3 = vec + 3; lambda[2] = d[0]; e1 = vec + 6; }
        else { lambda[0] = d[2]; e3 = vec; lambda[1] = d[1]; e2 = vec + 3; lambda[2] = d[0]; e1 = vec + 6; }
      }
      e1[0] = c[0];
      e1[1] = c[1];
      e1[2] = c[2];
      const double t2 = g * (v[1] * cs - v[2] * sn);
      const double t3 = g * (v[1] * sn + v[2] * cs);
      e2[0] = -v[0] * t2;      e3[0] = -v[0] * t3;
      e2[1] = cs - v[1] * t2;  e3[1] = sn - v[1] * t3;
      e2[2] = -sn - v[2] * t2; e3[2] = cs - v[2] * t3;
      if (k == 2) {
        swapd(&e2[0], &e2[1]);
        swapd(&e3[0], &e3[1]);
      } else if (k == 3) {
        swapd(&e2[0], &e2[2]);
        swapd(&e3[0], &e3[2]);
      }
    }
  }
  lambda[0] *= mult;
  lambda[1] *= mult;
  lambda[2] *= mult;
}

/* kernels::CalcEigenvalues<2> (kernels.hpp:1110-1136) */
static void eig_sym2(const double* A, double* lambda, double* vec) {
  double d0 = A[0], d3 = A[3], c, s;
  jacobi2(A[2], &d0, &d3, &c, &s);
  if (d0 <= d3) {
    lambda[0] = d0; lambda[1] = d3;
    vec[0] = c; vec[1] = -s; vec[2] = s; vec[3] = c;
  } else {
    lambda[0] = d3; lambda[1] = d0;
    vec[0] = s; vec[1] = c; vec[2] = c; vec[3] = -s;
  }
}

/* mfem::DenseMatrix::CalcEigenvalues (densemat.cpp:1309-1329) */
void orc_eig_sym(int dim, const double* A, double* lambda, double* vec) {
  if (dim == 2) eig_sym2(A, lambda, vec);
  else eig_sym3(A, lambda, vec);
}

/* f(A) = V f(Lambda) V^T as LogarithmicStrain / the exponential map build it: mfem::MultADAt (densemat.cpp:2762-2776) */
void orc_mult_ADAt(int dim, const double* V, const double* D, double* out) {
  for (int i = 0; i < dim; i++)
    for (int j = 0; j <= i; j++) {
      double t = 0.;
      for (int k = 0; k < dim; k++) t += D[k] * V[i + k * dim] * V[j + k * dim];
      out[j + i * dim] = out[i + j * dim] = t;
    }
}
