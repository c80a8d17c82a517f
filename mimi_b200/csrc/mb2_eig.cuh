// Device-side spectrum of a symmetric 2x2 / 3x3 matrix in the operation order of the solver the reference's J2Log
// material calls: mfem::DenseMatrix::CalcEigenvalues (third_party/mfem/linalg/densemat.cpp:1309-1329) =
// kernels::CalcEigenvalues<2|3> (third_party/mfem/linalg/kernels.hpp:1110-1340, helpers :441-1065), used by
// LogarithmicStrain (src/mimi/materials/material_utils.hpp:92-114) and by the exponential map of
// J2Log::PlasticStress (src/mimi/materials/materials.hpp:700-712).
//
// 3x3: power-of-two scaling, extreme root of the characteristic cubic (trigonometric form), unit near-kernel vector z
// of A - root I through a pivoted Householder QR, deflation by the reflection that maps P z to e_1, one Jacobi rotation
// of the remaining 2x2 block, back-transformation, ascending order. Only the reference's pivoting mode 0 is reachable.
// Everything is scalar register code with static indexing; sqrt and division are IEEE (nvcc defaults), hypot / acos /
// cos come from libdevice (the reference: glibc), hence the stated tolerance of the J2Log parity tests.
#pragma once
#include <cfloat>

namespace mb2 {

__device__ __forceinline__ void swapd(double& a, double& b) {
  const double t = a;
  a = b;
  b = t;
}

// Jacobi rotation of [[a, b], [b, d]] (kernels.hpp:463-494)
__device__ __forceinline__ void eig_jacobi2(double b, double& a, double& d, double& c, double& s) {
  const double big = sqrt(1. / DBL_EPSILON);
  if (b == 0.0) {
    c = 1.;
    s = 0.;
    return;
  }
  double t;
  const double zeta = (d - a) / (2 * b);
  const double az = fabs(zeta);
  if (az < big) t = copysign(1. / (az + sqrt(1. + zeta * zeta)), zeta);
  else t = copysign(0.5 / az, zeta);
  c = sqrt(1. / (1. + t * t));
  s = c * t;
  t *= b;
  a -= t;
  d += t;
}

// near-kernel vector of the general 2x2 [[p11, p12], [p21, p22]], returned in (p11, p22); true = zero matrix
// (kernels.hpp:519-659, mode 0)
__device__ __forceinline__ bool eig_null2(double& p11, double& p12, double& p21, double& p22) {
  double n1 = fabs(p11) + fabs(p21);
  double n2 = fabs(p22) + fabs(p12);
  const bool swap_cols = n2 > n1;
  double mu;
  if (!swap_cols) {
    if (n1 == 0.) return true;
    if (fabs(p11) > fabs(p21)) {
      swapd(p11, p21);
      swapd(p12, p22);
    }
  } else if (fabs(p12) > fabs(p22)) {
    swapd(p11, p22);
    swapd(p12, p21);
  } else {
    swapd(p11, p12);
    swapd(p21, p22);
  }
  n1 = hypot(p11, p21);
  if (p21 != 0.) {
    mu = copysign(n1, p11);
    n1 = -p21 * (p21 / (p11 + mu));
    p11 = mu;
    if (fabs(n1) <= fabs(p21)) {
      n1 = n1 / p21;
      mu = (2. / (1. + n1 * n1)) * (n1 * p12 + p22);
      p22 = p22 - mu;
      p12 = p12 - mu * n1;
    } else {
      n2 = p21 / n1;
      mu = (2. / (1. + n2 * n2)) * (p12 + n2 * p22);
      p22 = p22 - mu * n2;
      p12 = p12 - mu;
    }
  }
  mu = -p12 / p11;
  n2 = 1. / (1. + fabs(mu));
  if (fabs(p11) <= n2 * fabs(p22)) {
    p22 = 0.;
    p11 = 1.;
  } else {
    p22 = n2;
    p11 = mu * n2;
  }
  if (swap_cols) swapd(p11, p22);
  return false;
}

// (kernels.hpp:662-678): x1 carries the largest magnitude
__device__ __forceinline__ void eig_unit3_lead(double x1, double x2, double x3, double& n1, double& n2, double& n3) {
  const double m = fabs(x1);
  double r = x2 / m;
  double t = 1. + r * r;
  r = x3 / m;
  t = sqrt(1. / (t + r * r));
  n1 = copysign(t, x1);
  t /= m;
  n2 = x2 * t;
  n3 = x3 * t;
}

// (z1, z2, z3) <- z / |z| (kernels.hpp:681-707)
__device__ __forceinline__ void eig_unit3(double& z1, double& z2, double& z3) {
  const double x1 = z1, x2 = z2, x3 = z3;
  if (fabs(x1) >= fabs(x2)) {
    if (fabs(x1) >= fabs(x3)) {
      if (x1 != 0.) eig_unit3_lead(x1, x2, x3, z1, z2, z3);
      else z1 = z2 = z3 = 0.;
      return;
    }
  } else if (fabs(x2) >= fabs(x3)) {
    eig_unit3_lead(x2, x1, x3, z2, z1, z3);
    return;
  }
  eig_unit3_lead(x3, x1, x2, z3, z1, z2);
}

// unit near-kernel vector of the general 3x3 [[d1, c12, c13], [c21, d2, c23], [c31, c32, d3]] (pivot column in
// front): Householder on column 1, eig_null2 on the trailing block (kernels.hpp:709-814). z in (d1, d2, d3);
// returns the kernel dimension (1, or 2: z orthogonal to the kernel).
__device__ __forceinline__ int eig_null3_general(double& d1, double& d2, double& d3, double& c12, double& c13,
                                                 double& c23, double& c21, double& c31, double& c32) {
  int kdim;
  double mu, n1, n2, n3, s1, s2, s3;
  s1 = hypot(c21, c31);
  n1 = hypot(d1, s1);
  if (s1 != 0.) {
    mu = copysign(n1, d1);
    n1 = -s1 * (s1 / (d1 + mu));
    d1 = mu;
    const bool n1_ge_c21 = fabs(n1) >= fabs(c21);
    if (n1_ge_c21 && fabs(n1) >= fabs(c31)) {
      s2 = c21 / n1;
      s3 = c31 / n1;
      mu = 2. / (1. + s2 * s2 + s3 * s3);
      n2 = mu * (c12 + s2 * d2 + s3 * c32);
      n3 = mu * (c13 + s2 * c23 + s3 * d3);
      c12 = c12 - n2;
      d2 = d2 - s2 * n2;
      c32 = c32 - s3 * n2;
      c13 = c13 - n3;
      c23 = c23 - s2 * n3;
      d3 = d3 - s3 * n3;
    } else if (!n1_ge_c21 && fabs(c21) >= fabs(c31)) {
      s1 = n1 / c21;
      s3 = c31 / c21;
      mu = 2. / (1. + s1 * s1 + s3 * s3);
      n2 = mu * (s1 * c12 + d2 + s3 * c32);
      n3 = mu * (s1 * c13 + c23 + s3 * d3);
      c12 = c12 - s1 * n2;
      d2 = d2 - n2;
      c32 = c32 - s3 * n2;
      c13 = c13 - s1 * n3;
      c23 = c23 - n3;
      d3 = d3 - s3 * n3;
    } else {
      s1 = n1 / c31;
      s2 = c21 / c31;
      mu = 2. / (1. + s1 * s1 + s2 * s2);
      n2 = mu * (s1 * c12 + s2 * d2 + c32);
      n3 = mu * (s1 * c13 + s2 * c23 + d3);
      c12 = c12 - s1 * n2;
      d2 = d2 - s2 * n2;
      c32 = c32 - n2;
      c13 = c13 - s1 * n3;
      c23 = c23 - s2 * n3;
      d3 = d3 - n3;
    }
  }
  if (eig_null2(d2, c23, c32, d3)) {
    d2 = c12 / d1;
    d3 = c13 / d1;
    d1 = 1.;
    kdim = 2;
  } else {
    d1 = -(c12 * d2 + c13 * d3) / d1;
    kdim = 1;
  }
  eig_unit3(d1, d2, d3);
  return kdim;
}

// unit near-kernel vector of the symmetric [[d1, o12, o13], [o12, d2, o23], [o13, o23, d3]] (kernels.hpp:816-942):
// column of largest l1 norm first, row with the smallest pivot entry on top, then eig_null3_general.
// z overwrites (d1, d2, d3); 3 = zero matrix.
__device__ __forceinline__ int eig_null3_sym(double o12, double o13, double o23, double& d1, double& d2, double& d3) {
  double c12 = o12, c13 = o13, c23 = o23;
  double c21, c31, c32;
  int col, row;
  c32 = fabs(d1) + fabs(c12) + fabs(c13);
  c31 = fabs(d2) + fabs(c12) + fabs(c23);
  c21 = fabs(d3) + fabs(c13) + fabs(c23);
  if (c32 >= c21) col = (c32 >= c31) ? 1 : 2;
  else col = (c31 >= c21) ? 2 : 3;
  if (col == 1) {
    if (c32 == 0.) return 3;
  } else if (col == 2) {
    if (c31 == 0.) return 3;
    swapd(c13, c23);
    swapd(d1, d2);
  } else {
    if (c21 == 0.) return 3;
    swapd(c12, c23);
    swapd(d1, d3);
  }
  if (fabs(d1) <= fabs(c13)) row = (fabs(d1) <= fabs(c12)) ? 1 : 2;
  else row = (fabs(c12) <= fabs(c13)) ? 2 : 3;
  double g11, g12, g13, g21, g22, g23, g31, g32, g33;
  if (row == 1) {
    g11 = d1;  g12 = c12; g13 = c13;
    g21 = c12; g22 = d2;  g23 = c23;
    g31 = c13; g32 = c23; g33 = d3;
  } else if (row == 2) {  // rows 1 <-> 2
    g11 = c12; g12 = d2;  g13 = c23;
    g21 = d1;  g22 = c12; g23 = c13;
    g31 = c13; g32 = c23; g33 = d3;
  } else {  // rows 1 <-> 3
    g11 = c13; g12 = c23; g13 = d3;
    g21 = c12; g22 = d2;  g23 = c23;
    g31 = d1;  g32 = c12; g33 = c13;
  }
  const int kdim = eig_null3_general(g11, g22, g33, g12, g13, g23, g21, g31, g32);
  d1 = g11;
  d2 = g22;
  d3 = g33;
  if (col == 2) swapd(d1, d2);
  else if (col == 3) swapd(d1, d3);
  return kdim;
}

// deflation (kernels.hpp:944-1065, mode 0): B = Q P A P Q = diag(b1, [[b2, b23], [b23, b3]]); P swaps entries 1 and k
// (|z_k| smallest), Q = I - g v v^T maps P z to e_1. b1, b2, b3 -> (d1, d2, d3), b23 -> o23. Returns k (1-based).
__device__ __forceinline__ int eig_deflate3(double& d1, double& d2, double& d3, double& o12, double& o13, double& o23,
                                            double& z1, double& z2, double& z3, double& v1, double& v2, double& v3,
                                            double& g) {
  int k;
  double s, w1, w2, w3;
  if (fabs(z1) <= fabs(z3)) k = (fabs(z1) <= fabs(z2)) ? 1 : 2;
  else k = (fabs(z2) <= fabs(z3)) ? 2 : 3;
  if (k == 2) {
    swapd(o13, o23);
    swapd(d1, d2);
    swapd(z1, z2);
  } else if (k == 3) {
    swapd(o12, o23);
    swapd(d1, d3);
    swapd(z1, z3);
  }
  s = hypot(z2, z3);
  if (s == 0.) {
    v1 = v2 = v3 = 0.;
    g = 1.;
  } else {
    g = copysign(1., z1);
    v1 = -s * (s / (z1 + g));
    g = fabs(v1);
    if (fabs(z2) > g) g = fabs(z2);
    if (fabs(z3) > g) g = fabs(z3);
    v1 = v1 / g;
    v2 = z2 / g;
    v3 = z3 / g;
    g = 2. / (v1 * v1 + v2 * v2 + v3 * v3);
    w1 = g * (d1 * v1 + o12 * v2 + o13 * v3);
    w2 = g * (o12 * v1 + d2 * v2 + o23 * v3);
    w3 = g * (o13 * v1 + o23 * v2 + d3 * v3);
    s = (g / 2) * (v1 * w1 + v2 * w2 + v3 * w3);
    w1 -= s * v1;
    w2 -= s * v2;
    w3 -= s * v3;
    d1 -= 2 * v1 * w1;
    d2 -= 2 * v2 * w2;
    o23 -= v2 * w3 + v3 * w2;
    d3 -= 2 * v3 * w3;
  }
  if (k == 2) swapd(z1, z2);
  else if (k == 3) swapd(z1, z3);
  return k;
}

// kernels::CalcEigenvalues<3> (kernels.hpp:1141-1340): A column-major (upper triangle read), lambda ascending,
// eigenvector i in vec[3 i .. 3 i + 2].
static __device__ __noinline__ void eig_sym3(const double* A, double* lambda, double* vec) {
  double d1 = A[0], d2 = A[4], d3 = A[8];
  double o12 = A[3], o13 = A[6], o23 = A[7];
  double mult;
  {
    double vmax = fabs(d1);
    if (vmax < fabs(d2)) vmax = fabs(d2);
    if (vmax < fabs(d3)) vmax = fabs(d3);
    if (vmax < fabs(o12)) vmax = fabs(o12);
    if (vmax < fabs(o13)) vmax = fabs(o13);
    if (vmax < fabs(o23)) vmax = fabs(o23);
    if (vmax > 0.) {  // 2^e with vmax / 2^e in [0.5, 1) (kernels.hpp:497-517)
      int e;
      mult = frexp(vmax, &e);
      if (e == DBL_MAX_EXP) mult *= 2.0;
      mult = vmax / mult;
    } else {
      mult = 1.;
    }
  }
  d1 /= mult; d2 /= mult; d3 /= mult;
  o12 /= mult; o13 /= mult; o23 /= mult;
  double aa = (d1 + d2 + d3) / 3;
  double c1 = d1 - aa, c2 = d2 - aa, c3 = d3 - aa;
  const double Q = (2 * (o12 * o12 + o13 * o13 + o23 * o23) + c1 * c1 + c2 * c2 + c3 * c3) / 6;
  double R = (c1 * (o23 * o23 - c2 * c3) + o12 * (o12 * c3 - 2 * o13 * o23) + o13 * o13 * c2) / 2;
  bool spherical = Q <= 0.;
  if (!spherical) {
    const double sqrtQ = sqrt(Q);
    const double sqrtQ3 = Q * sqrtQ;
    double r;
    if (fabs(R) >= sqrtQ3) {
      r = (R < 0.) ? 2 * sqrtQ : -2 * sqrtQ;
    } else {
      R = R / sqrtQ3;
      if (R < 0.) r = -2 * sqrtQ * cos((acos(R) + 2.0 * 3.14159265358979323846) / 3);
      else r = -2 * sqrtQ * cos(acos(R) / 3);
    }
    aa += r;
    c1 = d1 - aa;
    c2 = d2 - aa;
    c3 = d3 - aa;
    spherical = eig_null3_sym(o12, o13, o23, c1, c2, c3) == 3;
  }
  if (spherical) {
    lambda[0] = lambda[1] = lambda[2] = aa;
#pragma unroll
    for (int i = 0; i < 9; i++) vec[i] = (i % 4 == 0) ? 1. : 0.;
  } else {
    double v1, v2, v3, g, cs, sn;
    const int k = eig_deflate3(d1, d2, d3, o12, o13, o23, c1, c2, c3, v1, v2, v3, g);
    eig_jacobi2(o23, d2, d3, cs, sn);
    // slots of the eigenvectors of d1 / d2 / d3 in ascending order (ties resolved as in kernels.hpp:1262-1306)
    int s1, s2;  // the third takes the slot left over
    if (d1 <= d2) {
      if (d2 <= d3) { s1 = 0; s2 = 1; }
      else if (d1 <= d3) { s1 = 0; s2 = 2; }
      else { s1 = 1; s2 = 2; }
    } else {
      if (d1 <= d3) { s2 = 0; s1 = 1; }
      else if (d2 <= d3) { s2 = 0; s1 = 2; }
      else { s2 = 1; s1 = 2; }
    }
    const double t2 = g * (v2 * cs - v3 * sn);
    const double t3 = g * (v2 * sn + v3 * cs);
    double e2[3] = {-v1 * t2, cs - v2 * t2, -sn - v3 * t2};
    double e3[3] = {-v1 * t3, sn - v2 * t3, cs - v3 * t3};
    if (k == 2) {
      swapd(e2[0], e2[1]);
      swapd(e3[0], e3[1]);
    } else if (k == 3) {
      swapd(e2[0], e2[2]);
      swapd(e3[0], e3[2]);
    }
    const double e1[3] = {c1, c2, c3};
#pragma unroll
    for (int slot = 0; slot < 3; slot++) {
      lambda[slot] = (s1 == slot) ? d1 : ((s2 == slot) ? d2 : d3);
#pragma unroll
      for (int j = 0; j < 3; j++) vec[3 * slot + j] = (s1 == slot) ? e1[j] : ((s2 == slot) ? e2[j] : e3[j]);
    }
  }
  lambda[0] *= mult;
  lambda[1] *= mult;
  lambda[2] *= mult;
}

// kernels::CalcEigenvalues<2> (kernels.hpp:1110-1136)
__device__ __forceinline__ void eig_sym2(const double* A, double* lambda, double* vec) {
  double d0 = A[0], d3 = A[3], c, s;
  eig_jacobi2(A[2], d0, d3, c, s);
  if (d0 <= d3) {
    lambda[0] = d0; lambda[1] = d3;
    vec[0] = c; vec[1] = -s; vec[2] = s; vec[3] = c;
  } else {
    lambda[0] = d3; lambda[1] = d0;
    vec[0] = s; vec[1] = c; vec[2] = c; vec[3] = -s;
  }
}

template <int DIM>
__device__ __forceinline__ void eig_sym(const double* A, double* lambda, double* vec) {
  if (DIM == 2) eig_sym2(A, lambda, vec);
  else eig_sym3(A, lambda, vec);
}

// mfem::MultADAt (densemat.cpp:2762-2776): out = V diag(D) V^T, lower triangle computed and mirrored
template <int DIM>
__device__ __forceinline__ void mult_ADAt(const double* V, const double* D, double* out) {
#pragma unroll
  for (int i = 0; i < DIM; i++)
#pragma unroll
    for (int j = 0; j <= i; j++) {
      double t = 0.;
#pragma unroll
      for (int k = 0; k < DIM; k++) t += D[k] * V[i + k * DIM] * V[j + k * DIM];
      out[j + i * DIM] = out[i + j * DIM] = t;
    }
}

}  // namespace mb2
