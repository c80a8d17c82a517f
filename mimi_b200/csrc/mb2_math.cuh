// Device-side small dense kernels and material laws, written in the reference's exact operation
// order so that FP64 results are bit-identical to its MFEM/host path.
// This translation unit set is compiled with -fmad=false: no mul+add may be contracted.
#pragma once
#include <cstdint>

#include "mb2_eig.cuh"

namespace mb2 {

enum : int { MAT_NEOHOOKEAN = 0, MAT_STVK = 1, MAT_J2LINEAR = 2, MAT_J2 = 3, MAT_J2SIMO = 4, MAT_J2LOG = 5 };
// hardening laws (src/mimi/materials/material_hardening.hpp); hp[] layout:
//  POWERLAW {sigma_y, n, eps0}   VOCE {sigma_y, sigma_sat, strain_constant}
//  JC* {A, B, n, C, eps0_dot, T_ref, T_melt, m, temperature_contribution_ (constant-temperature law)}
enum : int { HARD_POWERLAW = 0, HARD_VOCE = 1, HARD_JC = 2, HARD_JC_RATE = 3, HARD_JC_THERMO_RATE = 4, HARD_JC_CONST_T = 5 };

// MaterialBase parameters (materials.hpp:24-40) plus the J2Linear members (materials.hpp:147-166); the sqrt
// constants are computed on the host by std::sqrt exactly like the reference's member initialisers.
struct MatP {
  double lambda = 0, mu = 0;
  double K = 0, G = 0, iso = 0, kin = 0, sigma_y = 0, sqrt_6 = 0, sqrt_3_2 = 0, sqrt_2_3 = 0;
  // J2 family (materials.hpp:259-404): hardening law, thermo members, dt_ pushed by the integrator, density_
  int hard_kind = 0;
  double hp[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  double dt = 0, heat_fraction = 0, density = 0, specific_heat = 0, T0 = 0;
  int* flag = nullptr;  // device failure flag: 1 = root not bracketed, 2 = scalar solve did not converge
  // optional measurement counters (mb2_set_option "j2_counters"): [0] return mappings run (= yielded stress evaluations),
  // [1] hardening-law evaluations inside their bracketed Newton solves (ScalarSolve residual calls)
  unsigned long long* cnt = nullptr;
};

// per-quadrature-point history doubles: J2Linear = plastic_strain[9], beta[9], accumulated_plastic_strain
//                                   J2 = plastic_strain[9], accumulated_plastic_strain, temperature
//                                   J2Simo = be_old[9], F_old[9], accumulated_plastic_strain, temperature
//                                   J2Log = Fp_inv[9], accumulated_plastic_strain, temperature
__host__ __device__ constexpr int mat_state_size(int mat) {
  return mat == MAT_J2LINEAR ? 19 : ((mat == MAT_J2 || mat == MAT_J2LOG) ? 11 : (mat == MAT_J2SIMO ? 20 : 0));
}

// mfem::kernels::MultABt on DIM x DIM operands (third_party/mfem/linalg/kernels.hpp:363-384):
// C zeroed, then k outer, j, i inner: C(i,j) += A(i,k) * B(j,k).
template <int DIM>
__device__ __forceinline__ void mult_ABt(const double* A, const double* B, double* C) {
#pragma unroll
  for (int i = 0; i < DIM * DIM; i++) C[i] = 0.0;
#pragma unroll
  for (int k = 0; k < DIM; k++)
#pragma unroll
    for (int j = 0; j < DIM; j++) {
      const double bjk = B[j + DIM * k];
#pragma unroll
      for (int i = 0; i < DIM; i++) C[i + DIM * j] += A[i + DIM * k] * bjk;
    }
}

// mfem::MultAtB (third_party/mfem/linalg/densemat.cpp:3051-3068): C(i,j) = sum_k A(k,i) B(k,j), d from 0.0.
template <int DIM>
__device__ __forceinline__ void mult_AtB(const double* A, const double* B, double* C) {
#pragma unroll
  for (int j = 0; j < DIM; j++)
#pragma unroll
    for (int i = 0; i < DIM; i++) {
      double d = 0.0;
#pragma unroll
      for (int k = 0; k < DIM; k++) d += A[k + DIM * i] * B[k + DIM * j];
      C[i + DIM * j] = d;
    }
}

// mfem::kernels::Mult (kernels.hpp:341-357): A = B*C; zero, then j, k, i.
template <int DIM>
__device__ __forceinline__ void mult_AB(const double* B, const double* C, double* A) {
#pragma unroll
  for (int i = 0; i < DIM * DIM; i++) A[i] = 0.0;
#pragma unroll
  for (int j = 0; j < DIM; j++)
#pragma unroll
    for (int k = 0; k < DIM; k++)
#pragma unroll
      for (int i = 0; i < DIM; i++) A[i + DIM * j] += B[i + DIM * k] * C[k + DIM * j];
}

// mfem::DenseMatrix::Det (densemat.cpp:510-525)
template <int DIM>
__device__ __forceinline__ double det_small(const double* d) {
  if (DIM == 2) return d[0] * d[3] - d[1] * d[2];
  return d[0] * (d[4] * d[8] - d[5] * d[7]) + d[3] * (d[2] * d[7] - d[1] * d[8]) +
         d[6] * (d[1] * d[5] - d[2] * d[4]);
}

// mfem::kernels::CalcInverse<DIM> (kernels.hpp:244-251; adjugate + its own determinant expression,
// tmatrix.hpp:331-362 for 2x2, :483-525 for 3x3), every entry then multiplied by 1.0/det.
template <int DIM>
__device__ __forceinline__ void inverse_small(const double* A, double* B) {
  if (DIM == 2) {
    B[0] = A[3];
    B[2] = -A[2];
    B[1] = -A[1];
    B[3] = A[0];
    const double det = A[0] * A[3] - A[1] * A[2];
    const double s = 1.0 / det;
#pragma unroll
    for (int i = 0; i < 4; i++) B[i] *= s;
    return;
  }
#define MB2_a(i, j) A[(i) + 3 * (j)]
#define MB2_b(i, j) B[(i) + 3 * (j)]
  MB2_b(0, 0) = MB2_a(1, 1) * MB2_a(2, 2) - MB2_a(1, 2) * MB2_a(2, 1);
  MB2_b(0, 1) = MB2_a(0, 2) * MB2_a(2, 1) - MB2_a(0, 1) * MB2_a(2, 2);
  MB2_b(0, 2) = MB2_a(0, 1) * MB2_a(1, 2) - MB2_a(0, 2) * MB2_a(1, 1);
  MB2_b(1, 0) = MB2_a(1, 2) * MB2_a(2, 0) - MB2_a(1, 0) * MB2_a(2, 2);
  MB2_b(1, 1) = MB2_a(0, 0) * MB2_a(2, 2) - MB2_a(0, 2) * MB2_a(2, 0);
  MB2_b(1, 2) = MB2_a(0, 2) * MB2_a(1, 0) - MB2_a(0, 0) * MB2_a(1, 2);
  MB2_b(2, 0) = MB2_a(1, 0) * MB2_a(2, 1) - MB2_a(1, 1) * MB2_a(2, 0);
  MB2_b(2, 1) = MB2_a(0, 1) * MB2_a(2, 0) - MB2_a(0, 0) * MB2_a(2, 1);
  MB2_b(2, 2) = MB2_a(0, 0) * MB2_a(1, 1) - MB2_a(0, 1) * MB2_a(1, 0);
  const double det = (MB2_a(0, 0) * MB2_b(0, 0) + MB2_a(1, 0) * MB2_b(0, 1) + MB2_a(2, 0) * MB2_b(0, 2));
  const double s = 1.0 / det;
#pragma unroll
  for (int i = 0; i < 9; i++) B[i] *= s;
#undef MB2_a
#undef MB2_b
}

// J2Linear::PlasticStress<ACC> (src/mimi/materials/materials.hpp:187-238) on st = {plastic_strain[9], beta[9],
// accumulated_plastic_strain}. ACC == false: sigma out, state read-only (EvaluateCauchy); ACC == true: the history
// is committed (Accumulate, materials.hpp:245-248). Helpers in the reference's order: ElasticStrain
// (material_utils.hpp:60-85: copy F, Symmetrize, -I, -plastic strain), Dev (:22-56), Norm (:117-127).
template <int DIM, bool ACC>
__device__ __forceinline__ void j2linear_plastic_stress(const MatP& m, const double* F, double* st, double* sigma) {
  constexpr int DD = DIM * DIM;
  double eps[DD], s[DD], eta[DD];
  double* plastic_strain = st;
  double* beta = st + 9;
#pragma unroll
  for (int i = 0; i < DD; i++) eps[i] = F[i];
#pragma unroll
  for (int i = 0; i < DIM; i++)
#pragma unroll
    for (int j = 0; j < i; j++) {
      const double a = 0.5 * (eps[i * DIM + j] + eps[j * DIM + i]);
      eps[j * DIM + i] = eps[i * DIM + j] = a;
    }
#pragma unroll
  for (int i = 0; i < DIM; i++) eps[(DIM + 1) * i] -= 1.;
#pragma unroll
  for (int i = 0; i < DD; i++) eps[i] -= plastic_strain[i];
  double tr = 0.0;
#pragma unroll
  for (int i = 0; i < DIM; i++) tr += eps[i + i * DIM];
  const double p = m.K * tr;
  const double factor = 2.0 * m.G;
  if (DIM == 2) {
    const double trd = (eps[0] + eps[3]) / 2.0;
    s[0] = (eps[0] - trd) * factor;
    s[1] = eps[1] * factor;
    s[2] = eps[2] * factor;
    s[3] = (eps[3] - trd) * factor;
  } else {
    const double trd = (eps[0] + eps[4] + eps[8]) / 3.0;
#pragma unroll
    for (int i = 0; i < DD; i++) s[i] = ((i % (DIM + 1) == 0) ? (eps[i] - trd) : eps[i]) * factor;
  }
#pragma unroll
  for (int i = 0; i < DD; i++) eta[i] = s[i] + -1. * beta[i];  // mfem::Add(s, beta, -1., eta)
  double a = 0.0;
#pragma unroll
  for (int i = 0; i < DD; i++) a += eta[i] * eta[i];
  const double eta_norm = sqrt(a);
  const double q = m.sqrt_3_2 * eta_norm;
  const double phi = q - (m.sigma_y + m.iso * st[18]);
  if (phi > 0.) {
    const double inc = phi / (3. * m.G + m.kin + m.iso);
    const double inv = 1. / eta_norm;
#pragma unroll
    for (int i = 0; i < DD; i++) eta[i] *= inv;
    if (!ACC) {
      const double c = -m.sqrt_6 * m.G * inc;
#pragma unroll
      for (int i = 0; i < DD; i++) s[i] += c * eta[i];
    } else {
      st[18] += inc;
      const double c1 = m.sqrt_3_2 * inc;
#pragma unroll
      for (int i = 0; i < DD; i++) plastic_strain[i] += c1 * eta[i];
      const double c2 = m.sqrt_2_3 * m.kin * inc;
#pragma unroll
      for (int i = 0; i < DD; i++) beta[i] += c2 * eta[i];
    }
  }
  if (!ACC) {
#pragma unroll
    for (int i = 0; i < DD; i++) sigma[i] = s[i] + p * ((i % (DIM + 1) == 0) ? 1.0 : 0.0);
  }
}

// ---- J2 with nonlinear isotropic hardening (src/mimi/materials/materials.hpp:259-404) ----
// ADScalar<double,1> operator sequences of the hardening laws (src/mimi/utils/ad.inl). pow / exp / log are CUDA's
// libdevice functions (<= 1 ulp) where the reference calls glibc (< 1 ulp): values agree to the last bit or two, which
// the tests state as a tolerance for this material family (every other operation keeps the reference's order).
struct AD1 {
  double v, d;
};
__device__ __forceinline__ AD1 ad_pow(AD1 b, double power) {  // ad.inl:261-279
  const double tmp = pow(b.v, power - 1.0);
  b.v *= tmp;
  b.d *= power * tmp;
  return b;
}
__device__ __forceinline__ AD1 hardening_evaluate(const MatP& m, AD1 eps) {
  AD1 r;
  if (m.hard_kind == HARD_POWERLAW) {  // sigma_y * pow(1.0 + eps / eps0, 1.0 / n)
    const double inv = 1 / m.hp[2];
    AD1 t = {eps.v * inv, eps.d * inv};
    t.v += 1.0;
    t = ad_pow(t, 1.0 / m.hp[1]);
    r.v = t.v * m.hp[0];
    r.d = t.d * m.hp[0];
    return r;
  }
  if (m.hard_kind == HARD_VOCE) {  // sigma_sat - (sigma_sat - sigma_y) * exp(-eps / strain_constant)
    const double inv = 1 / m.hp[2];
    AD1 t = {-eps.v * inv, -eps.d * inv};
    const double tmp = exp(t.v);
    t.v = 1. * tmp;
    t.d = t.d * tmp;
    const double c = m.hp[1] - m.hp[0];
    t.v *= c;
    t.d *= c;
    r.v = -t.v + m.hp[1];
    r.d = -t.d;
    return r;
  }
  if (fabs(eps.v) < 1.e-13) {  // JohnsonCookHardening::Evaluate, material_hardening.hpp:143-150
    r.v = m.hp[0];
    r.d = 0.0;
    return r;
  }
  AD1 t = ad_pow(eps, m.hp[2]);
  t.v *= m.hp[1];
  t.d *= m.hp[1];
  r.v = t.v + m.hp[0];
  r.d = t.d;
  return r;
}
__device__ __forceinline__ double hardening_rate(const MatP& m, double rate) {  // material_hardening.hpp:186-197
  if (m.hard_kind < HARD_JC_RATE) return 1.;
  double v = 1.0;
  if (rate > m.hp[4]) v += m.hp[3] * log(rate / m.hp[4]);
  return v;
}
__device__ __forceinline__ double hardening_thermo(const MatP& m, double T) {  // material_hardening.hpp:262-279
  if (m.hard_kind < HARD_JC_THERMO_RATE) return 1.;
  if (m.hard_kind == HARD_JC_CONST_T) return m.hp[8];
  double v = 1.0;
  if (T < m.hp[5]) {
  } else if (T > m.hp[6]) {
    v = 0.0;
  } else {
    v -= pow((T - m.hp[5]) / (m.hp[6] - m.hp[5]), m.hp[7]);
  }
  return v;
}
// return-mapping residual (materials.hpp:345-351) for a double and for an ADScalar argument (delta / dt is a true
// division in the first and a multiplication by 1/dt in the second, ad.inl:184-193)
// RC = {q (or s_effective), eqps_old, thermo, simo, be_trace}: J2Simo's residual is q - G * delta * be_trace - ...
// (materials.hpp:517-523) instead of q - 3 G delta - ...
struct RC {
  double q, eqps_old, thermo;
  bool simo;
  double bt;
};
__device__ __forceinline__ double j2_residual_d(const MatP& m, const RC& c, double de) {
  const double q = c.q, eqps_old = c.eqps_old, thermo = c.thermo;
  const AD1 e = {eqps_old + de, 0.0};
  const AD1 H = hardening_evaluate(m, e);
  const double cc = hardening_rate(m, de / m.dt) * thermo;
  if (c.simo) return -(H.v * cc) + (q - m.G * de * c.bt);
  return -(H.v * cc) + (q - 3.0 * m.G * de);
}
__device__ __forceinline__ AD1 j2_residual_ad(const MatP& m, const RC& c, double x) {
  const double q = c.q, eqps_old = c.eqps_old, thermo = c.thermo;
  const double t1 = 3.0 * m.G;
  AD1 a = {-(x * t1) + q, -(1.0 * t1)};
  if (c.simo) {
    a.v = -((x * m.G) * c.bt) + q;
    a.d = -((1.0 * m.G) * c.bt);
  }
  const AD1 e = {x + eqps_old, 1.0};
  AD1 H = hardening_evaluate(m, e);
  const double inv_dt = 1 / m.dt;
  const double cc = hardening_rate(m, x * inv_dt) * thermo;
  H.v *= cc;
  H.d *= cc;
  a.v -= H.v;
  a.d -= H.d;
  return a;
}
// mimi::solvers::ScalarSolve (src/mimi/solvers/newton.hpp:53-169), x0 = 0; failures raise the device flag
static __device__ __forceinline__ double j2_scalar_solve(const MatP& m, const RC& c, double lower, double upper, double xtol,
                                               double rtol, int max_iter) {
  const double fl = j2_residual_d(m, c, lower), fh = j2_residual_d(m, c, upper);
  if (fabs(fl) < xtol) return lower;
  if (fabs(fh) < xtol) return upper;
  if (fl * fh > 0.) {
    if (m.flag) *m.flag = 1;
    return 0.0;
  }
  double xl = lower, xh = upper;
  if (fl > 0) {
    xl = upper;
    xh = lower;
  }
  double x = 0.0;
  if (x < lower || x > upper) x = 0.5 * (lower + upper);
  double delta_x_old = fabs(upper - lower), delta_x = delta_x_old;
  AD1 R = j2_residual_ad(m, c, x);
  double fval = R.v, df = R.d;
  int it = 0;
  bool converged = false;
  while (!converged) {
    if (it == max_iter) {
      if (m.flag) *m.flag = 2;
      break;
    }
    if ((x - xh) * df - fval > 0 || (x - xl) * df - fval < 0 || fabs(2. * fval) > fabs(delta_x_old * df)) {
      delta_x_old = delta_x;
      delta_x = 0.5 * (xh - xl);
      x = xl + delta_x;
    } else {
      delta_x_old = delta_x;
      delta_x = fval / df;
      x -= delta_x;
    }
    R = j2_residual_ad(m, c, x);
    fval = R.v;
    df = R.d;
    converged = (fabs(delta_x) < xtol) || (fabs(fval) < rtol);
    if (fval < 0) xl = x; else xh = x;
    ++it;
  }
  if (m.cnt) {
    atomicAdd(&m.cnt[0], 1ULL);
    atomicAdd(&m.cnt[1], (unsigned long long)(it + 3));  // two bracket residuals, the first AD residual, one per iteration
  }
  return x;
}
// J2::PlasticStress<ACC> (materials.hpp:311-391) on st = {plastic_strain[9], accumulated_plastic_strain, temperature}
template <int DIM, bool ACC>
__device__ __forceinline__ void j2_plastic_stress(const MatP& m, const double* F, double* st, double* sigma) {
  constexpr int DD = DIM * DIM;
  double eps[DD], s[DD];
#pragma unroll
  for (int i = 0; i < DD; i++) eps[i] = F[i];
#pragma unroll
  for (int i = 0; i < DIM; i++)
#pragma unroll
    for (int j = 0; j < i; j++) {
      const double a = 0.5 * (eps[i * DIM + j] + eps[j * DIM + i]);
      eps[j * DIM + i] = eps[i * DIM + j] = a;
    }
#pragma unroll
  for (int i = 0; i < DIM; i++) eps[(DIM + 1) * i] -= 1.;
#pragma unroll
  for (int i = 0; i < DD; i++) eps[i] -= st[i];
  double tr = 0.0;
#pragma unroll
  for (int i = 0; i < DIM; i++) tr += eps[i + i * DIM];
  const double p = m.K * tr;
  const double factor = 2.0 * m.G;
  {
    double trd;
    if (DIM == 2) trd = (eps[0] + eps[3]) / 2.0;
    else trd = (eps[0] + eps[4] + eps[8]) / 3.0;
#pragma unroll
    for (int i = 0; i < DD; i++) s[i] = ((i % (DIM + 1) == 0) ? (eps[i] - trd) : eps[i]) * factor;
  }
  double a = 0.0;
#pragma unroll
  for (int i = 0; i < DD; i++) a += s[i] * s[i];
  const double q = m.sqrt_3_2 * sqrt(a);
  const double eqps_old = st[9];
  const double thermo = hardening_thermo(m, st[10]);
  const RC rc = {q, eqps_old, thermo, false, 0.0};
  const double tolerance = m.hp[0] * 1.e-10;  // SigmaY() * k_tol
  if (j2_residual_d(m, rc, 0.0) > tolerance) {
    const AD1 e0 = {eqps_old, 0.0};
    const double upper = (q - hardening_evaluate(m, e0).v * thermo) / (3.0 * m.G);
    const double de = j2_scalar_solve(m, rc, 0.0, upper, 1.e-10, tolerance, 100);
    const double al = 1.5 / q;
    if (!ACC) {
      const double cc = -2.0 * m.G * de;
#pragma unroll
      for (int i = 0; i < DD; i++) s[i] += cc * (al * s[i]);
    } else {
      st[9] += de;
#pragma unroll
      for (int i = 0; i < DD; i++) st[i] += de * (al * s[i]);
      if (m.hard_kind == HARD_JC_THERMO_RATE) st[10] += m.heat_fraction * q * de / (m.density * m.specific_heat);
    }
  }
  if (!ACC) {
#pragma unroll
    for (int i = 0; i < DD; i++) sigma[i] = s[i] + p * ((i % (DIM + 1) == 0) ? 1.0 : 0.0);
  }
}

// Dev (material_utils.hpp:22-56)
template <int DIM>
__device__ __forceinline__ void dev_mat(const double* A, double factor, double* D) {
  constexpr int DD = DIM * DIM;
  double trd;
  if (DIM == 2) trd = (A[0] + A[3]) / 2.0;
  else trd = (A[0] + A[4] + A[8]) / 3.0;
#pragma unroll
  for (int i = 0; i < DD; i++) D[i] = ((i % (DIM + 1) == 0) ? (A[i] - trd) : A[i]) * factor;
}

// J2Simo::PlasticStress<ACC> (materials.hpp:455-562) on st = {be_old[9], F_old[9], eqps, temperature}; Finv / det_F
// are NonlinearSolidWorkData::FInv / DetF. ACC == false: P out (EvaluatePK1, materials.hpp:564-570). cbrt: libdevice.
template <int DIM, bool ACC>
__device__ __forceinline__ void j2simo_plastic_stress(const MatP& m, const double* F, const double* Finv, double det_F,
                                                   double* st, double* P) {
  constexpr int DD = DIM * DIM;
  double w0[DD], w1[DD], w2[DD];
  double* be_old = st;
  double* F_old = st + 9;
  mult_AB<DIM>(F_old, Finv, w0);
  inverse_small<DIM>(w0, w1);
  const double cb = cbrt(det_small<DIM>(w1));
#pragma unroll
  for (int i = 0; i < DD; i++) w1[i] *= cb;
  mult_AB<DIM>(w1, be_old, w2);
  mult_ABt<DIM>(w2, w1, w0);
  double* be = w0;
  double* s = w1;
  double* Np = w2;
  dev_mat<DIM>(be, m.G, s);
  double a = 0.0;
#pragma unroll
  for (int i = 0; i < DD; i++) a += s[i] * s[i];
  const double s_norm = sqrt(a);
  if (fabs(s_norm) < 2.220446049250313e-16) {
    const double v = sqrt(1. / 2.);
#pragma unroll
    for (int i = 0; i < DD; i++) Np[i] = (i % (DIM + 1) == 0) ? v : 0.0;
  } else {
    const double al = sqrt(3. / 2.) / s_norm;
#pragma unroll
    for (int i = 0; i < DD; i++) Np[i] = al * s[i];
  }
  double s_eff = 0.0;
#pragma unroll
  for (int i = 0; i < DD; i++) s_eff += Np[i] * s[i];
  double be_trace = 0.0;
#pragma unroll
  for (int i = 0; i < DIM; i++) be_trace += be[i + i * DIM];
  const RC rc = {s_eff, st[18], hardening_thermo(m, st[19]), true, be_trace};
  const double tolerance = m.hp[0] * 1.e-10;
  if (j2_residual_d(m, rc, 0.0) > tolerance) {
    const AD1 e0 = {rc.eqps_old, 0.0};
    const double upper = (s_eff - hardening_evaluate(m, e0).v * rc.thermo) / (m.G * be_trace);
    const double de = j2_scalar_solve(m, rc, 0.0, upper, 1.e-10, tolerance, 100);
    const double cc = -2. / 3. * de * be_trace;
#pragma unroll
    for (int i = 0; i < DD; i++) be[i] += cc * Np[i];
    dev_mat<DIM>(be, m.G, s);
    if (ACC) {
      st[18] += de;
      if (m.hard_kind == HARD_JC_THERMO_RATE) st[19] += m.heat_fraction * s_eff * de / (m.density * m.specific_heat);
    }
  }
  if (!ACC) {
    const double al = m.K * (det_F * det_F - 1.) * .5;
    double* tau = w2;
#pragma unroll
    for (int i = 0; i < DD; i++) tau[i] = s[i] + al * ((i % (DIM + 1) == 0) ? 1.0 : 0.0);
    mult_ABt<DIM>(tau, Finv, P);
  } else {
#pragma unroll
    for (int i = 0; i < DD; i++) F_old[i] = F[i];
#pragma unroll
    for (int i = 0; i < DD; i++) be_old[i] = be[i];
  }
}

// J2Log::PlasticStress<ACC> (materials.hpp:636-739) on st = {Fp_inv[9], eqps, temperature}. J2Log has no EvaluatePK1
// of its own: what reaches the residual is MaterialBase::EvaluatePK1 (materials.cpp:60-71) on alternative_stress_ =
// M = s + p / det F * I (materials.hpp:727), P = det F * M * F^-T; the F_e^-T M F_e^T F^-T product written to
// tmp.stress_ (materials.hpp:728-734) is overwritten by that call and the exponential map only feeds it when
// ACC == false, so neither is executed here. log / exp / hypot / acos / cos: libdevice.
template <int DIM, bool ACC>
__device__ __forceinline__ void j2log_plastic_stress(const MatP& m, const double* F, const double* Finv, double det_F,
                                                  double* st, double* P) {
  constexpr int DD = DIM * DIM;
  double w0[DD], w1[DD], V[DD], lam[DIM];
  double* Fp_inv = st;
  mult_AB<DIM>(F, Fp_inv, w0);   // F_e
  mult_AtB<DIM>(w0, w0, w1);     // C_e (LogarithmicStrain, material_utils.hpp:92-114)
  eig_sym<DIM>(w1, lam, V);
#pragma unroll
  for (int i = 0; i < DIM; i++) lam[i] = log(lam[i]);
  mult_ADAt<DIM>(V, lam, w0);    // E_e
#pragma unroll
  for (int i = 0; i < DD; i++) w0[i] *= 0.5;
  double tr = 0.0;
#pragma unroll
  for (int i = 0; i < DIM; i++) tr += w0[i + i * DIM];
  const double p = m.K * tr;
  double* s = w1;
  dev_mat<DIM>(w0, 2.0 * m.G, s);
  double a = 0.0;
#pragma unroll
  for (int i = 0; i < DD; i++) a += s[i] * s[i];
  const double q = m.sqrt_3_2 * sqrt(a);
  const RC rc = {q, st[9], hardening_thermo(m, st[10]), false, 0.0};
  const double tolerance = m.hp[0] * 1.e-10;
  if (j2_residual_d(m, rc, 0.0) > tolerance) {
    const AD1 e0 = {rc.eqps_old, 0.0};
    const double upper = (q - hardening_evaluate(m, e0).v * rc.thermo) / (3.0 * m.G);
    const double de = j2_scalar_solve(m, rc, 0.0, upper, 1.e-10, tolerance, 100);
    const double al = 1.5 / q;
    double* Np = w0;
#pragma unroll
    for (int i = 0; i < DD; i++) Np[i] = al * s[i];
    const double cc = -2.0 * m.G * de;
#pragma unroll
    for (int i = 0; i < DD; i++) s[i] += cc * Np[i];
    if (ACC) {
#pragma unroll
      for (int i = 0; i < DD; i++) Np[i] = -de * Np[i];   // increment
      eig_sym<DIM>(Np, lam, V);
#pragma unroll
      for (int i = 0; i < DIM; i++) lam[i] = exp(lam[i]);
      double ex[DD], old[DD];
      mult_ADAt<DIM>(V, lam, ex);
      st[9] += de;
#pragma unroll
      for (int i = 0; i < DD; i++) old[i] = Fp_inv[i];
      mult_AB<DIM>(old, ex, Fp_inv);
    }
  }
  if (!ACC) {
    const double al = p / det_F;
    double* M = w0;
#pragma unroll
    for (int i = 0; i < DD; i++) M[i] = s[i] + al * ((i % (DIM + 1) == 0) ? 1.0 : 0.0);
    mult_ABt<DIM>(M, Finv, P);
#pragma unroll
    for (int i = 0; i < DD; i++) P[i] *= det_F;
  }
}

// First Piola-Kirchhoff stress P(F). MAT is a compile-time material id; st = this quadrature point's history
// (read-only here; nullptr for stateless materials).
//  NeoHookean: CompressibleOgdenNeoHookean::EvaluateCauchy (src/mimi/materials/materials.cpp:96-118)
//              followed by MaterialBase::EvaluatePK1 (materials.cpp:60-71).
//  StVK:       StVenantKirchhoff::EvaluatePK1 (materials.cpp:73-94).
//  J2Linear:   J2Linear::EvaluateCauchy (materials.hpp:240-243) + MaterialBase::EvaluatePK1.
template <int DIM, int MAT>
__device__ __forceinline__ void material_pk1(const MatP& m, const double* F, const double* st, double* P) {
  constexpr int DD = DIM * DIM;
  const double lambda = m.lambda, mu = m.mu;
  if (MAT == MAT_STVK) {
    double C[DD], E[DD], S[DD];
    mult_AtB<DIM>(F, F, C);
#pragma unroll
    for (int i = 0; i < DD; i++) {
      const double Ii = (i % (DIM + 1) == 0) ? 1.0 : 0.0;
      E[i] = .5 * C[i] + -.5 * Ii;
    }
    double tr = 0.0;
#pragma unroll
    for (int i = 0; i < DIM; i++) tr += E[i + i * DIM];
    const double alpha = lambda * tr;
    const double beta = 2 * mu;
#pragma unroll
    for (int i = 0; i < DD; i++) {
      const double Ii = (i % (DIM + 1) == 0) ? 1.0 : 0.0;
      S[i] = alpha * Ii + beta * E[i];
    }
    mult_AB<DIM>(F, S, P);
    return;
  }
  double B[DD], sig[DD], Finv[DD];
  const double det_F = det_small<DIM>(F);
  if (MAT == MAT_J2LOG) {
    double stl[11];
#pragma unroll
    for (int i = 0; i < DD; i++) stl[i] = st[i];
    stl[9] = st[9];
    stl[10] = st[10];
    inverse_small<DIM>(F, Finv);
    j2log_plastic_stress<DIM, false>(m, F, Finv, det_F, stl, P);
    return;
  }
  if (MAT == MAT_J2SIMO) {
    double stl[20];
#pragma unroll
    for (int i = 0; i < DD; i++) {
      stl[i] = st[i];
      stl[9 + i] = st[9 + i];
    }
    stl[18] = st[18];
    stl[19] = st[19];
    inverse_small<DIM>(F, Finv);
    j2simo_plastic_stress<DIM, false>(m, F, Finv, det_F, stl, P);
    return;
  }
  if (MAT == MAT_J2LINEAR) {
    double stl[19];
#pragma unroll
    for (int i = 0; i < DD; i++) {
      stl[i] = st[i];
      stl[9 + i] = st[9 + i];
    }
    stl[18] = st[18];
    j2linear_plastic_stress<DIM, false>(m, F, stl, sig);
  } else if (MAT == MAT_J2) {
    double stl[11];
#pragma unroll
    for (int i = 0; i < DD; i++) stl[i] = st[i];
    stl[9] = st[9];
    stl[10] = st[10];
    j2_plastic_stress<DIM, false>(m, F, stl, sig);
  } else {
    const double mu_over_det_F = mu / det_F;
    mult_ABt<DIM>(F, F, B);
    const double beta = -mu_over_det_F + lambda * (det_F - 1.);
#pragma unroll
    for (int i = 0; i < DD; i++) {
      const double Ii = (i % (DIM + 1) == 0) ? 1.0 : 0.0;
      sig[i] = mu_over_det_F * B[i] + beta * Ii;
    }
  }
  inverse_small<DIM>(F, Finv);
  mult_ABt<DIM>(sig, Finv, P);
#pragma unroll
  for (int i = 0; i < DD; i++) P[i] *= det_F;
}

}  // namespace mb2
