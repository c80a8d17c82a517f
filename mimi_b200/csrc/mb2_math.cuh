// Device-side small dense kernels and material laws, written in the reference's exact operation
// order so that FP64 results are bit-identical to its MFEM/host path.
// This translation unit set is compiled with -fmad=false: no mul+add may be contracted.
#pragma once
#include <cstdint>

namespace mb2 {

enum : int { MAT_NEOHOOKEAN = 0, MAT_STVK = 1, MAT_J2LINEAR = 2 };

// MaterialBase parameters (materials.hpp:24-40) plus the J2Linear members (materials.hpp:147-166); the sqrt
// constants are computed on the host by std::sqrt exactly like the reference's member initialisers.
struct MatP {
  double lambda = 0, mu = 0;
  double K = 0, G = 0, iso = 0, kin = 0, sigma_y = 0, sqrt_6 = 0, sqrt_3_2 = 0, sqrt_2_3 = 0;
};

// per-quadrature-point history doubles: J2Linear = plastic_strain[9], beta[9], accumulated_plastic_strain
__host__ __device__ constexpr int mat_state_size(int mat) { return mat == MAT_J2LINEAR ? 19 : 0; }

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
  if (MAT == MAT_J2LINEAR) {
    double stl[19];
#pragma unroll
    for (int i = 0; i < DD; i++) {
      stl[i] = st[i];
      stl[9 + i] = st[9 + i];
    }
    stl[18] = st[18];
    j2linear_plastic_stress<DIM, false>(m, F, stl, sig);
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
