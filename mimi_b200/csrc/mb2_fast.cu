// MB2_TANGENT_ANALYTIC -- the non-parity "fast" tangent mode (north_star part 3).
//
// The reference's tangent is a forward difference of the element residual (nonlinear_solid.cpp:48-76); the parity
// kernels of mb2_assembly.cu reproduce it bit for bit and pay 82 residual evaluations per element for that. This file
// assembles what that difference approximates,
//     K_e[(i,j),(l,m)] = sum_q sum_{k,n} dN(q;i,k) * (w detJ)_q dP_jk/dF_mn (F_q) * dN(q;l,n)        ("B^T C B"),
// from the closed-form material tangent dP/dF of CompressibleOgdenNeoHookean (materials.cpp:96-118 + 60-71: P = mu F +
// (-mu + lambda J (J - 1)) F^-T) and StVenantKirchhoff (materials.cpp:73-94: P = F S). It is NOT bit-comparable with the
// reference (FMA contraction allowed, other summation order, no O(h) truncation / O(eps/h) rounding of the difference
// quotient); it is graded against the reference's K at the finite-difference noise level and by identical Newton
// iteration counts (tests/test_gpu_fast.py). This is the one translation unit compiled WITHOUT -fmad=false.
//
// p = 2 hexahedra (27 nodes, 64 points): the contraction is one GEMM per element,
//     Out[(jm,i), l] = sum_{(q,n)} T[(jm,i), (q,n)] * dN[(q,n), l],   T = sum_k dN(q;i,k) C_q[(j,k),(m,n)],
// 162 x 192 x 27 for the 6 component pairs j <= m (C has the major symmetry, so K^{mj} = (K^{jm})^T), run either on the
// FP64 tensor cores (mma.sync.m8n8k4.f64: "DMMA") or as a register-tiled DFMA loop on the CUDA cores; which one is the
// library's choice was decided by ncu (profiles/r02_fast_*). Other element types use the generic kernel below.
#include <cstdio>

#include "mb2_ctx.h"

namespace mb2 {

// P = first Piola-Kirchhoff stress, C[(j + DIM k) + DD (m + DIM n)] = dP_jk / dF_mn
template <int DIM, int MAT>
__device__ __forceinline__ void pk1_and_tangent(const MatP& mp, const double* F, double* P, double* C) {
  constexpr int DD = DIM * DIM;
  const double lambda = mp.lambda, mu = mp.mu;
  if (MAT == MAT_STVK) {
    // S = lambda tr(E) I + 2 mu E, E = (F^T F - I) / 2, P = F S
    // dP_jk/dF_mn = delta_jm S_nk + lambda F_jk F_mn + mu (F_jn F_mk + (F F^T)_jm delta_kn)
    double S[DD], B[DD];
    double tr = 0.0;
#pragma unroll
    for (int a = 0; a < DIM; a++)
#pragma unroll
      for (int b = 0; b < DIM; b++) {
        double cab = 0.0, bab = 0.0;
#pragma unroll
        for (int p = 0; p < DIM; p++) {
          cab += F[p + DIM * a] * F[p + DIM * b];   // (F^T F)_ab
          bab += F[a + DIM * p] * F[b + DIM * p];   // (F F^T)_ab
        }
        const double eab = 0.5 * (cab - (a == b ? 1.0 : 0.0));
        S[a + DIM * b] = 2.0 * mu * eab;
        B[a + DIM * b] = bab;
        if (a == b) tr += eab;
      }
#pragma unroll
    for (int a = 0; a < DIM; a++) S[a + DIM * a] += lambda * tr;
#pragma unroll
    for (int j = 0; j < DIM; j++)
#pragma unroll
      for (int k = 0; k < DIM; k++) {
        double p = 0.0;
#pragma unroll
        for (int a = 0; a < DIM; a++) p += F[j + DIM * a] * S[a + DIM * k];
        P[j + DIM * k] = p;
      }
#pragma unroll
    for (int n = 0; n < DIM; n++)
#pragma unroll
      for (int m = 0; m < DIM; m++)
#pragma unroll
        for (int k = 0; k < DIM; k++)
#pragma unroll
          for (int j = 0; j < DIM; j++) {
            double c = lambda * F[j + DIM * k] * F[m + DIM * n] + mu * F[j + DIM * n] * F[m + DIM * k];
            if (j == m) c += S[n + DIM * k];
            if (k == n) c += mu * B[j + DIM * m];
            C[(j + DIM * k) + DD * (m + DIM * n)] = c;
          }
    return;
  }
  // NeoHookean: P = mu F + g G, G = F^-T, g = -mu + lambda J (J - 1)
  // dP_jk/dF_mn = mu delta_jm delta_kn + lambda (2J - 1) J G_jk G_mn - g G_jn G_mk
  double Finv[DD], G[DD];
  const double J = det_small<DIM>(F);
  inverse_small<DIM>(F, Finv);
#pragma unroll
  for (int j = 0; j < DIM; j++)
#pragma unroll
    for (int k = 0; k < DIM; k++) G[j + DIM * k] = Finv[k + DIM * j];
  const double g = -mu + lambda * J * (J - 1.0);
  const double c1 = lambda * (2.0 * J - 1.0) * J;
#pragma unroll
  for (int i = 0; i < DD; i++) P[i] = mu * F[i] + g * G[i];
#pragma unroll
  for (int n = 0; n < DIM; n++)
#pragma unroll
    for (int m = 0; m < DIM; m++)
#pragma unroll
      for (int k = 0; k < DIM; k++)
#pragma unroll
        for (int j = 0; j < DIM; j++) {
          double c = c1 * G[j + DIM * k] * G[m + DIM * n] - g * G[j + DIM * n] * G[m + DIM * k];
          if (j == m && k == n) c += mu;
          C[(j + DIM * k) + DD * (m + DIM * n)] = c;
        }
}

__host__ __device__ constexpr int largest_divisor_le(int n, int cap) {
  int best = 1;
  for (int d = 1; d <= cap; d++)
    if (n % d == 0) best = d;
  return best;
}

// ---------------------------------------------------------------------------------------------
// Generic analytic kernel: any (DIM, ND, NQ). One CTA per element. Thread (row r = (i, j), column group g) owns CB
// consecutive columns (l, m) of one component m per pass; the quadrature points are streamed in chunks of QC.
// ---------------------------------------------------------------------------------------------
template <int DIM, int ND, int NQ>
struct FastGenCfg {
  static constexpr int TD = DIM * ND, DD = DIM * DIM;
  static constexpr int G = (256 / TD) > 0 ? (256 / TD) : 1;          // column groups handled concurrently
  static constexpr int NT = (TD * G + 31) / 32 * 32;
  static constexpr int CB = largest_divisor_le(ND, 16);              // columns per thread and pass (never straddles m)
  static constexpr int NCB = TD / CB;                                // column blocks
  static constexpr int NPASS = (NCB + G - 1) / G;
  static constexpr int QC = NQ < 8 ? NQ : 8;
  static constexpr int NCH = (NQ + QC - 1) / QC;
  static constexpr int KP = (TD * DIM + 3) / 4 * 4;
  static constexpr int OFF_UE = 0, OFF_DN = OFF_UE + TD, OFF_C = OFF_DN + QC * TD, OFF_B = OFF_C + QC * DD * DD;
  static constexpr int SMEM_DOUBLES = OFF_B + QC * DD;
};

template <int DIM, int ND, int NQ, int MAT>
__global__ void __launch_bounds__(FastGenCfg<DIM, ND, NQ>::NT) element_analytic_kernel(const ElemArgs g) {
  using Cfg = FastGenCfg<DIM, ND, NQ>;
  constexpr int TD = Cfg::TD, DD = Cfg::DD, NT = Cfg::NT, CB = Cfg::CB, QC = Cfg::QC, KP = Cfg::KP;
  __shared__ double smem[Cfg::SMEM_DOUBLES];
  double* sue = smem + Cfg::OFF_UE;
  double* sdN = smem + Cfg::OFF_DN;   // [QC][DIM][ND]  (the global layout of QuadData::dN_dX)
  double* sC = smem + Cfg::OFF_C;     // [QC][DD*DD]   (w detJ) dP/dF
  double* sb = smem + Cfg::OFF_B;     // [QC][DD]      (w detJ) P
  const int t = threadIdx.x;
  const int64_t e = g.e0 + blockIdx.x;
  const double* dNe = g.dNdX + e * (int64_t)(NQ * TD);
  if (t < TD) sue[t] = g.u[g.vdofs[e * TD + t]];
  const int r = t % TD, i = r % ND, j = r / ND;
  double racc = 0.0;
#pragma unroll 1
  for (int pass = 0; pass < Cfg::NPASS; pass++) {
    const int cbi = pass * Cfg::G + t / TD;
    const bool active = (t < TD * Cfg::G) && (cbi < Cfg::NCB);
    const int m = active ? (cbi * CB) / ND : 0, l0 = active ? (cbi * CB) % ND : 0;
    double acc[CB];
#pragma unroll
    for (int c = 0; c < CB; c++) acc[c] = 0.0;
#pragma unroll 1
    for (int ch = 0; ch < Cfg::NCH; ch++) {
      const int q0 = ch * QC;
      const int nqc = (NQ - q0 < QC) ? NQ - q0 : QC;
      __syncthreads();
      for (int idx = t; idx < nqc * TD; idx += NT) sdN[idx] = dNe[q0 * TD + idx];
      __syncthreads();
      if (t < nqc) {
        const double* dN = sdN + t * TD;
        double F[DD], P[DD];
        double* C = sC + t * DD * DD;   // written in place: 81 doubles per thread do not fit the register file
#pragma unroll
        for (int b = 0; b < DIM; b++)
#pragma unroll
          for (int a = 0; a < DIM; a++) {
            double d = (a == b) ? 1.0 : 0.0;
            for (int k = 0; k < ND; k++) d += sue[k + a * ND] * dN[k + b * ND];
            F[a + DIM * b] = d;
          }
        pk1_and_tangent<DIM, MAT>(g.mp, F, P, C);
        const double aq = g.wdetJ[e * NQ + q0 + t];
#pragma unroll
        for (int x = 0; x < DD * DD; x++) C[x] *= aq;
#pragma unroll
        for (int x = 0; x < DD; x++) sb[t * DD + x] = aq * P[x];
      }
      __syncthreads();
      if (pass == 0 && t < TD) {
        for (int qs = 0; qs < nqc; qs++)
#pragma unroll
          for (int k = 0; k < DIM; k++) racc += sdN[qs * TD + i + k * ND] * sb[qs * DD + j + DIM * k];
      }
      if (active) {
        for (int qs = 0; qs < nqc; qs++) {
          const double* dN = sdN + qs * TD;
          const double* Cq = sC + qs * DD * DD;
          double tn[DIM];
#pragma unroll
          for (int n = 0; n < DIM; n++) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < DIM; k++) s += dN[i + k * ND] * Cq[(j + DIM * k) + DD * (m + DIM * n)];
            tn[n] = s;
          }
#pragma unroll
          for (int c = 0; c < CB; c++)
#pragma unroll
            for (int n = 0; n < DIM; n++) acc[c] += tn[n] * dN[l0 + c + n * ND];
        }
      }
    }
    if (active) {
      // K_e block layout of mb2_assembly.cu: [row node i][KP], column = perturbed dof (l + m ND) * DIM + row component j
      double* out = g.KeT + (e - g.e0) * (int64_t)(ND * KP) + (int64_t)i * KP;
#pragma unroll
      for (int c = 0; c < CB; c++) out[(l0 + c + m * ND) * DIM + j] = acc[c];
    }
  }
  if (t < TD) g.Re[e * TD + t] = racc;
}

// ---------------------------------------------------------------------------------------------
// p = 2 hexahedra: 27 nodes, 64 quadrature points.
// ---------------------------------------------------------------------------------------------
struct FastP2 {
  static constexpr int DIM = 3, ND = 27, NQ = 64, TD = 81, DD = 9;
  static constexpr int NT = 256, NW = 8;
  static constexpr int NDP = 28;                // dN row pitch: conflict-free B fragments need NDP = 4 or 12 (mod 16)
  static constexpr int NJM = 6;                 // component pairs j <= m
  static constexpr int MR = NJM * ND;           // 162 GEMM rows (jm, i)
  static constexpr int MT = (MR + 7) / 8;       // 21 row tiles
  static constexpr int NTL = 4;                 // column tiles (l = 0..31, 27 used)
  static constexpr int QC = 4, KC = QC * DIM;   // chunk of quadrature points; GEMM depth per chunk
  static constexpr int LDR = 172;               // T is kept k-major, T[kk][row]: pitch = 4 or 12 (mod 16) for conflict-free A
                                                // fragments, and consecutive rows are consecutive words for the T-phase stores
  static constexpr int TBUF = KC * LDR;
  static constexpr int NCH = NQ / QC;
  static constexpr int KP = (TD * DIM + 3) / 4 * 4;  // 244
  static constexpr int OFF_DN = 0;                               // [NQ*DIM][NDP] + 4 pad
  static constexpr int OFF_C = OFF_DN + NQ * DIM * NDP + 4;      // [NQ][NJM][9]   (w detJ) C^{jm}[k][n]
  static constexpr int OFF_T = OFF_C + NQ * NJM * 9;             // [2][KC][LDR]
  static constexpr int OFF_B = OFF_T + 2 * TBUF;                 // [NQ][9]  (w detJ) P
  static constexpr int OFF_UE = OFF_B + NQ * 9;                  // [81]
  static constexpr int SMEM_DOUBLES = OFF_UE + 84;
  static constexpr size_t SMEM_BYTES = sizeof(double) * SMEM_DOUBLES;
  static_assert(NQ * NJM * 9 + 2 * TBUF >= ND * KP, "K_e staging aliases the C and T buffers");
  static_assert(OFF_C % 2 == 0 && OFF_T % 2 == 0 && LDR >= MT * 8, "alignment / size");
  static_assert(2 * (SMEM_BYTES + 1024) <= 227 * 1024, "two CTAs per SM");
};

__device__ __forceinline__ void dmma_m8n8k4(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

// component pair index -> (j, m), j <= m
__device__ __forceinline__ void jm_of(int p, int& j, int& m) {
  j = (p < 3) ? 0 : (p < 5 ? 1 : 2);
  m = (p < 3) ? p : (p < 5 ? p - 2 : 2);
}

// Third form of the p = 2 contraction (the library's default): the A fragments of the tensor-core GEMM are built in
// REGISTERS. ncu of the second form (profiles/r02_fast_dmma_vs_dfma_summary.txt): 36 % of the DMMA peak, the T = dN C stage
// between the MMA bursts 28 % of the stall samples, the CTA barrier per chunk 19 %, shared memory the busiest pipe (22
// LDS / STS per k-step). With the GEMM depth ordered (n, q) -- a k-step = one n, four consecutive points -- a thread's A
// element T[(p,i),(q,n)] = sum_k dN(q;i,k) C_q^p[k][n] needs the three dN(q;i,:) it keeps for all three n and three C
// values: no T buffers, no barrier inside the loop (warps run through the 64 points independently), 12 LDS per k-step.
struct FastP2V3 {
  static constexpr int DIM = 3, ND = 27, NQ = 64, TD = 81, NDP = 28, NJM = 6, MR = NJM * ND, MT = (MR + 7) / 8, NTL = 4;
  static constexpr int CQP = 74;  // pitch of one point's C block [p][n][4]: 72 + 2, so that the LDS.128 of the four points
                                  // of a k-step fall into four different bank groups
  static constexpr int KP = (TD * DIM + 3) / 4 * 4;
  static constexpr int OFF_DN = 0;                               // [NQ*DIM][NDP] + 4 pad
  static constexpr int OFF_C = OFF_DN + NQ * DIM * NDP + 4;      // [NQ][CQP]
  static constexpr int OFF_B = OFF_C + NQ * CQP;                 // [NQ][9]  (w detJ) P
  static constexpr int OFF_UE = OFF_B + NQ * 9;                  // [81]
  static constexpr int OFF_SCR = OFF_UE + 84;                    // per-point stage scratch
  static constexpr int SMEM_DOUBLES = OFF_SCR + 4 * NQ * 9 + 2 * NQ;
  static constexpr size_t SMEM_BYTES = sizeof(double) * SMEM_DOUBLES;
  static_assert(OFF_B >= ND * KP, "K_e staging aliases the dN and C buffers");
  static_assert(OFF_C % 2 == 0 && CQP % 2 == 0, "16-byte aligned C rows");
  static_assert(2 * (SMEM_BYTES + 1024) <= 227 * 1024, "two CTAs per SM");
};

// Everything before the contraction, shared by the p = 2 kernels: staging, the three per-point stages, the residual.
// scr: 4 * 64 * 9 + 2 * 64 doubles of scratch. Ends without a barrier: sdN, sC, sb are complete (the last barrier follows
// the C stage); the residual only reads them.
template <int MAT, bool V3>
__device__ __forceinline__ void fast_p2_prepare(const ElemArgs& g, const int64_t e, const int t, double* sdN, double* sC, double* sb,
                                                double* sue, double* scr) {
  constexpr int DIM = 3, ND = 27, NQ = 64, TD = 81, DD = 9, NDP = FastP2::NDP;
  // ---- stage the element
  if (t < TD) sue[t] = g.u[g.vdofs[e * TD + t]];
  {
    const double* src = g.dNdX + e * (int64_t)(NQ * TD);
    for (int idx = t; idx < NQ * TD; idx += 256) {
      const int row = idx / ND, i = idx - row * ND;  // row = q*DIM + k
      sdN[row * NDP + i] = __ldg(src + idx);
    }
    for (int row = t; row < NQ * DIM; row += 256) sdN[row * NDP + ND] = 0.0;
    if (t < 4) sdN[NQ * DIM * NDP + t] = 0.0;
  }
  __syncthreads();
  // ---- per quadrature point, in three stages that keep all 256 threads busy (the stress + tangent of one point is
  // ~1 kflop: one thread per point left 6 of 8 warps waiting at a barrier, ncu profiles/r02_fast_dmma_v1_summary.txt):
  //  (1) F entries, one per thread:  F(a, b) = delta_ab + sum_k u_a[k] dN(q; k, b)
  //  (2) one thread per point: P and the point's ingredients of dP/dF in the unified form
  //        (w detJ) dP_jk/dF_mn = alpha X_jk X_mn + beta X_jn X_mk + delta_jm Y_nk + Z_jm delta_kn
  //      NeoHookean: X = F^-T, alpha = w lambda (2J-1) J, beta = -w g, Y = w mu I, Z = 0;  StVK: X = F, alpha = w lambda,
  //      beta = w mu, Y = w S, Z = w mu F F^T   (w = w detJ; the stages' scratch aliases the T buffers)
  //  (3) one thread per (point, component pair j <= m): the 9 entries C^{jm}[k][n]
  double* sF = scr;                      // [NQ][9]
  double* sX = scr + NQ * 9;              // [NQ][9]
  double* sY = scr + 2 * NQ * 9;          // [NQ][9]
  double* sZ = scr + 3 * NQ * 9;          // [NQ][9]
  double* sab = scr + 4 * NQ * 9;         // [NQ][2]
  for (int task = t; task < NQ * DD; task += 256) {
    const int q = task / DD, x = task - q * DD;
    const int a = x % DIM, b = x / DIM;
    double d = (a == b) ? 1.0 : 0.0;
    const double* dN = sdN + (q * DIM + b) * NDP;
    const double* ua = sue + a * ND;
#pragma unroll
    for (int k = 0; k < ND; k++) d += ua[k] * dN[k];
    sF[task] = d;
  }
  __syncthreads();
  if (t < NQ) {
    const double lambda = g.mp.lambda, mu = g.mp.mu;
    const double aq = g.wdetJ[e * NQ + t];
    double F[DD], P[DD];
#pragma unroll
    for (int x = 0; x < DD; x++) F[x] = sF[t * DD + x];
    if (MAT == MAT_STVK) {
      double S[DD], B[DD];
      double tr = 0.0;
#pragma unroll
      for (int a = 0; a < DIM; a++)
#pragma unroll
        for (int b = 0; b < DIM; b++) {
          double cab = 0.0, bab = 0.0;
#pragma unroll
          for (int pp = 0; pp < DIM; pp++) {
            cab += F[pp + DIM * a] * F[pp + DIM * b];
            bab += F[a + DIM * pp] * F[b + DIM * pp];
          }
          const double eab = 0.5 * (cab - (a == b ? 1.0 : 0.0));
          S[a + DIM * b] = 2.0 * mu * eab;
          B[a + DIM * b] = bab;
          if (a == b) tr += eab;
        }
#pragma unroll
      for (int a = 0; a < DIM; a++) S[a + DIM * a] += lambda * tr;
#pragma unroll
      for (int j = 0; j < DIM; j++)
#pragma unroll
        for (int k = 0; k < DIM; k++) {
          double pv = 0.0;
#pragma unroll
          for (int a = 0; a < DIM; a++) pv += F[j + DIM * a] * S[a + DIM * k];
          P[j + DIM * k] = pv;
        }
#pragma unroll
      for (int x = 0; x < DD; x++) {
        sX[t * DD + x] = F[x];
        sY[t * DD + x] = aq * S[x];
        sZ[t * DD + x] = aq * mu * B[x];
      }
      sab[2 * t] = aq * lambda;
      sab[2 * t + 1] = aq * mu;
    } else {
      double Finv[DD];
      const double J = det_small<DIM>(F);
      inverse_small<DIM>(F, Finv);
      const double gg = -mu + lambda * J * (J - 1.0);
#pragma unroll
      for (int j = 0; j < DIM; j++)
#pragma unroll
        for (int k = 0; k < DIM; k++) {
          const double G = Finv[k + DIM * j];
          P[j + DIM * k] = mu * F[j + DIM * k] + gg * G;
          sX[t * DD + j + DIM * k] = G;
          sY[t * DD + j + DIM * k] = (j == k) ? aq * mu : 0.0;
          sZ[t * DD + j + DIM * k] = 0.0;
        }
      sab[2 * t] = aq * lambda * (2.0 * J - 1.0) * J;
      sab[2 * t + 1] = -aq * gg;
    }
#pragma unroll
    for (int x = 0; x < DD; x++) sb[t * DD + x] = aq * P[x];
  }
  __syncthreads();
  for (int task = t; task < NQ * FastP2::NJM; task += 256) {
    const int q = task / FastP2::NJM, p = task - q * FastP2::NJM;
    int j, m;
    jm_of(p, j, m);
    const double* X = sX + q * DD;
    const double* Y = sY + q * DD;
    const double alpha = sab[2 * q], beta = sab[2 * q + 1];
    const double zjm = sZ[q * DD + j + DIM * m];
    double xj[DIM], xm[DIM];
#pragma unroll
    for (int k = 0; k < DIM; k++) {
      xj[k] = X[j + DIM * k];
      xm[k] = X[m + DIM * k];
    }
    // C layout: V3 = [q] pitch CQP, [(p, n)][4] (k contiguous: one LDS.128 + one LDS.64 per A-fragment element);
    // otherwise [q][p][k][n]
    double* o = V3 ? sC + q * FastP2V3::CQP + p * 12 : sC + task * 9;
#pragma unroll
    for (int k = 0; k < DIM; k++)
#pragma unroll
      for (int n = 0; n < DIM; n++) {
        double c = alpha * xj[k] * xm[n] + beta * xj[n] * xm[k];
        if (j == m) c += Y[n + DIM * k];
        if (k == n) c += zjm;
        o[V3 ? n * 4 + k : k * 3 + n] = c;
      }
  }
  __syncthreads();
  // ---- residual R(i, j) = sum_q sum_k dN(q; i, k) (w detJ P)(j, k)
  if (t < TD) {
    const int i = t % ND, j = t / ND;
    double r = 0.0;
#pragma unroll 4
    for (int q = 0; q < NQ; q++)
#pragma unroll
      for (int k = 0; k < DIM; k++) r += sdN[(q * DIM + k) * NDP + i] * sb[q * DD + j + DIM * k];
    g.Re[e * TD + t] = r;
  }

}

template <int MAT, int GEMM>  // GEMM: 0 = FP64 tensor cores (DMMA), 1 = CUDA-core DFMA register tiles
__global__ void __launch_bounds__(256, 2) element_fast_p2_kernel(const ElemArgs g) {
  using Cfg = FastP2;
  constexpr int DIM = 3, ND = 27, NQ = 64, TD = 81, DD = 9, NDP = Cfg::NDP, LDR = Cfg::LDR, QC = Cfg::QC, KP = Cfg::KP;
  extern __shared__ __align__(16) double smem[];
  double* sdN = smem + Cfg::OFF_DN;
  double* sC = smem + Cfg::OFF_C;
  double* sT = smem + Cfg::OFF_T;
  double* sb = smem + Cfg::OFF_B;
  double* sue = smem + Cfg::OFF_UE;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int64_t e = g.e0 + blockIdx.x;

  fast_p2_prepare<MAT, false>(g, e, t, sdN, sC, sb, sue, sT);   // the stages' scratch aliases the T buffers
  // the rows 162 .. LDR-1 of both T buffers feed the padding rows of the last row tiles: zero once
  for (int idx = t; idx < 2 * Cfg::KC * (Cfg::LDR - Cfg::MR); idx += 256) {
    const int kk = idx / (Cfg::LDR - Cfg::MR), x = idx - kk * (Cfg::LDR - Cfg::MR);
    sT[kk * Cfg::LDR + Cfg::MR + x] = 0.0;
  }
  // T phase of chunk ch: T[(qs, n)][(p, i)] = sum_k dN(q; i, k) C_q^{p}[k][n]; task = (p, qs) per warp, lanes over the node i
  auto t_phase = [&](int ch) {
    double* Tb = sT + (ch & 1) * Cfg::TBUF;
    const int i = lane;
#pragma unroll
    for (int r = 0; r < Cfg::NJM * QC / 8; r++) {
      const int pq = warp + 8 * r;
      const int p = pq / QC, qs = pq - p * QC;
      const int q = ch * QC + qs;
      const double* c = sC + (q * Cfg::NJM + p) * 9;
      if (i < ND) {
        const double d0 = sdN[(q * DIM + 0) * NDP + i], d1 = sdN[(q * DIM + 1) * NDP + i], d2 = sdN[(q * DIM + 2) * NDP + i];
        double* o = Tb + (qs * DIM) * LDR + p * ND + i;
#pragma unroll
        for (int n = 0; n < DIM; n++) o[n * LDR] = d0 * c[n] + d1 * c[3 + n] + d2 * c[6 + n];
      }
    }
  };
  static_assert((Cfg::NJM * QC) % 8 == 0, "T-phase tasks per warp");

  if (GEMM == 0) {
    // ---- DMMA: warp w owns row tiles w, w + 8, w + 16 (the last only for w < 5) x all 4 column tiles
    const int gid = lane >> 2, tig = lane & 3;
    constexpr int NMT = 3;
    double acc[NMT][Cfg::NTL][2];
#pragma unroll
    for (int a = 0; a < NMT; a++)
#pragma unroll
      for (int b = 0; b < Cfg::NTL; b++) acc[a][b][0] = acc[a][b][1] = 0.0;
    const bool has3 = warp + 16 < Cfg::MT;   // the third tile of warps 5..7 is padding: computed (keeps the MMA loop
                                             // branch-free and the sub-partitions balanced), never stored
    t_phase(0);
    __syncthreads();
#pragma unroll 1
    for (int ch = 0; ch < Cfg::NCH; ch++) {
      if (ch + 1 < Cfg::NCH) t_phase(ch + 1);
      const double* Tb = sT + (ch & 1) * Cfg::TBUF;
#pragma unroll
      for (int ks = 0; ks < Cfg::KC / 4; ks++) {
        double bf[Cfg::NTL], af[NMT];
        // B(k, col) = dN[(q, n) = chunk row 4 ks + tig][l = 8 nt + gid]
        const double* brow = sdN + (ch * Cfg::KC + ks * 4 + tig) * NDP + gid;
#pragma unroll
        for (int nt = 0; nt < Cfg::NTL; nt++) bf[nt] = brow[nt * 8];
#pragma unroll
        for (int a = 0; a < NMT; a++) af[a] = Tb[(ks * 4 + tig) * LDR + (warp + 8 * a) * 8 + gid];
#pragma unroll
        for (int a = 0; a < NMT; a++)
#pragma unroll
          for (int nt = 0; nt < Cfg::NTL; nt++) dmma_m8n8k4(acc[a][nt][0], acc[a][nt][1], af[a], bf[nt]);
      }
      __syncthreads();
    }
    // ---- K_e staged in shared memory (aliases C and T), K^{mj} = (K^{jm})^T mirrored
    double* sK = sC;
#pragma unroll
    for (int a = 0; a < NMT; a++) {
      if (a < 2 || has3) {
        const int row = (warp + 8 * a) * 8 + gid;
        if (row < Cfg::MR) {
          const int p = row / ND, i = row - p * ND;
          int j, m;
          jm_of(p, j, m);
#pragma unroll
          for (int nt = 0; nt < Cfg::NTL; nt++)
#pragma unroll
            for (int x = 0; x < 2; x++) {
              const int l = nt * 8 + tig * 2 + x;
              if (l < ND) {
                const double v = acc[a][nt][x];
                sK[i * KP + (l + m * ND) * DIM + j] = v;
                if (j != m) sK[l * KP + (i + j * ND) * DIM + m] = v;
              }
            }
        }
      }
    }
  } else {
    // ---- CUDA cores: thread (row group of 6, column group of 3) = 27 x 9 = 243 threads, 18 DFMA accumulators
    const int rg = t / 9, cg = t - rg * 9;
    const bool act = t < 243;
    double acc[6][3];
#pragma unroll
    for (int a = 0; a < 6; a++)
#pragma unroll
      for (int b = 0; b < 3; b++) acc[a][b] = 0.0;
    t_phase(0);
    __syncthreads();
#pragma unroll 1
    for (int ch = 0; ch < Cfg::NCH; ch++) {
      if (ch + 1 < Cfg::NCH) t_phase(ch + 1);
      const double* Tb = sT + (ch & 1) * Cfg::TBUF;
      if (act) {
#pragma unroll
        for (int kk = 0; kk < Cfg::KC; kk++) {
          const double* brow = sdN + (ch * Cfg::KC + kk) * NDP + cg * 3;
          const double b0 = brow[0], b1 = brow[1], b2 = brow[2];
#pragma unroll
          for (int a = 0; a < 6; a++) {
            const double av = Tb[kk * LDR + rg * 6 + a];
            acc[a][0] += av * b0;
            acc[a][1] += av * b1;
            acc[a][2] += av * b2;
          }
        }
      }
      __syncthreads();
    }
    double* sK = sC;
    if (act) {
#pragma unroll
      for (int a = 0; a < 6; a++) {
        const int row = rg * 6 + a;
        const int p = row / ND, i = row - p * ND;
        int j, m;
        jm_of(p, j, m);
#pragma unroll
        for (int b = 0; b < 3; b++) {
          const int l = cg * 3 + b;
          const double v = acc[a][b];
          sK[i * KP + (l + m * ND) * DIM + j] = v;
          if (j != m) sK[l * KP + (i + j * ND) * DIM + m] = v;
        }
      }
    }
  }
  __syncthreads();
  // ---- coalesced 16-byte stores of the element matrix (pad column KP - 1 carries a zero)
  {
    const double* sK = sC;
    double2* out = reinterpret_cast<double2*>(g.KeT + (e - g.e0) * (int64_t)(ND * KP));
    for (int idx = t; idx < ND * KP / 2; idx += 256) {
      const int col = (2 * idx) % KP;
      double2 v = *reinterpret_cast<const double2*>(sK + 2 * idx);
      if (col + 1 >= TD * DIM) v.y = 0.0;
      out[idx] = v;
    }
  }
}

template <int MAT>
__global__ void __launch_bounds__(256, 2) element_fast_p2v3_kernel(const __grid_constant__ ElemArgs g) {
  using Cfg = FastP2V3;
  constexpr int DIM = 3, ND = 27, TD = 81, NDP = Cfg::NDP, KP = Cfg::KP;
  extern __shared__ __align__(16) double smem[];
  double* sdN = smem + Cfg::OFF_DN;
  double* sC = smem + Cfg::OFF_C;
  double* sb = smem + Cfg::OFF_B;
  double* sue = smem + Cfg::OFF_UE;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int64_t e = g.e0 + blockIdx.x;
  fast_p2_prepare<MAT, true>(g, e, t, sdN, sC, sb, sue, smem + Cfg::OFF_SCR);

  // warp w owns row tiles w, w + 8, w + 16 x all 4 column tiles; row = (p, i), rows >= 162 read the zero pad column of dN.
  // Measured and dropped (48^3 elements, this form 13.48 ms): rows grouped so that a warp serves one component pair and
  // loads its C values once per k-step (-20 % shared-memory wavefronts) with the residual spread over 243 threads: 13.76 ms;
  // the three pure-padding row tiles of warps 5 .. 7 skipped behind warp-uniform branches: 13.95 ms. The loop waits for the
  // FP64 / tensor pipe ("math" is the top stall), not for shared memory; profiles/r02_fast_v3_summary.txt.
  const int gid = lane >> 2, tig = lane & 3;
  constexpr int NMT = 3;
  int ia[NMT], pa12[NMT];
#pragma unroll
  for (int a = 0; a < NMT; a++) {
    const int row = (warp + 8 * a) * 8 + gid;
    const int p = row < Cfg::MR ? row / ND : 0;
    ia[a] = row < Cfg::MR ? row - p * ND : ND;
    pa12[a] = p * 12;
  }
  double acc[NMT][Cfg::NTL][2];
#pragma unroll
  for (int a = 0; a < NMT; a++)
#pragma unroll
    for (int b = 0; b < Cfg::NTL; b++) acc[a][b][0] = acc[a][b][1] = 0.0;
#pragma unroll 2
  for (int qg = 0; qg < Cfg::NQ / 4; qg++) {
    const int q = qg * 4 + tig;                 // the k index of a fragment element is the point
    const double* dq = sdN + q * (DIM * NDP);
    const double* cq = sC + q * Cfg::CQP;
    double dn[NMT][DIM];
#pragma unroll
    for (int a = 0; a < NMT; a++)
#pragma unroll
      for (int k = 0; k < DIM; k++) dn[a][k] = dq[k * NDP + ia[a]];
#pragma unroll
    for (int n = 0; n < DIM; n++) {
      double af[NMT], bf[Cfg::NTL];
#pragma unroll
      for (int a = 0; a < NMT; a++) {
        const double2 c01 = *reinterpret_cast<const double2*>(cq + pa12[a] + n * 4);
        const double c2 = cq[pa12[a] + n * 4 + 2];
        af[a] = dn[a][0] * c01.x + dn[a][1] * c01.y + dn[a][2] * c2;
      }
      // B(k, col) = dN(q; l = 8 nt + gid, n)
      const double* brow = dq + n * NDP + gid;
#pragma unroll
      for (int nt = 0; nt < Cfg::NTL; nt++) bf[nt] = brow[nt * 8];
#pragma unroll
      for (int a = 0; a < NMT; a++)
#pragma unroll
        for (int nt = 0; nt < Cfg::NTL; nt++) dmma_m8n8k4(acc[a][nt][0], acc[a][nt][1], af[a], bf[nt]);
    }
  }
  __syncthreads();  // every warp has left the loop (and the residual): dN and C may be overwritten
  // ---- K_e staged in shared memory (aliases dN and C), K^{mj} = (K^{jm})^T mirrored
  double* sK = smem;
#pragma unroll
  for (int a = 0; a < NMT; a++) {
    const int row = (warp + 8 * a) * 8 + gid;
    if (row < Cfg::MR) {
      const int p = row / ND, i = row - p * ND;
      int j, m;
      jm_of(p, j, m);
#pragma unroll
      for (int nt = 0; nt < Cfg::NTL; nt++)
#pragma unroll
        for (int x = 0; x < 2; x++) {
          const int l = nt * 8 + tig * 2 + x;
          if (l < ND) {
            const double v = acc[a][nt][x];
            sK[i * KP + (l + m * ND) * DIM + j] = v;
            if (j != m) sK[l * KP + (i + j * ND) * DIM + m] = v;
          }
        }
    }
  }
  __syncthreads();
  // ---- coalesced 16-byte stores of the element matrix (pad column KP - 1 carries a zero)
  {
    double2* out = reinterpret_cast<double2*>(g.KeT + (e - g.e0) * (int64_t)(ND * KP));
    for (int idx = t; idx < ND * KP / 2; idx += 256) {
      const int col = (2 * idx) % KP;
      double2 v = *reinterpret_cast<const double2*>(sK + 2 * idx);
      if (col + 1 >= TD * DIM) v.y = 0.0;
      out[idx] = v;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------------
template <int DIM, int ND, int NQ>
static int launch_generic(mb2_ctx* c, const ElemArgs& a, cudaStream_t st) {
  using Cfg = FastGenCfg<DIM, ND, NQ>;
  const unsigned grid = (unsigned)(a.e1 - a.e0);
  if (c->mat_kind == MAT_NEOHOOKEAN) element_analytic_kernel<DIM, ND, NQ, MAT_NEOHOOKEAN><<<grid, Cfg::NT, 0, st>>>(a);
  else element_analytic_kernel<DIM, ND, NQ, MAT_STVK><<<grid, Cfg::NT, 0, st>>>(a);
  MB2_CUDA(cudaGetLastError());
  return MB2_OK;
}

template <int MAT, int GEMM>
static int launch_p2(mb2_ctx* c, const ElemArgs& a, cudaStream_t st) {
  auto kern = element_fast_p2_kernel<MAT, GEMM>;
  MB2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FastP2::SMEM_BYTES));
  kern<<<(unsigned)(a.e1 - a.e0), 256, FastP2::SMEM_BYTES, st>>>(a);
  MB2_CUDA(cudaGetLastError());
  return MB2_OK;
}

template <int MAT>
static int launch_p2v3(mb2_ctx* c, const ElemArgs& a, cudaStream_t st) {
  auto kern = element_fast_p2v3_kernel<MAT>;
  MB2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FastP2V3::SMEM_BYTES));
  kern<<<(unsigned)(a.e1 - a.e0), 256, FastP2V3::SMEM_BYTES, st>>>(a);
  MB2_CUDA(cudaGetLastError());
  return MB2_OK;
}

int launch_element_analytic(mb2_ctx* c, const ElemArgs& a, cudaStream_t st) {
  if (c->mat_kind != MAT_NEOHOOKEAN && c->mat_kind != MAT_STVK) {
    c->err = "MB2_TANGENT_ANALYTIC: closed-form tangents exist for CompressibleOgdenNeoHookean and StVenantKirchhoff only";
    return MB2_ERR_UNSUPPORTED;
  }
  const int dim = c->dim, nd = c->nd, nq = c->nq;
  if (dim == 3 && nd == 27 && nq == 64 && c->elem_variant != 1) {
    const bool nh = c->mat_kind == MAT_NEOHOOKEAN;
    // fast_gemm: 0 = tensor cores, A fragments built in registers (default); 1 = CUDA-core DFMA tiles; 2 = tensor cores
    // fed from T buffers in shared memory (the second form: same summation order as 1, hence the same bits)
    if (c->fast_gemm == 1) return nh ? launch_p2<MAT_NEOHOOKEAN, 1>(c, a, st) : launch_p2<MAT_STVK, 1>(c, a, st);
    if (c->fast_gemm == 2) return nh ? launch_p2<MAT_NEOHOOKEAN, 0>(c, a, st) : launch_p2<MAT_STVK, 0>(c, a, st);
    return nh ? launch_p2v3<MAT_NEOHOOKEAN>(c, a, st) : launch_p2v3<MAT_STVK>(c, a, st);
  }
  if (dim == 3 && nd == 27 && nq == 64) return launch_generic<3, 27, 64>(c, a, st);
  if (dim == 3 && nd == 64 && nq == 125) return launch_generic<3, 64, 125>(c, a, st);
  if (dim == 3 && nd == 8 && nq == 27) return launch_generic<3, 8, 27>(c, a, st);
  if (dim == 2 && nd == 16 && nq == 25) return launch_generic<2, 16, 25>(c, a, st);
  if (dim == 2 && nd == 9 && nq == 16) return launch_generic<2, 9, 16>(c, a, st);
  if (dim == 2 && nd == 4 && nq == 9) return launch_generic<2, 4, 9>(c, a, st);
  c->err = "MB2_TANGENT_ANALYTIC: no kernel instantiated for this element type";
  return MB2_ERR_UNSUPPORTED;
}

// FP64 tensor-core micro-benchmark (roofline denominator of the DMMA form): dependent-chain-free m8n8k4 issues
__global__ void __launch_bounds__(256) dmma_peak_kernel(int iters, double a, double b, double* out) {
  double acc[8][2];
#pragma unroll
  for (int i = 0; i < 8; i++) acc[i][0] = acc[i][1] = (double)(threadIdx.x + i) * 1e-3;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) dmma_m8n8k4(acc[i][0], acc[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += acc[i][0] + acc[i][1];
  if (s == 123.456) out[0] = s;
}

int launch_dmma_peak(mb2_ctx* c, double* tflops) {
  double* d_out = nullptr;
  MB2_CUDA(cudaMalloc(&d_out, 8));
  const int iters = 2048, threads = 256, blocks = c->sm_count * 8;
  float best = 1e30f;
  for (int rep = 0; rep < 4; rep++) {
    MB2_CUDA(cudaEventRecord(c->ev[0], c->stream));
    dmma_peak_kernel<<<blocks, threads, 0, c->stream>>>(iters, 1e-3, 1e-3, d_out);
    MB2_CUDA(cudaEventRecord(c->ev[1], c->stream));
    MB2_CUDA(cudaEventSynchronize(c->ev[1]));
    float ms = 0;
    MB2_CUDA(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
    if (rep > 0 && ms < best) best = ms;
  }
  MB2_CUDA(cudaFree(d_out));
  // one m8n8k4 = 8 x 8 x 4 FMA = 512 flop per warp
  const double flops = 512.0 * 8.0 * iters * (double)(threads / 32) * blocks;
  *tflops = flops / (best * 1e-3) / 1e12;
  return MB2_OK;
}

}  // namespace mb2
