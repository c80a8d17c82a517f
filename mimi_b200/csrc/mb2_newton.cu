// the O(nnz) / O(NV) sweeps that surround the integrators in one Newton iteration, on
// the device-resident CSR, so that an implicit step needs no 9 GB host round trip per iteration at C3.
//
// Reference sequence replaced: mimi::operators::NonlinearSolid::ResidualAndGrad
// (src/mimi/operators/nonlinear_solid.cpp:240-283) with mimi::forms::Nonlinear::AddMultGrad
// (src/mimi/forms/nonlinear.hpp:53-116):
//   temp_x = x + fac0 a                       mfem::add                      (linalg/vector.cpp:338-366)
//   y = M a  (+ S (v + fac1 a))               SparseMatrix::Mult / AddMult   (linalg/sparsemat.cpp:916-926)
//   J = M                                     std::copy_n of the mass values
//   domain integrators: y += R, J += fac0 K   -> mb2_assemble_residual_and_grad_device
//   y[ess] = 0; J.EliminateRowCol(ess)        (forms/nonlinear.hpp:111-115; sparsemat.cpp:2140-2180, DIAG_ONE)
//   contact integrators, then again y[ess] = 0 / EliminateRowCol
//   J += fac1 S                               SparseMatrix::Add              (sparsemat.cpp:3234-3256)
//   y -= rhs; y[ess] = 0                      Vector::Add(-1.0, rhs)
// Every kernel keeps the reference's per-entry expression (row sums sequential in column order), so y and J are
// bit-identical to the reference's (tests/test_gpu_newton.py against dumps of the compiled reference).
#include <dlfcn.h>
#include <mutex>
#include <string>
#include <algorithm>
#include <cmath>
#include <vector>

#include "mb2_ctx.h"

namespace mb2 {

struct NewtonState {
  double* d_M = nullptr;     // nnz mass values
  double* d_S = nullptr;     // nnz viscosity values (optional)
  double* d_rhs = nullptr;   // NV (optional)
  double* d_a = nullptr;     // NV acceleration iterate
  double* d_tv = nullptr;    // NV v + fac1 a
  double* d_v = nullptr;     // NV
  double* d_x = nullptr;     // NV
  int32_t* d_ess = nullptr;  // essential tdofs
  int64_t n_ess = 0;
  // linear solve: GMRES(m) + Jacobi on the resident Jacobian
  int64_t* d_diag = nullptr;  // position of the diagonal entry of every row
  double* d_c = nullptr;      // solution of the last solve
  double* d_w = nullptr;      // NV work
  double* d_t = nullptr;      // NV work
  double* d_part = nullptr;   // block partials of the dot products
  std::vector<double*> d_vk;  // Krylov basis, m + 1 vectors
  // direct option: cuSOLVER's sparse QR (mb2_newton_solve_direct_device)
  int32_t* d_I32 = nullptr;   // row offsets as int32 (cuSOLVER's index type)
  void* sp_handle = nullptr;  // cusolverSpHandle_t
  void* sp_descr = nullptr;   // cusparseMatDescr_t
};

static NewtonState* NS(mb2_ctx* c) { return static_cast<NewtonState*>(c->newton); }

static void free_direct(NewtonState* S);  // cuSOLVER handles of the direct option (defined with it below)

void free_newton(mb2_ctx* c) {
  NewtonState* S = NS(c);
  if (!S) return;
  free_direct(S);
  for (void* p : {(void*)S->d_M, (void*)S->d_S, (void*)S->d_rhs, (void*)S->d_a, (void*)S->d_tv, (void*)S->d_v, (void*)S->d_x,
                  (void*)S->d_ess, (void*)S->d_diag, (void*)S->d_c, (void*)S->d_w, (void*)S->d_t, (void*)S->d_part})
    if (p) cudaFree(p);
  for (double* p : S->d_vk)
    if (p) cudaFree(p);
  delete S;
  c->newton = nullptr;
}

// z = x + alpha y   (mfem::add)
__global__ void k_add_scaled(int64_t n, const double* __restrict__ x, double alpha, const double* __restrict__ y,
                             double* __restrict__ z) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) z[i] = x[i] + alpha * y[i];
}

// y[i] (+)= 1.0 * sum_j A[j] x[J[j]], j ascending from d = 0.0. One warp per row would change the summation order:
// one THREAD per row keeps it (rows hold <= 375 entries at p = 2; the sweep is a small fraction of an assembly).
__global__ void k_spmv_rows(int64_t nrows, const int64_t* __restrict__ I, const int32_t* __restrict__ J,
                            const double* __restrict__ A, const double* __restrict__ x, double* __restrict__ y, int add) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= nrows) return;
  double d = 0.0;
  for (int64_t j = I[i]; j < I[i + 1]; j++) d += A[j] * x[J[j]];
  y[i] = (add ? y[i] : 0.0) + 1.0 * d;
}

__global__ void k_axpy(int64_t n, double a, const double* __restrict__ x, double* __restrict__ y) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) y[i] += a * x[i];
}

// y[ess] = 0 and SparseMatrix::EliminateRowCol(rc, DIAG_ONE): row rc -> e_rc, column rc zeroed through the symmetric
// pattern (columns are sorted: binary search instead of the reference's linear scan; the written values are the same).
__global__ void k_eliminate(int64_t n_ess, const int32_t* __restrict__ ess, const int64_t* __restrict__ I,
                            const int32_t* __restrict__ J, double* __restrict__ A, double* __restrict__ y, int* flag) {
  const int64_t e = blockIdx.x;
  if (e >= n_ess) return;
  const int32_t rc = ess[e];
  if (threadIdx.x == 0) y[rc] = 0.0;
  for (int64_t j = I[rc] + threadIdx.x; j < I[rc + 1]; j += blockDim.x) {
    const int32_t col = J[j];
    if (col == rc) {
      A[j] = 1.0;
    } else {
      A[j] = 0.0;
      int64_t lo = I[col], hi = I[col + 1] - 1;
      bool found = false;
      while (lo <= hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (J[mid] == rc) {
          A[mid] = 0.0;
          found = true;
          break;
        }
        if (J[mid] < rc) lo = mid + 1; else hi = mid - 1;
      }
      if (!found && I[col + 1] > I[col]) *flag = 3;  // non-symmetric pattern ("EliminateRowCol() #2")
    }
  }
}

__global__ void k_zero_list(int64_t n, const int32_t* __restrict__ idx, double* __restrict__ y) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) y[idx[i]] = 0.0;
}

static inline unsigned nb(int64_t n) { return (unsigned)((n + 255) / 256); }

int contact_assemble(mb2_ctx* c, bool grad, double gf);  // mb2_contact.cu
bool contact_active(mb2_ctx* c);

int newton_residual_and_grad(mb2_ctx* c, const double* x, const double* v, const double* a, double fac0, double fac1) {
  NewtonState* S = NS(c);
  const int64_t NV = c->NV, nnz = c->nnz;
  cudaStream_t st = c->stream;
  MB2_CUDA(cudaMemcpyAsync(S->d_x, x, sizeof(double) * NV, cudaMemcpyHostToDevice, st));
  MB2_CUDA(cudaMemcpyAsync(S->d_a, a, sizeof(double) * NV, cudaMemcpyHostToDevice, st));
  k_add_scaled<<<nb(NV), 256, 0, st>>>(NV, S->d_x, fac0, S->d_a, c->d_u);               // temp_x
  k_spmv_rows<<<nb(NV), 256, 0, st>>>(NV, c->d_I, c->d_J, S->d_M, S->d_a, c->d_r, 0);    // y = M a
  if (S->d_S) {
    if (!v) { c->err = "viscosity is set: v must be given"; return MB2_ERR_ARG; }
    MB2_CUDA(cudaMemcpyAsync(S->d_v, v, sizeof(double) * NV, cudaMemcpyHostToDevice, st));
    k_add_scaled<<<nb(NV), 256, 0, st>>>(NV, S->d_v, fac1, S->d_a, S->d_tv);
    k_spmv_rows<<<nb(NV), 256, 0, st>>>(NV, c->d_I, c->d_J, S->d_S, S->d_tv, c->d_r, 1);
  }
  MB2_CUDA(cudaMemcpyAsync(c->d_A, S->d_M, sizeof(double) * nnz, cudaMemcpyDeviceToDevice, st));  // J = M
  MB2_CUDA(cudaGetLastError());
  int rc = launch_residual_and_grad(c, fac0, nullptr, nullptr);
  if (rc) return rc;
  int launches = c->last_launches + 3;
  if (S->n_ess) {
    k_eliminate<<<(unsigned)S->n_ess, 128, 0, st>>>(S->n_ess, S->d_ess, c->d_I, c->d_J, c->d_A, c->d_r, c->d_flag);
    launches++;
  }
  if (contact_active(c)) {
    if ((rc = contact_assemble(c, true, fac0))) return rc;
    launches += c->last_launches;
    if (S->n_ess) {
      k_eliminate<<<(unsigned)S->n_ess, 128, 0, st>>>(S->n_ess, S->d_ess, c->d_I, c->d_J, c->d_A, c->d_r, c->d_flag);
      launches++;
    }
  }
  if (S->d_S) k_axpy<<<nb(nnz), 256, 0, st>>>(nnz, fac1, S->d_S, c->d_A);                  // J += fac1 S
  if (S->d_rhs) k_axpy<<<nb(NV), 256, 0, st>>>(NV, -1.0, S->d_rhs, c->d_r);                // y -= rhs
  if (S->n_ess) k_zero_list<<<nb(S->n_ess), 256, 0, st>>>(S->n_ess, S->d_ess, c->d_r);     // y[dirichlet] = 0
  MB2_CUDA(cudaGetLastError());
  c->last_launches = launches + 3;
  return MB2_OK;
}


// ---- the reference's iterative linear solver on the resident CSR --------------------------------
// mfem::GMRESSolver left-preconditioned by mfem::DSmoother (PyNonlinearSolid::Setup with use_iterative_solver,
// src/mimi/py/py_nonlinear_solid.cpp:330-339; third_party/mfem/linalg/solvers.cpp:931-1160, sparsemat.cpp:2683-2720):
// modified Gram-Schmidt, Givens rotations, restart m, iterative_mode == false. SpMV keeps the reference's row sums; the
// dot products are fixed-shape two-level tree reductions (deterministic, but not the reference's sequential sum), so the
// iterates agree with the reference to rounding, not bitwise: tolerance parity (tests/test_gpu_newton.py).
__global__ void k_find_diag(int64_t n, const int64_t* __restrict__ I, const int32_t* __restrict__ J, int64_t* __restrict__ diag,
                            int* flag) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int64_t lo = I[i], hi = I[i + 1] - 1, pos = -1;
  while (lo <= hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (J[mid] == i) { pos = mid; break; }
    if (J[mid] < i) lo = mid + 1; else hi = mid - 1;
  }
  diag[i] = pos;
  if (pos < 0) *flag = 4;  // "Diagonal not found in SparseMatrix::DiagScale"
}
// x[i] = 1.0 * b[i] / A[diag(i)]   (SparseMatrix::DiagScale, sc = 1)
__global__ void k_diag_scale(int64_t n, const int64_t* __restrict__ diag, const double* __restrict__ A,
                             const double* __restrict__ b, double* __restrict__ x) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) x[i] = 1.0 * b[i] / A[diag[i]];
}
__global__ void k_scale_to(int64_t n, double a, const double* __restrict__ x, double* __restrict__ y) {  // y = a x
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) y[i] = a * x[i];
}
__global__ void k_sub(int64_t n, const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ z) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) z[i] = a[i] - b[i];
}
constexpr int DOT_CHUNK = 4096;  // elements per block: thread t sums t, t + 256, ... in order, then a shared-memory tree
__global__ void __launch_bounds__(256) k_dot_partial(int64_t n, const double* __restrict__ a, const double* __restrict__ b,
                                                     double* __restrict__ part) {
  __shared__ double sh[256];
  const int64_t base = (int64_t)blockIdx.x * DOT_CHUNK;
  double d = 0.0;
  for (int k = threadIdx.x; k < DOT_CHUNK; k += 256) {
    const int64_t i = base + k;
    if (i < n) d += a[i] * b[i];
  }
  sh[threadIdx.x] = d;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) part[blockIdx.x] = sh[0];
}
__global__ void __launch_bounds__(256) k_dot_final(int64_t nparts, const double* __restrict__ part, double* __restrict__ out) {
  __shared__ double sh[256];
  double d = 0.0;
  for (int64_t k = threadIdx.x; k < nparts; k += 256) d += part[k];
  sh[threadIdx.x] = d;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = sh[0];
}

static int dev_dot(mb2_ctx* c, NewtonState* S, const double* a, const double* b, double* out) {
  const int64_t n = c->NV, nparts = (n + DOT_CHUNK - 1) / DOT_CHUNK;
  k_dot_partial<<<(unsigned)nparts, 256, 0, c->stream>>>(n, a, b, S->d_part);
  k_dot_final<<<1, 256, 0, c->stream>>>(nparts, S->d_part, S->d_part + nparts);
  MB2_CUDA(cudaMemcpyAsync(out, S->d_part + nparts, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}

int newton_solve_gmres(mb2_ctx* c, double rel_tol, double abs_tol, int max_iter, int m, int* final_iter, double* final_norm,
                       int* converged_out) {
  NewtonState* S = NS(c);
  const int64_t n = c->NV;
  cudaStream_t st = c->stream;
  const int64_t nparts = (n + DOT_CHUNK - 1) / DOT_CHUNK;
  if (!S->d_diag) {
    MB2_CUDA(cudaMalloc(&S->d_diag, sizeof(int64_t) * n));
    k_find_diag<<<nb(n), 256, 0, st>>>(n, c->d_I, c->d_J, S->d_diag, c->d_flag);
    for (double** p : {&S->d_c, &S->d_w, &S->d_t})
      if (!*p) MB2_CUDA(cudaMalloc(p, sizeof(double) * n));   // d_c may exist already (direct option)
  }
  if (!S->d_part) MB2_CUDA(cudaMalloc(&S->d_part, sizeof(double) * (nparts + 1)));
  if ((int)S->d_vk.size() < m + 1) S->d_vk.resize(m + 1, nullptr);
  const int ldh = m + 1;
  std::vector<double> H((size_t)ldh * m, 0.0), s(m + 1, 0.0), cs(m + 1, 0.0), sn(m + 1, 0.0), y(m + 1, 0.0);
  double* r = S->d_t;   // preconditioned residual
  double* w = S->d_w;
  double* x = S->d_c;
  const double* b = c->d_r;
  const double* A = c->d_A;
  auto update = [&](int k) {  // Update(x, k, H, s, v): back substitution on the host, x += sum_j y_j v_j
    for (int i = 0; i <= m; i++) y[i] = s[i];
    for (int i = k; i >= 0; i--) {
      y[i] /= H[i + (size_t)i * ldh];
      for (int j = i - 1; j >= 0; j--) y[j] -= H[j + (size_t)i * ldh] * y[i];
    }
    for (int j = 0; j <= k; j++) k_axpy<<<nb(n), 256, 0, st>>>(n, y[j], S->d_vk[j], x);
  };
  int rc, i = 0, j = 0, converged = 0;
  MB2_CUDA(cudaMemsetAsync(x, 0, sizeof(double) * n, st));
  k_diag_scale<<<nb(n), 256, 0, st>>>(n, S->d_diag, A, b, r);
  double dd = 0.0;
  if ((rc = dev_dot(c, S, r, r, &dd))) return rc;
  double beta = std::sqrt(dd);
  if (!std::isfinite(beta)) { c->err = "GMRES: beta is not finite"; return MB2_ERR_NUMERIC; }
  *final_norm = std::max(rel_tol * beta, abs_tol);
  if (beta <= *final_norm) {
    *final_norm = beta;
    *final_iter = 0;
    converged = 1;
    goto finish;
  }
  for (j = 1; j <= max_iter;) {
    if (!S->d_vk[0]) MB2_CUDA(cudaMalloc(&S->d_vk[0], sizeof(double) * n));
    k_scale_to<<<nb(n), 256, 0, st>>>(n, 1.0 / beta, r, S->d_vk[0]);
    std::fill(s.begin(), s.end(), 0.0);
    s[0] = beta;
    for (i = 0; i < m && j <= max_iter; i++, j++) {
      k_spmv_rows<<<nb(n), 256, 0, st>>>(n, c->d_I, c->d_J, A, S->d_vk[i], r, 0);
      k_diag_scale<<<nb(n), 256, 0, st>>>(n, S->d_diag, A, r, w);
      for (int k = 0; k <= i; k++) {
        double h = 0.0;
        if ((rc = dev_dot(c, S, w, S->d_vk[k], &h))) return rc;
        H[k + (size_t)i * ldh] = h;
        k_axpy<<<nb(n), 256, 0, st>>>(n, -h, S->d_vk[k], w);
      }
      if ((rc = dev_dot(c, S, w, w, &dd))) return rc;
      const double hn = std::sqrt(dd);
      if (!std::isfinite(hn)) { c->err = "GMRES: Norm(w) is not finite"; return MB2_ERR_NUMERIC; }
      H[i + 1 + (size_t)i * ldh] = hn;
      if (!S->d_vk[i + 1]) MB2_CUDA(cudaMalloc(&S->d_vk[i + 1], sizeof(double) * n));
      k_scale_to<<<nb(n), 256, 0, st>>>(n, 1.0 / hn, w, S->d_vk[i + 1]);
      for (int k = 0; k < i; k++) {  // ApplyPlaneRotation
        double& dx = H[k + (size_t)i * ldh];
        double& dy = H[k + 1 + (size_t)i * ldh];
        const double t = cs[k] * dx + sn[k] * dy;
        dy = -sn[k] * dx + cs[k] * dy;
        dx = t;
      }
      {  // GeneratePlaneRotation
        const double dx = H[i + (size_t)i * ldh], dy = H[i + 1 + (size_t)i * ldh];
        if (dy == 0.0) {
          cs[i] = 1.0;
          sn[i] = 0.0;
        } else if (std::fabs(dy) > std::fabs(dx)) {
          const double t = dx / dy;
          sn[i] = 1.0 / std::sqrt(1.0 + t * t);
          cs[i] = t * sn[i];
        } else {
          const double t = dy / dx;
          cs[i] = 1.0 / std::sqrt(1.0 + t * t);
          sn[i] = t * cs[i];
        }
      }
      {
        double& dx = H[i + (size_t)i * ldh];
        double& dy = H[i + 1 + (size_t)i * ldh];
        double t = cs[i] * dx + sn[i] * dy;
        dy = -sn[i] * dx + cs[i] * dy;
        dx = t;
        t = cs[i] * s[i] + sn[i] * s[i + 1];
        s[i + 1] = -sn[i] * s[i] + cs[i] * s[i + 1];
        s[i] = t;
      }
      const double resid = std::fabs(s[i + 1]);
      if (!std::isfinite(resid)) { c->err = "GMRES: resid is not finite"; return MB2_ERR_NUMERIC; }
      if (resid <= *final_norm) {
        update(i);
        *final_norm = resid;
        *final_iter = j;
        converged = 1;
        goto finish;
      }
    }
    update(i - 1);
    k_spmv_rows<<<nb(n), 256, 0, st>>>(n, c->d_I, c->d_J, A, x, r, 0);
    k_sub<<<nb(n), 256, 0, st>>>(n, b, r, w);
    k_diag_scale<<<nb(n), 256, 0, st>>>(n, S->d_diag, A, w, r);
    if ((rc = dev_dot(c, S, r, r, &dd))) return rc;
    beta = std::sqrt(dd);
    if (beta <= *final_norm) {
      *final_norm = beta;
      *final_iter = j;
      converged = 1;
      goto finish;
    }
  }
  *final_norm = beta;
  *final_iter = max_iter;
finish:
  MB2_CUDA(cudaGetLastError());
  MB2_CUDA(cudaStreamSynchronize(st));
  *converged_out = converged;
  return MB2_OK;
}

// NonlinearSolid::Mult(a, y) (src/mimi/operators/nonlinear_solid.cpp:172-205): the residual-only evaluation the line
// search of LineSearchNewton makes twice per iteration (solvers/newton.cpp:138-157), and its norm:
//   temp_x = x + fac0 a;  y = M a;  stiffness->AddMult(temp_x, y) (+ y[ess] = 0, forms/nonlinear.hpp:53-82);
//   y += S (v + fac1 a);  contact->AddMult(temp_x, y) (+ y[ess] = 0);  y -= rhs;  y[dirichlet] = 0;  ||y||_2
// y bit-identical to the reference; the norm is a deterministic tree reduction (the reference: sequential sum).
int newton_residual(mb2_ctx* c, const double* x, const double* v, const double* a, double fac0, double fac1,
                    double* norm_out) {
  NewtonState* S = NS(c);
  const int64_t NV = c->NV;
  cudaStream_t st = c->stream;
  MB2_CUDA(cudaMemcpyAsync(S->d_x, x, sizeof(double) * NV, cudaMemcpyHostToDevice, st));
  MB2_CUDA(cudaMemcpyAsync(S->d_a, a, sizeof(double) * NV, cudaMemcpyHostToDevice, st));
  k_add_scaled<<<nb(NV), 256, 0, st>>>(NV, S->d_x, fac0, S->d_a, c->d_u);
  k_spmv_rows<<<nb(NV), 256, 0, st>>>(NV, c->d_I, c->d_J, S->d_M, S->d_a, c->d_r, 0);
  MB2_CUDA(cudaGetLastError());
  int rc = launch_residual(c);
  if (rc) return rc;
  int launches = c->last_launches + 2;
  if (S->n_ess) k_zero_list<<<nb(S->n_ess), 256, 0, st>>>(S->n_ess, S->d_ess, c->d_r);
  if (S->d_S) {
    if (!v) { c->err = "viscosity is set: v must be given"; return MB2_ERR_ARG; }
    MB2_CUDA(cudaMemcpyAsync(S->d_v, v, sizeof(double) * NV, cudaMemcpyHostToDevice, st));
    k_add_scaled<<<nb(NV), 256, 0, st>>>(NV, S->d_v, fac1, S->d_a, S->d_tv);
    k_spmv_rows<<<nb(NV), 256, 0, st>>>(NV, c->d_I, c->d_J, S->d_S, S->d_tv, c->d_r, 1);
    launches += 2;
  }
  if (contact_active(c)) {
    if ((rc = contact_assemble(c, false, 0.0))) return rc;
    launches += c->last_launches;
    if (S->n_ess) k_zero_list<<<nb(S->n_ess), 256, 0, st>>>(S->n_ess, S->d_ess, c->d_r);
  }
  if (S->d_rhs) k_axpy<<<nb(NV), 256, 0, st>>>(NV, -1.0, S->d_rhs, c->d_r);
  if (S->n_ess) k_zero_list<<<nb(S->n_ess), 256, 0, st>>>(S->n_ess, S->d_ess, c->d_r);
  MB2_CUDA(cudaGetLastError());
  c->last_launches = launches + 3;
  if (norm_out) {
    if (!S->d_part) MB2_CUDA(cudaMalloc(&S->d_part, sizeof(double) * ((NV + DOT_CHUNK - 1) / DOT_CHUNK + 1)));
    double dd = 0.0;
    if ((rc = dev_dot(c, S, c->d_r, c->d_r, &dd))) return rc;
    *norm_out = std::sqrt(dd);
    if ((rc = consume_flag(c))) return rc;  // the norm already synchronised: a failed return mapping surfaces here
  }
  return MB2_OK;
}

// ---- direct option of the linear solve: the reference's default is mfem::UMFPackSolver (py_nonlinear_solid.cpp:340-343).
// A sparse direct factorisation is library territory, not a kernel of this path: the resident CSR is handed to cuSOLVER's
// sparse QR (cusolverSpDcsrlsvqr, device path, symrcm reordering), loaded with dlopen like NCCL so that the library keeps no
// link-time dependency. For the mesh sizes a direct solve is used at; large meshes take the iterative option above.
struct SpApi {
  void *lib_solver = nullptr, *lib_sparse = nullptr;
  int (*Create)(void**) = nullptr;
  int (*Destroy)(void*) = nullptr;
  int (*SetStream)(void*, cudaStream_t) = nullptr;
  int (*CreateMatDescr)(void**) = nullptr;
  int (*DestroyMatDescr)(void*) = nullptr;
  int (*csrlsvqr)(void*, int, int, void*, const double*, const int*, const int*, const double*, double, int, double*, int*) = nullptr;
};
static SpApi g_sp;
static int load_sp(std::string& err) {
  if (g_sp.csrlsvqr) return MB2_OK;
  for (const char* n : {"libcusolver.so.11", "libcusolver.so.12", "libcusolver.so"})
    if ((g_sp.lib_solver = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
  for (const char* n : {"libcusparse.so.12", "libcusparse.so.11", "libcusparse.so"})
    if ((g_sp.lib_sparse = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
  if (!g_sp.lib_solver || !g_sp.lib_sparse) { err = "direct solve: cannot load libcusolver / libcusparse"; return MB2_ERR_UNSUPPORTED; }
#define MB2_SP(field, lib, name)                                                          \
  g_sp.field = reinterpret_cast<decltype(g_sp.field)>(dlsym(g_sp.lib, name));            \
  if (!g_sp.field) { err = std::string("direct solve: missing symbol ") + name; return MB2_ERR_UNSUPPORTED; }
  MB2_SP(Create, lib_solver, "cusolverSpCreate")
  MB2_SP(Destroy, lib_solver, "cusolverSpDestroy")
  MB2_SP(SetStream, lib_solver, "cusolverSpSetStream")
  MB2_SP(CreateMatDescr, lib_sparse, "cusparseCreateMatDescr")
  MB2_SP(DestroyMatDescr, lib_sparse, "cusparseDestroyMatDescr")
  MB2_SP(csrlsvqr, lib_solver, "cusolverSpDcsrlsvqr")
#undef MB2_SP
  return MB2_OK;
}
static void free_direct(NewtonState* S) {
  if (S->d_I32) cudaFree(S->d_I32);
  if (S->sp_descr && g_sp.DestroyMatDescr) g_sp.DestroyMatDescr(S->sp_descr);
  if (S->sp_handle && g_sp.Destroy) g_sp.Destroy(S->sp_handle);
  S->d_I32 = nullptr;
  S->sp_descr = S->sp_handle = nullptr;
}
__global__ void k_i64_to_i32(int64_t n, const int64_t* __restrict__ a, int32_t* __restrict__ b) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) b[i] = (int32_t)a[i];
}
int newton_solve_direct(mb2_ctx* c, int* singular_row) {
  NewtonState* S = NS(c);
  const int64_t n = c->NV;
  if (c->nnz > 2147483647LL || n > 2147483647LL) { c->err = "direct solve: cuSOLVER indexes with 32 bits (use the iterative option)"; return MB2_ERR_UNSUPPORTED; }
  int rc = load_sp(c->err);
  if (rc) return rc;
  if (!S->sp_handle) {
    if (g_sp.Create(&S->sp_handle) || g_sp.CreateMatDescr(&S->sp_descr)) { c->err = "direct solve: cusolverSpCreate / cusparseCreateMatDescr failed"; return MB2_ERR_CUDA; }
  }
  if (g_sp.SetStream(S->sp_handle, c->stream)) { c->err = "direct solve: cusolverSpSetStream failed"; return MB2_ERR_CUDA; }
  if (!S->d_I32) {
    MB2_CUDA(cudaMalloc(&S->d_I32, sizeof(int32_t) * (n + 1)));
    k_i64_to_i32<<<nb(n + 1), 256, 0, c->stream>>>(n + 1, c->d_I, S->d_I32);
    MB2_CUDA(cudaGetLastError());
  }
  if (!S->d_c) MB2_CUDA(cudaMalloc(&S->d_c, sizeof(double) * n));
  int sing = -1;
  const int st = g_sp.csrlsvqr(S->sp_handle, (int)n, (int)c->nnz, S->sp_descr, c->d_A, S->d_I32, c->d_J, c->d_r, 1e-300, 1, S->d_c, &sing);
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  if (st) { c->err = "direct solve: cusolverSpDcsrlsvqr returned status " + std::to_string(st); return MB2_ERR_CUDA; }
  if (singular_row) *singular_row = sing;
  if (sing >= 0) { c->err = "direct solve: the Jacobian is singular (row " + std::to_string(sing) + ")"; return MB2_ERR_NUMERIC; }
  return MB2_OK;
}

static int ensure(mb2_ctx* c) {
  if (!c->have_pattern) { c->err = "call mb2_build_pattern first"; return MB2_ERR_ARG; }
  if (!c->newton) {
    NewtonState* S = new NewtonState();
    c->newton = S;
    for (double** p : {&S->d_a, &S->d_tv, &S->d_v, &S->d_x}) MB2_CUDA(cudaMalloc(p, sizeof(double) * c->NV));
  }
  return MB2_OK;
}

}  // namespace mb2

using mb2::NewtonState;

extern "C" {

int mb2_set_mass(mb2_ctx* c, const double* M) {
  if (!c) return MB2_ERR_ARG;
  int rc = mb2::ensure(c);
  if (rc) return rc;
  if (!M) { c->err = "null mass values"; return MB2_ERR_ARG; }
  NewtonState* S = static_cast<NewtonState*>(c->newton);
  if (!S->d_M) MB2_CUDA(cudaMalloc(&S->d_M, sizeof(double) * c->nnz));
  MB2_CUDA(cudaMemcpyAsync(S->d_M, M, sizeof(double) * c->nnz, cudaMemcpyHostToDevice, c->stream));
  return MB2_OK;
}

static int adopt_A(mb2_ctx* c, double** dst, int form_system_matrix) {
  NewtonState* S = static_cast<NewtonState*>(c->newton);
  if (!c->d_A) { c->err = "no resident CSR values"; return MB2_ERR_ARG; }
  if (!*dst) MB2_CUDA(cudaMalloc(dst, sizeof(double) * c->nnz));
  MB2_CUDA(cudaMemcpyAsync(*dst, c->d_A, sizeof(double) * c->nnz, cudaMemcpyDeviceToDevice, c->stream));
  if (form_system_matrix && S->n_ess) {
    // BilinearForm::FormSystemMatrix(zero_dofs, .) (py_nonlinear_solid.cpp:173,190): EliminateRowCol(dof, DIAG_ONE) per dof
    MB2_CUDA(cudaMemsetAsync(c->d_flag, 0, sizeof(int), c->stream));
    mb2::k_eliminate<<<(unsigned)S->n_ess, 64, 0, c->stream>>>(S->n_ess, S->d_ess, c->d_I, c->d_J, *dst, S->d_tv, c->d_flag);
    MB2_CUDA(cudaGetLastError());
    int flag = 0;
    MB2_CUDA(cudaMemcpyAsync(&flag, c->d_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    MB2_CUDA(cudaStreamSynchronize(c->stream));
    if (flag) { c->err = "non-symmetric pattern while eliminating essential dofs"; return MB2_ERR_NUMERIC; }
  }
  return MB2_OK;
}

int mb2_newton_mass_from_A(mb2_ctx* c, int form_system_matrix) {
  if (!c) return MB2_ERR_ARG;
  int rc = mb2::ensure(c);
  if (rc) return rc;
  return adopt_A(c, &static_cast<NewtonState*>(c->newton)->d_M, form_system_matrix);
}

int mb2_newton_viscosity_from_A(mb2_ctx* c, int form_system_matrix) {
  if (!c) return MB2_ERR_ARG;
  int rc = mb2::ensure(c);
  if (rc) return rc;
  return adopt_A(c, &static_cast<NewtonState*>(c->newton)->d_S, form_system_matrix);
}

int mb2_newton_get_viscosity(mb2_ctx* c, double* Sv) {
  if (!c) return MB2_ERR_ARG;
  NewtonState* S = static_cast<NewtonState*>(c->newton);
  if (!S || !S->d_S || !Sv) { c->err = "no viscosity on the device"; return MB2_ERR_ARG; }
  MB2_CUDA(cudaMemcpyAsync(Sv, S->d_S, sizeof(double) * c->nnz, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}

int mb2_newton_rhs_from_r(mb2_ctx* c, int zero_essential) {
  if (!c) return MB2_ERR_ARG;
  int rc = mb2::ensure(c);
  if (rc) return rc;
  NewtonState* S = static_cast<NewtonState*>(c->newton);
  if (!c->d_r) { c->err = "no resident vector"; return MB2_ERR_ARG; }
  if (!S->d_rhs) MB2_CUDA(cudaMalloc(&S->d_rhs, sizeof(double) * c->NV));
  MB2_CUDA(cudaMemcpyAsync(S->d_rhs, c->d_r, sizeof(double) * c->NV, cudaMemcpyDeviceToDevice, c->stream));
  if (zero_essential && S->n_ess) {
    mb2::k_zero_list<<<mb2::nb(S->n_ess), 256, 0, c->stream>>>(S->n_ess, S->d_ess, S->d_rhs);
    MB2_CUDA(cudaGetLastError());
  }
  return MB2_OK;
}

int mb2_newton_get_mass(mb2_ctx* c, double* M) {
  if (!c) return MB2_ERR_ARG;
  NewtonState* S = static_cast<NewtonState*>(c->newton);
  if (!S || !S->d_M || !M) { c->err = "no mass on the device"; return MB2_ERR_ARG; }
  MB2_CUDA(cudaMemcpyAsync(M, S->d_M, sizeof(double) * c->nnz, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}

int mb2_set_viscosity(mb2_ctx* c, const double* Sv) {
  if (!c) return MB2_ERR_ARG;
  int rc = mb2::ensure(c);
  if (rc) return rc;
  NewtonState* S = static_cast<NewtonState*>(c->newton);
  if (!Sv) {
    if (S->d_S) cudaFree(S->d_S);
    S->d_S = nullptr;
    return MB2_OK;
  }
  if (!S->d_S) MB2_CUDA(cudaMalloc(&S->d_S, sizeof(double) * c->nnz));
  MB2_CUDA(cudaMemcpyAsync(S->d_S, Sv, sizeof(double) * c->nnz, cudaMemcpyHostToDevice, c->stream));
  return MB2_OK;
}

int mb2_set_rhs(mb2_ctx* c, const double* rhs) {
  if (!c) return MB2_ERR_ARG;
  int rc = mb2::ensure(c);
  if (rc) return rc;
  NewtonState* S = static_cast<NewtonState*>(c->newton);
  if (!rhs) {
    if (S->d_rhs) cudaFree(S->d_rhs);
    S->d_rhs = nullptr;
    return MB2_OK;
  }
  if (!S->d_rhs) MB2_CUDA(cudaMalloc(&S->d_rhs, sizeof(double) * c->NV));
  MB2_CUDA(cudaMemcpyAsync(S->d_rhs, rhs, sizeof(double) * c->NV, cudaMemcpyHostToDevice, c->stream));
  return MB2_OK;
}

int mb2_set_essential_dofs(mb2_ctx* c, const int32_t* tdofs, int64_t n) {
  if (!c) return MB2_ERR_ARG;
  int rc = mb2::ensure(c);
  if (rc) return rc;
  NewtonState* S = static_cast<NewtonState*>(c->newton);
  for (int64_t i = 0; i < n; i++)
    if (tdofs[i] < 0 || tdofs[i] >= c->NV) { c->err = "essential dof out of range"; return MB2_ERR_ARG; }
  if (S->d_ess) cudaFree(S->d_ess);
  S->d_ess = nullptr;
  S->n_ess = n;
  if (n) {
    MB2_CUDA(cudaMalloc(&S->d_ess, sizeof(int32_t) * n));
    MB2_CUDA(cudaMemcpy(S->d_ess, tdofs, sizeof(int32_t) * n, cudaMemcpyHostToDevice));
  }
  return MB2_OK;
}

int mb2_newton_residual_and_grad_device(mb2_ctx* c, const double* x, const double* v, const double* a, double fac0,
                                        double fac1) {
  if (!c) return MB2_ERR_ARG;
  NewtonState* S = static_cast<NewtonState*>(c->newton);
  if (!S || !S->d_M) { c->err = "call mb2_set_mass first"; return MB2_ERR_ARG; }
  if (!x || !a) { c->err = "null vector"; return MB2_ERR_ARG; }
  if (!c->d_dNdX || c->mat_kind < 0) { c->err = "tables / material missing"; return MB2_ERR_ARG; }
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::newton_residual_and_grad(c, x, v, a, fac0, fac1);
}

int mb2_newton_solve_gmres_device(mb2_ctx* c, double rel_tol, double abs_tol, int max_iter, int kdim, int* iterations,
                                  double* final_norm, int* converged) {
  if (!c) return MB2_ERR_ARG;
  int rc = mb2::ensure(c);
  if (rc) return rc;
  if (!iterations || !final_norm || !converged) { c->err = "null output"; return MB2_ERR_ARG; }
  if (max_iter < 1 || kdim < 1 || kdim > 1000) { c->err = "GMRES: max_iter >= 1 and 1 <= kdim <= 1000"; return MB2_ERR_ARG; }
  MB2_CUDA(cudaSetDevice(c->device));
  rc = mb2::newton_solve_gmres(c, rel_tol, abs_tol, max_iter, kdim, iterations, final_norm, converged);
  if (rc) return rc;
  // any latched device failure (a failed return mapping in the assembly that produced J, a bad elimination, a missing
  // diagonal) makes this solve meaningless: the reference would have thrown before reaching its linear solver
  return mb2::consume_flag(c);
}

int mb2_newton_solve_direct_device(mb2_ctx* c, int* singular_row) {
  if (!c) return MB2_ERR_ARG;
  int rc = mb2::ensure(c);
  if (rc) return rc;
  MB2_CUDA(cudaSetDevice(c->device));
  if ((rc = mb2::newton_solve_direct(c, singular_row))) return rc;
  return mb2::consume_flag(c);
}

// Host-CSR consumer without the per-step upload of the CSR: what operators::NonlinearSolid::ResidualAndGrad does around
// the domain integrator (operators/nonlinear_solid.cpp:257: J.data <- mass.data, then stiffness->AddMultGrad), with the mass
// resident since mb2_set_mass / mb2_newton_mass_from_A. r += R(u) (r travels both ways, NV doubles); A_out = M + gf K(u)
// leaves chunk by chunk behind the element kernels (D2H only).
int mb2_domain_residual_and_grad_on_mass(mb2_ctx* c, const double* u, double gf, double* r, double* A_out) {
  if (!c) return MB2_ERR_ARG;
  NewtonState* S = static_cast<NewtonState*>(c->newton);
  if (!S || !S->d_M) { c->err = "call mb2_set_mass (or mb2_newton_mass_from_A) first"; return MB2_ERR_ARG; }
  if (!u || !r || !A_out) { c->err = "null vector"; return MB2_ERR_ARG; }
  if (!c->d_dNdX || c->mat_kind < 0) { c->err = "tables / material missing"; return MB2_ERR_ARG; }
  MB2_CUDA(cudaSetDevice(c->device));
  int rc;
  if ((rc = mb2_set_u(c, u))) return rc;
  if ((rc = mb2_set_r(c, r))) return rc;
  MB2_CUDA(cudaMemcpyAsync(c->d_A, S->d_M, sizeof(double) * c->nnz, cudaMemcpyDeviceToDevice, c->stream));
  if ((rc = mb2::launch_residual_and_grad(c, gf, nullptr, A_out))) return rc;
  if ((rc = mb2_get_r(c, r))) return rc;
  return c->ns ? mb2_check_numeric(c) : MB2_OK;
}

int mb2_newton_get_solution(mb2_ctx* c, double* out) {
  if (!c) return MB2_ERR_ARG;
  NewtonState* S = static_cast<NewtonState*>(c->newton);
  if (!S || !S->d_c || !out) { c->err = "no linear solve has run"; return MB2_ERR_ARG; }
  MB2_CUDA(cudaMemcpyAsync(out, S->d_c, sizeof(double) * c->NV, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}

int mb2_newton_residual_device(mb2_ctx* c, const double* x, const double* v, const double* a, double fac0, double fac1,
                               double* norm_out) {
  if (!c) return MB2_ERR_ARG;
  NewtonState* S = static_cast<NewtonState*>(c->newton);
  if (!S || !S->d_M) { c->err = "call mb2_set_mass first"; return MB2_ERR_ARG; }
  if (!x || !a) { c->err = "null vector"; return MB2_ERR_ARG; }
  if (!c->d_dNdX || c->mat_kind < 0) { c->err = "tables / material missing"; return MB2_ERR_ARG; }
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::newton_residual(c, x, v, a, fac0, fac1, norm_out);
}

// ---------------------------------------------------------------------------------------------
// Allocation-free dot product (no counterpart in the reference; added for callers
// that cannot call cudaMalloc: CUDA graph capture, tight serving loops). The same two kernels as dev_dot above -- the
// deterministic reduction behind the Newton norm and the GMRES inner products -- with the block partials in a file-scope
// device array instead of NewtonState::d_part. Every GPU that loads the library holds the 128 KiB until exit.
// In host code `g_noalloc_part` is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress,
// resolved once per device with that device current (first call on the device), and is what the kernels receive.
// All calls share the array and it keeps the last call's partials: nothing needs clearing (k_dot_partial overwrites the
// nparts entries k_dot_final reads), but two calls that overlap on the GPU (different streams / contexts on one device)
// race on it.
// ---------------------------------------------------------------------------------------------
static_assert(MB2_NOALLOC_MAX_N % mb2::DOT_CHUNK == 0, "whole reduction blocks");
constexpr int64_t NOALLOC_MAX_PARTS = MB2_NOALLOC_MAX_N / mb2::DOT_CHUNK;  // 16384 doubles
static __device__ __align__(256) double g_noalloc_part[NOALLOC_MAX_PARTS];
constexpr int NOALLOC_MAX_DEVICES = 64;
static double* g_noalloc_ptr[NOALLOC_MAX_DEVICES];  // device address of g_noalloc_part per device, 0 = not resolved yet
static std::mutex g_noalloc_mutex;

int mb2_dot_noalloc(mb2_ctx* c, const double* d_a, const double* d_b, int64_t n, double* d_out) {
  if (!c) return MB2_ERR_ARG;
  if (!d_a || !d_b || !d_out || n < 0) { c->err = "mb2_dot_noalloc: null pointer or negative length"; return MB2_ERR_ARG; }
  if (n > (int64_t)MB2_NOALLOC_MAX_N) {
    c->err = "mb2_dot_noalloc: n = " + std::to_string(n) + " exceeds MB2_NOALLOC_MAX_N = " + std::to_string((int64_t)MB2_NOALLOC_MAX_N) +
             " (the static scratch holds " + std::to_string(NOALLOC_MAX_PARTS) + " block partials); use the allocating entries";
    return MB2_ERR_ARG;
  }
  if (c->device < 0 || c->device >= NOALLOC_MAX_DEVICES) { c->err = "mb2_dot_noalloc: device index out of range"; return MB2_ERR_ARG; }
  MB2_CUDA(cudaSetDevice(c->device));
  double* part = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[c->device]) {
      double* p = nullptr;
      MB2_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_part));  // this device's instance of the array
      g_noalloc_ptr[c->device] = p;
    }
    part = g_noalloc_ptr[c->device];
  }
  const int64_t nparts = (n + mb2::DOT_CHUNK - 1) / mb2::DOT_CHUNK;
  if (nparts) mb2::k_dot_partial<<<(unsigned)nparts, 256, 0, c->stream>>>(n, d_a, d_b, part);
  mb2::k_dot_final<<<1, 256, 0, c->stream>>>(nparts, part, d_out);
  MB2_CUDA(cudaGetLastError());
  return MB2_OK;
}

}  // extern "C"
