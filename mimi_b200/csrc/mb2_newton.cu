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
};

static NewtonState* NS(mb2_ctx* c) { return static_cast<NewtonState*>(c->newton); }

void free_newton(mb2_ctx* c) {
  NewtonState* S = NS(c);
  if (!S) return;
  for (void* p : {(void*)S->d_M, (void*)S->d_S, (void*)S->d_rhs, (void*)S->d_a, (void*)S->d_tv, (void*)S->d_v, (void*)S->d_x,
                  (void*)S->d_ess})
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

}  // extern "C"
