// Reference-side binding of libmimi_b200's linear solve: an mfem::Solver that takes the place of the
// mfem::GMRESSolver + mfem::DSmoother pair PyNonlinearSolid::Setup builds when `use_iterative_solver` is set
// (src/mimi/py/py_nonlinear_solid.cpp:330-339; Newton drives it as prec->SetOperator(J); prec->Mult(r, c),
// src/mimi/solvers/newton.cpp:125-132). Same algorithm and constants (restarted GMRES(50), Jacobi, rel 1e-8, abs 1e-12,
// 300 iterations) on the device: mb2_newton_solve_gmres_device (include/mimi_b200.h).
//
// SetOperator uploads the values of the Jacobian the operator composed on the host (its pattern is the library's: a copy
// of precomputed_->sparsity_pattern_, operators/nonlinear_solid.cpp:114-121); Mult uploads r and downloads c. A caller
// that composes the Jacobian on the device (mb2_newton_residual_and_grad_device) skips both uploads: SetResident(true).
//
// Test infrastructure compiles it into oracle/_ref/mimi_ref_b200 (harness mode `steps` with MIMI_REF_B200_GMRES=1).
#pragma once
#include <mfem.hpp>

#include "mimi/utils/print.hpp"
#include "mimi_b200.h"

namespace mimi::solvers {

class B200GMRESSolver : public mfem::Solver {
  mb2_ctx* ctx_;
  double rel_tol_{1e-8}, abs_tol_{1e-12};
  int max_iter_{300}, kdim_{50};
  bool resident_{false};
  mutable int iterations_{0}, converged_{0};
  mutable double final_norm_{0.0};

  void Check(int rc) const {
    if (rc) mimi::utils::PrintAndThrowError("mimi_b200:", mb2_last_error(ctx_));
  }

public:
  explicit B200GMRESSolver(mb2_ctx* ctx) : ctx_(ctx) {}
  void SetRelTol(const double t) { rel_tol_ = t; }
  void SetAbsTol(const double t) { abs_tol_ = t; }
  void SetMaxIter(const int n) { max_iter_ = n; }
  void SetKDim(const int k) { kdim_ = k; }
  void SetPrintLevel(int) {}
  /// J and r are already the device's resident CSR values / residual (device-side Newton sweeps)
  void SetResident(const bool on) { resident_ = on; }
  int GetNumIterations() const { return iterations_; }
  bool GetConverged() const { return converged_ != 0; }
  double GetFinalNorm() const { return final_norm_; }

  void SetOperator(const mfem::Operator& op) override {
    const auto* A = dynamic_cast<const mfem::SparseMatrix*>(&op);
    if (!A) mimi::utils::PrintAndThrowError("B200GMRESSolver: the operator must be the mfem::SparseMatrix Jacobian");
    height = width = A->Height();
    if (!resident_) Check(mb2_set_A(ctx_, A->GetData()));
  }
  void Mult(const mfem::Vector& b, mfem::Vector& x) const override {
    if (!resident_) Check(mb2_set_r(ctx_, b.GetData()));
    Check(mb2_newton_solve_gmres_device(ctx_, rel_tol_, abs_tol_, max_iter_, kdim_, &iterations_, &final_norm_, &converged_));
    Check(mb2_newton_get_solution(ctx_, x.GetData()));
  }
};

/// The direct counterpart: takes the place of the mfem::UMFPackSolver PyNonlinearSolid::Setup builds by default
/// (src/mimi/py/py_nonlinear_solid.cpp:340-343) -- mb2_newton_solve_direct_device (cuSOLVER sparse QR on the device CSR).
class B200DirectSolver : public mfem::Solver {
  mb2_ctx* ctx_;
  bool resident_{false};

  void Check(int rc) const {
    if (rc) mimi::utils::PrintAndThrowError("mimi_b200:", mb2_last_error(ctx_));
  }

public:
  explicit B200DirectSolver(mb2_ctx* ctx) : ctx_(ctx) {}
  void SetResident(const bool on) { resident_ = on; }
  void SetOperator(const mfem::Operator& op) override {
    const auto* A = dynamic_cast<const mfem::SparseMatrix*>(&op);
    if (!A) mimi::utils::PrintAndThrowError("B200DirectSolver: the operator must be the mfem::SparseMatrix Jacobian");
    height = width = A->Height();
    if (!resident_) Check(mb2_set_A(ctx_, A->GetData()));
  }
  void Mult(const mfem::Vector& b, mfem::Vector& x) const override {
    if (!resident_) Check(mb2_set_r(ctx_, b.GetData()));
    Check(mb2_newton_solve_direct_device(ctx_, nullptr));
    Check(mb2_newton_get_solution(ctx_, x.GetData()));
  }
};

}  // namespace mimi::solvers
