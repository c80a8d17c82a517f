// Reference-side binding of libmimi_b200 on SEVERAL GPUs from one process: a mimi::integrators::NonlinearBase subclass
// with NonlinearSolid's constructor signature (+ the device list) that owns one mb2_ctx per GPU. PyNonlinearSolid::Setup
// (src/mimi/py/py_nonlinear_solid.cpp:202-217) constructs it instead of NonlinearSolid; no MPI, no torchrun.
//   Prepare():  mb2_plan_partition per rank on the connectivity PrecomputedData holds (contiguous element ranges, node owner
//               = rank of its lowest element), mb2_plan_apply, the rank's slice of the precomputed tables, pattern, material.
//   AddDomainResidual[AndGrad]():  u to every GPU, mb2_mg_group_assemble (raw K_e / R_e of the interface elements travel
//               GPU to GPU by cudaMemcpyPeerAsync), every rank's owned CSR rows / residual entries back into the reference's
//               mfem::SparseMatrix / mfem::Vector. Bit-identical to the single-thread reference (tests/test_gpu_adapter.py).
#pragma once
#include <algorithm>
#include <memory>
#include <string>
#include <vector>

#include "b200_nonlinear_solid.hpp"

namespace mimi::integrators {

class B200MultiGpuNonlinearSolid : public NonlinearBase {
  struct Rank {
    mb2_ctx* ctx{nullptr};
    std::vector<int> rows;          // owned vdofs, ascending
    std::vector<int64_t> runs;      // (global value offset, count) pairs of the owned rows, merged
    int64_t nnz{0};
    std::vector<double> A, r;       // staging of the rank's compact value array / full-length residual
  };
  std::vector<Rank> ranks_;
  std::vector<mb2_ctx*> group_;
  std::vector<int> devices_;
  std::shared_ptr<mimi::materials::MaterialBase> material_;

  static void Check(mb2_ctx* c, int rc) {
    if (rc) mimi::utils::PrintAndThrowError("mimi_b200:", mb2_last_error(c));
  }

public:
  B200MultiGpuNonlinearSolid(const std::string& name, const std::shared_ptr<mimi::materials::MaterialBase>& material,
                             const std::shared_ptr<mimi::utils::PrecomputedData>& precomputed, const std::vector<int>& devices)
      : NonlinearBase(name, precomputed), devices_(devices), material_(material) {
    for (const int d : devices_) {
      Rank R;
      if (mb2_create(&R.ctx, d, nullptr)) mimi::utils::PrintAndThrowError("mimi_b200:", mb2_last_error(nullptr));
      ranks_.push_back(std::move(R));
      group_.push_back(ranks_.back().ctx);
    }
  }
  ~B200MultiGpuNonlinearSolid() override {
    for (auto& R : ranks_) mb2_destroy(R.ctx);
  }
  int NRanks() const { return (int)ranks_.size(); }

  void Prepare() override {
    auto& pre = *precomputed_;
    material_->Setup(pre.dim_);
    auto& eqd = pre.GetElementQuadData("domain");
    const int dim = pre.dim_, NE = pre.n_elements_, NV = pre.n_v_dofs_, world = NRanks();
    const int tdof = eqd[0].GetElementData().n_tdof, nd = eqd[0].GetElementData().n_dof, nq = eqd[0].n_quad;
    std::vector<int> vdofs(size_t(NE) * tdof);
    for (int e = 0; e < NE; ++e) std::copy_n(eqd[e].GetElementData().v_dofs.GetData(), tdof, &vdofs[size_t(e) * tdof]);
    const int* gI = pre.sparsity_pattern_->GetI();
    for (int rk = 0; rk < world; ++rk) {
      Rank& R = ranks_[rk];
      mb2_plan* plan = nullptr;
      if (mb2_plan_partition(vdofs.data(), dim, nd, NE, NV, world, rk, &plan)) mimi::utils::PrintAndThrowError("mimi_b200: partition planner failed");
      Check(R.ctx, mb2_plan_apply(R.ctx, plan, vdofs.data(), nq));
      const int64_t nl = mb2_plan_size(plan, 1);
      std::vector<int64_t> loc(nl);
      mb2_plan_local_elems(plan, loc.data());
      std::vector<double> w(size_t(nl) * nq), dj(size_t(nl) * nq), dX(size_t(nl) * nq * nd * dim);
      for (int64_t x = 0; x < nl; ++x)
        for (int q = 0; q < nq; ++q) {
          const auto& qd = eqd[(int)loc[x]].GetQuadData()[q];
          w[size_t(x) * nq + q] = qd.integration_weight;
          dj[size_t(x) * nq + q] = qd.det_dX_dxi;
          std::copy_n(qd.dN_dX.GetData(), nd * dim, &dX[(size_t(x) * nq + q) * nd * dim]);
        }
      Check(R.ctx, mb2_upload_quad_tables(R.ctx, dX.data(), w.data(), dj.data()));
      Check(R.ctx, mb2_build_pattern(R.ctx, &R.nnz));
      std::vector<uint8_t> owned(size_t(NV / dim));
      mb2_plan_owned_nodes(plan, owned.data());
      mb2_plan_free(plan);
      int64_t count = 0;
      for (int n = 0; n < NV / dim; ++n)
        if (owned[n])
          for (int c = 0; c < dim; ++c) {
            const int v = dim * n + c;
            R.rows.push_back(v);
            if (!R.runs.empty() && R.runs[R.runs.size() - 2] + R.runs.back() == gI[v]) R.runs.back() += gI[v + 1] - gI[v];
            else { R.runs.push_back(gI[v]); R.runs.push_back(gI[v + 1] - gI[v]); }
            count += gI[v + 1] - gI[v];
          }
      if (count != R.nnz) mimi::utils::PrintAndThrowError("mimi_b200: owned rows do not match the rank's pattern");
      R.A.resize(size_t(R.nnz));
      R.r.resize(size_t(NV));
      B200NonlinearSolid::SetMaterialOn(R.ctx, *material_, dt_);
    }
    Check(group_[0], mb2_mg_group_init(group_.data(), world));
  }

  void Assemble(const mfem::Vector& u, const double gf, mfem::Vector& r, mfem::SparseMatrix* A) const {
    auto& ranks = const_cast<std::vector<Rank>&>(ranks_);
    for (auto& R : ranks) {
      Check(R.ctx, mb2_set_time_step(R.ctx, dt_));
      Check(R.ctx, mb2_set_u(R.ctx, u.GetData()));
      Check(R.ctx, mb2_set_r(R.ctx, nullptr));
      if (A) {
        double* o = R.A.data();
        for (size_t k = 0; k < R.runs.size(); k += 2) o = std::copy_n(A->GetData() + R.runs[k], R.runs[k + 1], o);
        Check(R.ctx, mb2_set_A(R.ctx, R.A.data()));
      }
    }
    const int rc = mb2_mg_group_assemble(const_cast<mb2_ctx**>(group_.data()), NRanks(), gf, A ? 1 : 0);
    if (rc)
      for (auto* c : group_) Check(c, *mb2_last_error(c) ? rc : 0);
    for (auto& R : ranks) {
      Check(R.ctx, mb2_get_r(R.ctx, R.r.data()));
      for (const int v : R.rows) r[v] += R.r[size_t(v)];     // nonlinear_base.hpp:96-105: residual[i] += thread_local[i]
      if (A) {
        Check(R.ctx, mb2_get_A(R.ctx, R.A.data()));
        const double* s = R.A.data();
        for (size_t k = 0; k < R.runs.size(); k += 2) {
          std::copy_n(s, R.runs[k + 1], A->GetData() + R.runs[k]);
          s += R.runs[k + 1];
        }
      }
      Check(R.ctx, mb2_check_numeric(R.ctx));
    }
  }
  void AddDomainResidual(const mfem::Vector& u, mfem::Vector& r) const override { Assemble(u, 0.0, r, nullptr); }
  void AddDomainResidualAndGrad(const mfem::Vector& u, const double grad_factor, mfem::Vector& r,
                                mfem::SparseMatrix& A) const override {
    Assemble(u, grad_factor, r, &A);
  }
  /// history variables live with their elements on the rank that computes them
  void DomainPostTimeAdvance(const mfem::Vector& u) override {
    for (auto& R : ranks_) Check(R.ctx, mb2_post_time_advance(R.ctx, u.GetData()));
  }
};

}  // namespace mimi::integrators
