// Reference-side binding of libmimi_b200: a mimi::integrators::NonlinearBase subclass with NonlinearSolid's
// constructor signature (src/mimi/integrators/nonlinear_solid.hpp:45-50) that forwards the domain virtuals of
// src/mimi/integrators/nonlinear_base.hpp:14-154 to the C ABI of include/mimi_b200.h. A maintainer drops this file
// into src/mimi/integrators/ and constructs it in PyNonlinearSolid::Setup (src/mimi/py/py_nonlinear_solid.cpp:202-217)
// instead of NonlinearSolid; forms::Nonlinear, the operators, Newton, the ODE solvers and the pybind11 module stay
// as they are.
//
// It needs MFEM and the reference's headers, so it is NOT part of the product library. oracle/Makefile.ref compiles
// it, together with the unmodified reference sources, into oracle/_ref/mimi_ref_b200 (test infrastructure):
// tests/test_gpu_adapter.py runs that binary on the GPU box and compares what the reference's own call chain
// produces through this adapter with the golden vectors of the reference's CPU integrator.
#pragma once
#include <algorithm>
#include <memory>
#include <string>
#include <vector>

#include "mimi/integrators/nonlinear_base.hpp"
#include "mimi/materials/materials.hpp"
#include "mimi/utils/precomputed.hpp"
#include "mimi/utils/print.hpp"
#include "mimi_b200.h"

namespace mimi::integrators {

class B200NonlinearSolid : public NonlinearBase {
  mb2_ctx* ctx_{nullptr};
  std::shared_ptr<mimi::materials::MaterialBase> material_;
  mutable double* pinned_{nullptr};
  bool pin_jacobian_{false};

  void Check(int rc) const {  // same error convention as the rest of mimi (utils/print.hpp:49-56)
    if (rc) mimi::utils::PrintAndThrowError("mimi_b200:", mb2_last_error(ctx_));
  }

public:
  B200NonlinearSolid(const std::string& name,
                     const std::shared_ptr<mimi::materials::MaterialBase>& material,
                     const std::shared_ptr<mimi::utils::PrecomputedData>& precomputed)
      : NonlinearBase(name, precomputed),
        material_(material) {
    if (mb2_create(&ctx_, /*device*/ 0, /*stream*/ nullptr))
      mimi::utils::PrintAndThrowError("mimi_b200:", mb2_last_error(nullptr));
  }
  ~B200NonlinearSolid() override {
    if (pinned_) mb2_host_unregister(ctx_, pinned_);
    mb2_destroy(ctx_);
  }
  mb2_ctx* Context() const { return ctx_; }  // shared with the contact adapter
  /// operators::NonlinearSolid keeps ONE jacobian_ for the whole run (operators/nonlinear_solid.cpp:114-121): page-lock its
  /// value array once so that the rows stream through the copy engines. Off by default (callers with temporaries).
  void SetPinJacobian(const bool on) { pin_jacobian_ = on; }

  /// NonlinearSolid::Prepare (nonlinear_solid.cpp:21-46): material setup, then the element tables go to the device.
  /// This variant uploads the tables PrecomputedData already holds (precomputed.cpp:264-330); for meshes whose tables
  /// do not fit the host, mb2_set_nodes + mb2_generate_patch_tables rebuild them on the device bit-identically.
  void Prepare() override {
    auto& pre = *precomputed_;
    material_->Setup(pre.dim_);
    auto& eqd = pre.GetElementQuadData("domain");
    const int dim = pre.dim_, NE = pre.n_elements_;
    const int tdof = eqd[0].GetElementData().n_tdof, nd = eqd[0].GetElementData().n_dof, nq = eqd[0].n_quad;
    std::vector<int> vdofs(size_t(NE) * tdof);
    std::vector<double> w(size_t(NE) * nq), dj(size_t(NE) * nq), dX(size_t(NE) * nq * nd * dim);
    for (int e = 0; e < NE; ++e) {
      std::copy_n(eqd[e].GetElementData().v_dofs.GetData(), tdof, &vdofs[size_t(e) * tdof]);
      for (int q = 0; q < nq; ++q) {
        const auto& qd = eqd[e].GetQuadData()[q];  // precomputed.hpp:58-69
        w[size_t(e) * nq + q] = qd.integration_weight;
        dj[size_t(e) * nq + q] = qd.det_dX_dxi;
        std::copy_n(qd.dN_dX.GetData(), nd * dim, &dX[(size_t(e) * nq + q) * nd * dim]);
      }
    }
    Check(mb2_set_connectivity(ctx_, dim, nd, nq, NE, pre.n_v_dofs_, vdofs.data()));
    Check(mb2_upload_quad_tables(ctx_, dX.data(), w.data(), dj.data()));
    int64_t nnz = 0;
    Check(mb2_build_pattern(ctx_, &nnz));  // bit-identical to precomputed_->sparsity_pattern_
    if (pre.sparsity_pattern_ && nnz != pre.sparsity_pattern_->NumNonZeroElems())
      mimi::utils::PrintAndThrowError("mimi_b200: pattern size differs from PrepareSparsity()");
    SetMaterial();
  }

  /// Dispatch on MaterialBase::Name() (materials.hpp); user subclasses the library does not know throw.
  void SetMaterial() { SetMaterialOn(ctx_, *material_, dt_); }
  static void SetMaterialOn(mb2_ctx* ctx_, mimi::materials::MaterialBase& m, const double dt_) {
    namespace mm = mimi::materials;
    auto Check = [ctx_](int rc) {
      if (rc) mimi::utils::PrintAndThrowError("mimi_b200:", mb2_last_error(ctx_));
    };
    const std::string n = m.Name();
    if (n == "StVenantKirchhoff" || n.find("NeoHookean") != std::string::npos) {
      const double par[2] = {m.lambda_, m.mu_};
      Check(mb2_set_material(ctx_, n == "StVenantKirchhoff" ? MB2_MAT_STVK : MB2_MAT_NEOHOOKEAN, par, 2));
    } else if (n == "J2Linear") {
      auto& j = dynamic_cast<mm::J2Linear&>(m);
      const double par[7] = {m.lambda_, m.mu_, m.K_, m.G_, j.isotropic_hardening_, j.kinematic_hardening_, j.sigma_y_};
      Check(mb2_set_material(ctx_, MB2_MAT_J2LINEAR, par, 7));
    } else if (n == "J2" || n == "J2Simo" || n == "J2Log") {  // three classes with the same members
      auto fill = [&](auto& j, int kind) {
        double par[21] = {m.lambda_, m.mu_, m.K_, m.G_};
        auto& hb = *j.hardening_;
        double* hp = par + 5;
        if (auto* pl = dynamic_cast<mm::PowerLawHardening*>(&hb)) {
          par[4] = MB2_HARD_POWERLAW; hp[0] = pl->sigma_y_; hp[1] = pl->n_; hp[2] = pl->eps0_;
        } else if (auto* vo = dynamic_cast<mm::VoceHardening*>(&hb)) {
          par[4] = MB2_HARD_VOCE; hp[0] = vo->sigma_y_; hp[1] = vo->sigma_sat_; hp[2] = vo->strain_constant_;
        } else {  // the JohnsonCook family derives from JohnsonCookHardening
          auto& l = dynamic_cast<mm::JohnsonCookHardening&>(hb);
          hp[0] = l.A_; hp[1] = l.B_; hp[2] = l.n_; par[4] = MB2_HARD_JC;
          if (auto* r = dynamic_cast<mm::JohnsonCookRateDependentHardening*>(&l)) {
            hp[3] = r->C_; hp[4] = r->effective_plastic_strain_rate_; par[4] = MB2_HARD_JC_RATE;
          }
          if (auto* t = dynamic_cast<mm::JohnsonCookTemperatureAndRateDependentHardening*>(&l)) {
            hp[5] = t->reference_temperature_; hp[7] = t->m_;
            par[4] = dynamic_cast<mm::JohnsonCookConstantTemperatureHardening*>(&l) ? MB2_HARD_JC_CONST_T
                                                                                    : MB2_HARD_JC_THERMO_RATE;
          }
        }
        par[15] = dt_; par[16] = j.heat_fraction_; par[17] = m.density_; par[18] = j.specific_heat_;
        par[19] = j.initial_temperature_; par[20] = j.melting_temperature_;
        Check(mb2_set_material(ctx_, kind, par, 21));
      };
      if (n == "J2") fill(dynamic_cast<mm::J2&>(m), MB2_MAT_J2);
      else if (n == "J2Simo") fill(dynamic_cast<mm::J2Simo&>(m), MB2_MAT_J2SIMO);
      else fill(dynamic_cast<mm::J2Log&>(m), MB2_MAT_J2LOG);
    } else {
      mimi::utils::PrintAndThrowError("mimi_b200: unsupported material", n);
    }
  }

  void AddDomainResidual(const mfem::Vector& u, mfem::Vector& r) const override {
    Check(mb2_set_time_step(ctx_, dt_));  // nonlinear_solid.cpp:154: material_->dt_ = dt_
    Check(mb2_add_domain_residual(ctx_, u.GetData(), r.GetData()));
  }

  void AddDomainResidualAndGrad(const mfem::Vector& u, const double grad_factor, mfem::Vector& r,
                                mfem::SparseMatrix& A) const override {
    if (pin_jacobian_ && A.GetData() != pinned_) {
      if (pinned_) Check(mb2_host_unregister(ctx_, pinned_));
      Check(mb2_host_register(ctx_, A.GetData(), sizeof(double) * A.NumNonZeroElems()));
      pinned_ = A.GetData();
    }
    Check(mb2_set_time_step(ctx_, dt_));
    Check(mb2_add_domain_residual_and_grad(ctx_, u.GetData(), grad_factor, r.GetData(), A.GetData()));
  }

  /// commits the device-resident history (no-op for stateless materials)
  void DomainPostTimeAdvance(const mfem::Vector& u) override { Check(mb2_post_time_advance(ctx_, u.GetData())); }

  // ---- the forms PyNonlinearSolid::Setup assembles with MFEM (py_nonlinear_solid.cpp:154-191, 221-241), on the device ----
  /// knot vectors / control-point counts / knot-span indices of the single tensor-product patch (mesh/nurbs.hpp:556-879)
  struct Patch {
    int p{0};
    std::vector<int> ncp, ijk;
    std::vector<std::vector<double>> kv;
    std::vector<const double*> knots;
  };
  Patch DescribePatch(mfem::Mesh& mesh) const {
    mfem::NURBSExtension* ext = mesh.NURBSext;
    if (!ext || ext->GetNP() != 1) mimi::utils::PrintAndThrowError("mimi_b200: device forms need a single-patch NURBS mesh");
    const int dim = mesh.Dimension();
    mfem::Array<const mfem::KnotVector*> kvs;
    ext->GetPatchKnotVectors(0, kvs);
    Patch P;
    P.p = kvs[0]->GetOrder();
    P.kv.resize(dim);
    for (int d = 0; d < dim; ++d) {
      if (kvs[d]->GetOrder() != P.p) mimi::utils::PrintAndThrowError("mimi_b200: mixed degrees are not supported");
      P.ncp.push_back(kvs[d]->GetNCP());
      for (int i = 0; i < kvs[d]->Size(); ++i) P.kv[d].push_back((*kvs[d])[i]);
    }
    for (int d = 0; d < dim; ++d) P.knots.push_back(P.kv[d].data());
    mfem::Array<int> ijk(dim);
    for (int e = 0; e < mesh.GetNE(); ++e) {
      ext->GetElementIJK(e, ijk);
      for (int d = 0; d < dim; ++d) P.ijk.push_back(ijk[d]);
    }
    return P;
  }
  /// node coordinates (byVDIM) and NURBS weights -> mb2_set_nodes
  void SetNodes(mfem::Mesh& mesh) const {
    const mfem::GridFunction& nodes = *mesh.GetNodes();
    const int dim = mesh.Dimension(), nn = nodes.Size() / dim;
    std::vector<double> x(nodes.GetData(), nodes.GetData() + nodes.Size());
    if (nodes.FESpace()->GetOrdering() == mfem::Ordering::byNODES)
      for (int a = 0; a < nn; ++a)
        for (int c = 0; c < dim; ++c) x[size_t(a) * dim + c] = nodes(c * nn + a);
    Check(mb2_set_nodes(ctx_, x.data(), mesh.NURBSext->GetWeights().GetData()));
  }
  /// M <- VectorMassIntegrator(rho); M must carry the library's pattern (a copy of precomputed_->sparsity_pattern_)
  void AssembleMass(mfem::Mesh& mesh, const double rho, mfem::SparseMatrix& M) const {
    const Patch P = DescribePatch(mesh);
    SetNodes(mesh);
    Check(mb2_set_A(ctx_, nullptr));
    Check(mb2_assemble_mass_device(ctx_, P.p, P.ncp.data(), P.knots.data(), P.ijk.data(), -1, rho));
    Check(mb2_get_A(ctx_, M.GetData()));
  }
  /// S <- VectorDiffusionIntegrator(viscosity)
  void AssembleViscosity(mfem::Mesh& mesh, const double viscosity, mfem::SparseMatrix& S) const {
    const Patch P = DescribePatch(mesh);
    SetNodes(mesh);
    Check(mb2_set_A(ctx_, nullptr));
    Check(mb2_assemble_viscosity_device(ctx_, P.p, P.ncp.data(), P.knots.data(), P.ijk.data(), -1, viscosity));
    Check(mb2_get_A(ctx_, S.GetData()));
  }
  /// b <- VectorDomainLFIntegrator(constant force)
  void AssembleBodyForce(mfem::Mesh& mesh, const mfem::Vector& force, mfem::Vector& b) const {
    const Patch P = DescribePatch(mesh);
    SetNodes(mesh);
    Check(mb2_set_r(ctx_, nullptr));
    Check(mb2_assemble_body_force_device(ctx_, P.p, P.ncp.data(), P.knots.data(), P.ijk.data(), -1, force.GetData()));
    Check(mb2_get_r(ctx_, b.GetData()));
  }

  /// b <- ((b + f_1) + f_2) + ...: VectorBoundaryLFIntegrator(constant traction) on the boundary elements whose quadrature
  /// data the caller created under `bdr_name` at the integrator's rule:
  ///   precomputed->CreateBoundaryElementQuadData(bdr_name, marked, n); precomputed->PrecomputeElementQuadData(bdr_name, 2 * p, false);
  /// Call after AssembleBodyForce with the same b (LinearForm::Assemble adds the domain integrators first).
  void AddTraction(const std::string& bdr_name, const mfem::Vector& traction, mfem::Vector& b) const {
    auto& tq = precomputed_->GetElementQuadData(std::string(bdr_name));  // the accessor takes an rvalue (precomputed.hpp:183)
    const int NF = (int)tq.size(), ndf = tq[0].GetElementData().n_dof, tdf = tq[0].GetElementData().n_tdof, nqf = tq[0].n_quad;
    std::vector<int> vd(size_t(NF) * tdf);
    std::vector<double> w(size_t(NF) * nqf), dw(size_t(NF) * nqf), N(size_t(NF) * nqf * ndf);
    for (int f = 0; f < NF; ++f) {
      std::copy_n(tq[f].GetElementData().v_dofs.GetData(), tdf, &vd[size_t(f) * tdf]);
      for (int q = 0; q < nqf; ++q) {
        const auto& qd = tq[f].GetQuadData()[q];
        w[size_t(f) * nqf + q] = qd.integration_weight;
        dw[size_t(f) * nqf + q] = qd.det_dX_dxi;
        std::copy_n(qd.N.GetData(), ndf, &N[(size_t(f) * nqf + q) * ndf]);
      }
    }
    Check(mb2_set_r(ctx_, b.GetData()));
    Check(mb2_assemble_traction_device(ctx_, NF, ndf, nqf, vd.data(), N.data(), w.data(), dw.data(), traction.GetData()));
    Check(mb2_get_r(ctx_, b.GetData()));
  }

  /// history variables [NE][nq][mb2_state_size] (checkpoints, tests); empty for stateless materials
  std::vector<double> State() const {
    const int ns = mb2_state_size(ctx_);
    auto& eqd = precomputed_->GetElementQuadData("domain");
    std::vector<double> st(size_t(precomputed_->n_elements_) * eqd[0].n_quad * std::max(ns, 0));
    if (!st.empty()) Check(mb2_get_state(ctx_, st.data()));
    return st;
  }
};

}  // namespace mimi::integrators
