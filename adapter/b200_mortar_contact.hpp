// Reference-side binding of libmimi_b200's rigid-body contact: a mimi::integrators::NonlinearBase subclass that takes
// the place of mimi::integrators::MortarContact (src/mimi/integrators/mortar_contact.hpp:73-79, registered with
// AddBdrFaceIntegrator in src/mimi/py/py_nonlinear_solid.cpp:292-325) and forwards its boundary virtuals to the C ABI
// of include/mimi_b200.h. It shares the mb2_ctx of the domain adapter (adapter/b200_nonlinear_solid.hpp): same CSR
// pattern, same resident u / r / A.
//
// The rigid body: the library evaluates the closest-point query on the device, so the scene handed to the constructor
// must be describable by mb2_contact_set_rigid_body (affine patch, circle / sphere, Bezier curve / surface). A scene
// says so by also implementing B200RigidBody; for the reference's NearestDistanceToSplines
// (coefficients/nearest_distance.hpp:214-281) that is the control net of each splinepy Bezier it holds.
//
// Compiled by oracle/Makefile.ref into oracle/_ref/mimi_ref_b200 (test infrastructure) and exercised on the GPU by
// tests/test_gpu_adapter.py against the golden vectors of the reference's own MortarContact.
#pragma once
#include <algorithm>
#include <memory>
#include <string>
#include <vector>

#include "mimi/coefficients/nearest_distance.hpp"
#include "mimi/integrators/nonlinear_base.hpp"
#include "mimi/utils/precomputed.hpp"
#include "mimi/utils/print.hpp"
#include "mimi_b200.h"

namespace mimi::integrators {

/// what a NearestDistanceBase scene exposes so that the device can run its closest-point query
struct B200RigidBody {
  virtual ~B200RigidBody() = default;
  virtual int B200Kind() const = 0;                     // MB2_RIGID_*
  virtual std::vector<double> B200Params() const = 0;   // layout of mb2_contact_set_rigid_body for that kind
};

/// The reference's own scene class (coefficients/nearest_distance.hpp:214-281) describing itself to the device: the
/// control net / knots / weights are read from the splinepy core it holds (SplinepyCurrentProperties), so a script keeps
/// calling scene.add_spline(curv); scene.plant_kd_tree(resolution, nthreads) and moving curv.cps between steps.
/// Kind by spline type: Bezier -> MB2_RIGID_BEZIER, RationalBezier -> MB2_RIGID_RATIONAL_BEZIER, BSpline ->
/// MB2_RIGID_BSPLINE, NURBS -> MB2_RIGID_NURBS. The kd-tree resolution becomes the device's seed sampling (capped: the
/// device samples every Bezier patch per query instead of planting a tree once).
/// Describes the spline a NearestDistanceToSplines scene holds in the layout of mb2_contact_set_rigid_body.
/// n_seed: sampling of the device's initial guess per Bezier patch and direction.
struct B200SplineScene : mimi::coefficients::NearestDistanceToSplines {
  // the scene keeps its splines protected: reach them through a derived class, read-only
  static const std::vector<std::shared_ptr<splinepy::py::PySpline>>& Splines(
      const mimi::coefficients::NearestDistanceToSplines& s) {
    return s.*(&B200SplineScene::splines_);
  }
  static int Kind(const mimi::coefficients::NearestDistanceToSplines& s) {
    const auto& core = Splines(s).at(0)->Core();
    const bool knots = core->SplinepyHasKnotVectors(), rational = core->SplinepyIsRational();
    return knots ? (rational ? MB2_RIGID_NURBS : MB2_RIGID_BSPLINE) : (rational ? MB2_RIGID_RATIONAL_BEZIER : MB2_RIGID_BEZIER);
  }
  static std::vector<double> Params(const mimi::coefficients::NearestDistanceToSplines& s, const int resolution,
                                    const int max_seed_curve = 256, const int max_seed_surface = 32) {
    const auto& splines = Splines(s);
    if (splines.size() != 1) mimi::utils::PrintAndThrowError("mimi_b200: one spline per contact scene (as the reference asserts)");
    const auto& core = splines[0]->Core();
    const int pd = core->SplinepyParaDim(), dim = core->SplinepyDim(), n = core->SplinepyNumberOfControlPoints();
    const bool knots = core->SplinepyHasKnotVectors(), rational = core->SplinepyIsRational();
    int deg[3] = {0, 0, 0}, res[3] = {1, 1, 1};
    std::vector<std::vector<double>> kv;
    std::vector<double> cps(size_t(n) * dim), w(rational ? n : 0);
    core->SplinepyCurrentProperties(deg, knots ? &kv : nullptr, cps.data(), rational ? w.data() : nullptr);
    core->SplinepyControlMeshResolutions(res);
    const int cap = pd == 1 ? max_seed_curve : max_seed_surface;
    const int n_seed = std::max(1, std::min(resolution - 1, cap));
    std::vector<double> p = {double(deg[0]), pd > 1 ? double(deg[1]) : 0.0, double(n_seed)};
    if (knots) {
      p.push_back(res[0]);
      p.push_back(pd > 1 ? res[1] : 1);
      for (int d = 0; d < pd; ++d) p.insert(p.end(), kv[d].begin(), kv[d].end());
    }
    for (int i = 0; i < n; ++i) {  // splinepy: first parametric direction fastest = the library's [v][u] layout
      for (int c = 0; c < dim; ++c) p.push_back(cps[size_t(i) * dim + c]);
      if (rational) p.push_back(w[i]);
    }
    return p;
  }
};

/// The reference's own scene class (coefficients/nearest_distance.hpp:214-281) that also remembers the kd-tree resolution
/// of plant_kd_tree, which becomes the device's seed sampling (capped: the device samples every Bezier patch per query
/// instead of planting a tree once). A plain NearestDistanceToSplines (what the pybind11 module constructs) works too:
/// B200MortarContact describes it with the default sampling.
class B200NearestDistanceToSplines : public mimi::coefficients::NearestDistanceToSplines, public B200RigidBody {
  int resolution_{65};

public:
  int max_seed_curve_{256}, max_seed_surface_{32};
  void PlantKdTree(const int resolution, const int nthreads) {  // hides the base's: also remembers the sampling
    resolution_ = resolution;
    mimi::coefficients::NearestDistanceToSplines::PlantKdTree(resolution, nthreads);
  }
  int B200Kind() const override { return B200SplineScene::Kind(*this); }
  std::vector<double> B200Params() const override {
    return B200SplineScene::Params(*this, resolution_, max_seed_curve_, max_seed_surface_);
  }
};

class B200MortarContact : public NonlinearBase {
  mb2_ctx* ctx_;
  std::shared_ptr<mimi::coefficients::NearestDistanceBase> scene_;
  // recordings of the last assembly, named as in MortarContact (mortar_contact.hpp:32-61)
  mfem::Vector average_gap_, area_, average_pressure_;
  mimi::utils::Vector<int> marked_boundary_v_dofs_;

  void Check(int rc) const {
    if (rc) mimi::utils::PrintAndThrowError("mimi_b200:", mb2_last_error(ctx_));
  }
  void PushBody() const {  // every call: examples/nl_contact.py moves the body between steps
    if (const auto* body = dynamic_cast<const B200RigidBody*>(scene_.get())) {
      const std::vector<double> p = body->B200Params();
      Check(mb2_contact_set_rigid_body(ctx_, body->B200Kind(), p.data(), (int)p.size(), scene_->coefficient_));
    } else if (const auto* sp = dynamic_cast<const mimi::coefficients::NearestDistanceToSplines*>(scene_.get())) {
      // the reference's own scene as the pybind11 module builds it (py_nearest_distance.cpp:24-32)
      const std::vector<double> p = B200SplineScene::Params(*sp, 257);
      Check(mb2_contact_set_rigid_body(ctx_, B200SplineScene::Kind(*sp), p.data(), (int)p.size(), scene_->coefficient_));
    } else {
      mimi::utils::PrintAndThrowError("mimi_b200: the contact scene does not describe itself (B200RigidBody)");
    }
  }
  void Fetch() {
    Check(mb2_contact_get_nodal(ctx_, average_gap_.GetData(), area_.GetData(), average_pressure_.GetData()));
    double s[5];
    Check(mb2_contact_last_scalars(ctx_, s));
    last_area_ = s[0];
    last_pressure_ = s[1];
    last_force_[0] = s[2];
    last_force_[1] = s[3];
    last_force_[2] = s[4];
  }

public:
  double last_area_{0.0}, last_pressure_{0.0}, last_force_[3] = {0.0, 0.0, 0.0};

  B200MortarContact(const std::shared_ptr<mimi::coefficients::NearestDistanceBase>& nearest_distance_coeff,
                    const std::string& name, const std::shared_ptr<mimi::utils::PrecomputedData>& precomputed,
                    mb2_ctx* domain_ctx)
      : NonlinearBase(name, precomputed),
        ctx_(domain_ctx),
        scene_(nearest_distance_coeff) {}

  /// MortarContact::Prepare (mortar_contact.cpp:19-133), after CreateBoundaryElementQuadData + PrecomputeElementQuadData
  void Prepare() override {
    auto& pre = *precomputed_;
    auto& beqd = pre.GetElementQuadData(Name());
    const int NF = (int)beqd.size(), pd = pre.dim_ - 1;
    const int ndf = beqd[0].GetElementData().n_dof, tdf = beqd[0].GetElementData().n_tdof, nqf = beqd[0].n_quad;
    std::vector<int> vd(size_t(NF) * tdf);
    std::vector<double> N(size_t(NF) * nqf * ndf), dN(N.size() * pd), w(size_t(NF) * nqf);
    for (int f = 0; f < NF; ++f) {
      std::copy_n(beqd[f].GetElementData().v_dofs.GetData(), tdf, &vd[size_t(f) * tdf]);
      for (int q = 0; q < nqf; ++q) {
        const auto& qd = beqd[f].GetQuadData()[q];  // precomputed.hpp:58-69
        w[size_t(f) * nqf + q] = qd.integration_weight;
        std::copy_n(qd.N.GetData(), ndf, &N[(size_t(f) * nqf + q) * ndf]);
        std::copy_n(qd.dN_dxi.GetData(), ndf * pd, &dN[(size_t(f) * nqf + q) * ndf * pd]);
      }
    }
    // x_ref of the boundary nodes, from where MortarContact::Prepare takes it (the faces' point matrices,
    // mortar_contact.cpp:96-107), scattered into a byVDIM node array (only the marked nodes are read by the library)
    std::vector<double> x_ref(size_t(pre.n_v_dofs_), 0.0);
    for (int f = 0; f < NF; ++f) {
      const auto& ed = beqd[f].GetElementData();
      const mfem::DenseMatrix& p_mat = ed.element_trans->GetPointMat();
      for (int c = 0; c < pre.dim_; ++c)
        for (int a = 0; a < ndf; ++a) x_ref[vd[size_t(f) * tdf + c * ndf + a]] = p_mat(c, a);
    }
    Check(mb2_set_nodes(ctx_, x_ref.data(), nullptr));
    Check(mb2_contact_set_faces(ctx_, NF, ndf, nqf, vd.data(), N.data(), dN.data(), w.data()));
    int64_t nm = 0;
    Check(mb2_contact_n_marked_nodes(ctx_, &nm));
    average_gap_.SetSize((int)nm);
    area_.SetSize((int)nm);
    average_pressure_.SetSize((int)nm);
    marked_boundary_v_dofs_.resize(size_t(nm) * pre.dim_);
    Check(mb2_contact_get_marked_v_dofs(ctx_, marked_boundary_v_dofs_.data()));
  }

  void AddBoundaryResidual(const mfem::Vector& u, mfem::Vector& r) override {
    PushBody();
    Check(mb2_add_boundary_residual(ctx_, u.GetData(), r.GetData()));
    Fetch();
  }
  void AddBoundaryResidualAndGrad(const mfem::Vector& u, const double grad_factor, mfem::Vector& r,
                                  mfem::SparseMatrix& A) override {
    PushBody();
    Check(mb2_add_boundary_residual_and_grad(ctx_, u.GetData(), grad_factor, r.GetData(), A.GetData()));
    Fetch();
  }
  double GapNorm(const mfem::Vector& x, const int) override {
    double g = 0.0;
    PushBody();
    Check(mb2_contact_gap_norm(ctx_, x.GetData(), &g));
    return g;
  }
  void BoundaryPostTimeAdvance(const mfem::Vector&) override {}  // recordings are refreshed by every assembly (Fetch)

  const mfem::Vector& Gap() const { return average_gap_; }
  const mfem::Vector& Pressure() const { return average_pressure_; }
  const mfem::Vector& Area() const { return area_; }
  const mimi::utils::Vector<int>& MarkedVDofs() const { return marked_boundary_v_dofs_; }
};

}  // namespace mimi::integrators
