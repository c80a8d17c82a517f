// TEST INFRASTRUCTURE (oracle/_ref). Drives the UNMODIFIED reference integrator
// (mimi::integrators::NonlinearSolid, /root/reference/src/mimi/integrators/nonlinear_solid.cpp)
// exactly the way PyNonlinearSolid::Setup wires it (/root/reference/src/mimi/py/py_nonlinear_solid.cpp:57-217)
// and dumps / times what it computes. Nothing in the product links or executes this file.
//
//   mimi_ref dump <mesh> <elevate> <refine> <material> <nthreads> <gf> <u.bin|-> <outdir> [lite]
//   mimi_ref time <mesh> <elevate> <refine> <material> <nthreads> <reps> [warmup reps, untimed]
//   mimi_ref contact <mesh> <elevate> <refine> <material> <nthreads> <gf> <u.bin|-> <outdir> <bdr_attr> <penalty>
//            <body> <p0> <p1> ...      rigid body = analytic NearestDistanceBase subclass (below):
//            plane  o[dim] t0[dim] (t1[dim] in 3D) lo hi   closest point on the (clamped) affine patch o + s t0 + t t1
//            sphere c[dim] R                                  closest point on the circle / sphere
//            spline <description>  the reference's own NearestDistanceToSplines on a real splinepy Bezier / BSpline / NURBS
//                                  (the flat description is documented at struct SplineDesc below)
//   Dumps the marked faces' tables and what MortarContact (mortar_contact.cpp) computes on them.
//   mimi_ref proximity ... / steps_contact ...: see the mode blocks in main().
// material: neohookean | stvk | j2linear  (E=2100, nu=0.3, rho=1: tests/test_nonlinear_solid.py:59-62;
//           j2linear adds isotropic_hardening=80, kinematic_hardening=40, sigma_y=<MIMI_REF_SIGMA_Y or 2.0>)
// For materials with history the dump continues: DomainPostTimeAdvance(u) -> state.f64 [NE][nq][19]
// (plastic_strain 9, beta 9, accumulated), then u_b = 1.5 u -> r_b.f64, A_b.f64 on the accumulated state.
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include <mfem.hpp>
#include <omp.h>
#include <Python.h>

#include <splinepy/splines/splinepy_base.hpp>

#include "mimi/coefficients/nearest_distance.hpp"
#include "mimi/integrators/mortar_contact.hpp"
#include "mimi/integrators/nonlinear_solid.hpp"
#include "mimi/materials/materials.hpp"
#include "mimi/forms/nonlinear.hpp"
#include "mimi/operators/nonlinear_solid.hpp"
#include "mimi/solvers/newton.hpp"
#include "mimi/solvers/ode.hpp"
#include "mimi/utils/precomputed.hpp"

// -DMIMI_REF_USE_B200 (target oracle/_ref/mimi_ref_b200): the reference's call chain with its domain integrator replaced
// by the adapter a maintainer would add (adapter/b200_nonlinear_solid.hpp -> C ABI of libmimi_b200.so). Everything
// else in this file -- mesh, FE space, PrecomputedData, materials, mass form, the dumps -- stays the reference's.
#ifdef MIMI_REF_USE_B200
#include "b200_gmres_solver.hpp"
#include "b200_mortar_contact.hpp"
#include "b200_multigpu.hpp"
#include "b200_nonlinear_solid.hpp"
using DomainIntegrator = mimi::integrators::B200NonlinearSolid;
#define MIMI_REF_B200_BODY , mimi::integrators::B200RigidBody
#else
using DomainIntegrator = mimi::integrators::NonlinearSolid;
#define MIMI_REF_B200_BODY
#endif

// splinepy::py::PySpline::Core() is the one symbol of splinepy's *Python binding* (src/py/py_spline.cpp, a pybind11
// module this harness does not build) that the reference's contact scene calls
// (coefficients/nearest_distance.hpp:225-281, NearestDistanceToSplines). Everything behind it -- the spline classes and
// the closest-point search splinepy::proximity::Proximity::VerboseQuery (kd-tree seed, Newton -> SLSQP -> LM,
// third_party/splinepy/src/proximity/proximity.cpp:643-733) -- is the vendored splinepy core, compiled unmodified into
// oracle/_ref/libsplinepy_ref.a by oracle/Makefile.ref.
namespace splinepy::py {
PySpline::CoreSpline_& PySpline::Core() {
  if (!c_spline_) throw std::runtime_error("PySpline: core spline is not initialised");
  return c_spline_;
}
}  // namespace splinepy::py

namespace {

uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
// u_i = 1e-3 (U_i - 0.5), U_i from splitmix64(seed + i): SURVEY.md section 8(d)
double synth_u(uint64_t seed, uint64_t i) {
  const double U = (double)(splitmix64(seed + i) >> 11) * (1.0 / 9007199254740992.0);
  return 1.0e-3 * (U - 0.5);
}

template <typename T>
void dump(const std::string& dir, const char* name, const T* p, size_t n) {
  std::ofstream f(dir + "/" + name, std::ios::binary);
  f.write(reinterpret_cast<const char*>(p), sizeof(T) * n);
}

struct Setup {
  std::unique_ptr<mfem::Mesh> mesh;
  std::unique_ptr<mfem::FiniteElementSpace> fes;
  std::shared_ptr<mimi::utils::PrecomputedData> pre;
  std::shared_ptr<mimi::materials::MaterialBase> mat;
  std::shared_ptr<DomainIntegrator> integ;
  std::unique_ptr<mfem::BilinearForm> mass;
  double t_precompute = 0.0;
};

Setup build(const char* mesh_file, int elevate, int refine, const std::string& material, int nthreads,
            bool with_mass) {
  Setup s;
  // PySolid::SetupNTheads (py_solid.cpp:303-318): the reference sizes its OpenMP team itself. Without this a parent that
  // exports OMP_NUM_THREADS=1 (torch.distributed.run does) would run NThreadExe's `#pragma omp parallel for` on one thread.
  omp_set_num_threads(nthreads > 0 ? nthreads : 1);
  s.mesh = std::make_unique<mfem::Mesh>(mesh_file, 1, 1);  // py_solid.cpp:76
  if (elevate > 0) s.mesh->DegreeElevate(elevate, 50);     // py_solid.cpp:155
  for (int i = 0; i < refine; ++i) s.mesh->UniformRefinement();  // py_solid.cpp:177
  auto* fec = s.mesh->GetNodes()->OwnFEC();
  s.fes = std::make_unique<mfem::FiniteElementSpace>(s.mesh.get(), nullptr, fec, s.mesh->Dimension(),
                                                     mfem::Ordering::byVDIM);
  if (material == "neohookean") s.mat = std::make_shared<mimi::materials::CompressibleOgdenNeoHookean>();
  else if (material == "stvk") s.mat = std::make_shared<mimi::materials::StVenantKirchhoff>();
  else if (material == "j2linear") {
    auto m = std::make_shared<mimi::materials::J2Linear>();
    m->isotropic_hardening_ = 80.0;
    m->kinematic_hardening_ = 40.0;
    const char* sy = std::getenv("MIMI_REF_SIGMA_Y");
    m->sigma_y_ = sy ? std::atof(sy) : 2.0;
    s.mat = m;
  }
  else if (material.rfind("j2_", 0) == 0 || material.rfind("j2simo_", 0) == 0 || material.rfind("j2log_", 0) == 0) {
    // J2 with nonlinear isotropic hardening (materials.hpp:259-404); thermo values of tests/test_nonlinear_solid.py:16-33,
    // yield stresses scaled down so that the synthetic 1e-3 displacement field yields at part of the points
    namespace mm = mimi::materials;
    const bool simo = material.rfind("j2simo_", 0) == 0;
    const bool logm = material.rfind("j2log_", 0) == 0;
    std::shared_ptr<mm::HardeningBase>* hard_slot = nullptr;
    if (logm) {
      auto m = std::make_shared<mm::J2Log>();
      m->melting_temperature_ = 1500; m->initial_temperature_ = 20; m->specific_heat_ = 450; m->heat_fraction_ = 0.9;
      hard_slot = &m->hardening_;
      s.mat = m;
    } else if (simo) {
      auto m = std::make_shared<mm::J2Simo>();
      m->melting_temperature_ = 1500; m->initial_temperature_ = 20; m->specific_heat_ = 450; m->heat_fraction_ = 0.9;
      hard_slot = &m->hardening_;
      s.mat = m;
    } else {
      auto m = std::make_shared<mm::J2>();
      m->melting_temperature_ = 1500; m->initial_temperature_ = 20; m->specific_heat_ = 450; m->heat_fraction_ = 0.9;
      hard_slot = &m->hardening_;
      s.mat = m;
    }
    struct { std::shared_ptr<mm::HardeningBase>& hardening_; } mref{*hard_slot};
    auto* m = &mref;
    const std::string h = material.substr(simo ? 7 : (logm ? 6 : 3));
    auto jc_fill = [](auto& jc) { jc.A_ = 2.0; jc.B_ = 4.0; jc.n_ = 0.2835; };
    auto jcr_fill = [&](auto& jc) { jc_fill(jc); jc.C_ = 0.015; jc.effective_plastic_strain_rate_ = 1.0e-5; };
    if (h == "powerlaw") { auto hd = std::make_shared<mm::PowerLawHardening>(); hd->sigma_y_ = 2.0; hd->n_ = 3.0; hd->eps0_ = 0.01; m->hardening_ = hd; }
    else if (h == "voce") { auto hd = std::make_shared<mm::VoceHardening>(); hd->sigma_y_ = 2.0; hd->sigma_sat_ = 3.0; hd->strain_constant_ = 0.05; m->hardening_ = hd; }
    else if (h == "jc") { auto hd = std::make_shared<mm::JohnsonCookHardening>(); jc_fill(*hd); m->hardening_ = hd; }
    else if (h == "jcrate") { auto hd = std::make_shared<mm::JohnsonCookRateDependentHardening>(); jcr_fill(*hd); m->hardening_ = hd; }
    else if (h == "jcthermo") {
      auto hd = std::make_shared<mm::JohnsonCookTemperatureAndRateDependentHardening>();
      jcr_fill(*hd); hd->reference_temperature_ = 20; hd->m_ = 1.3558; m->hardening_ = hd;
    } else if (h == "jcconst") {
      auto hd = std::make_shared<mm::JohnsonCookConstantTemperatureHardening>();
      jcr_fill(*hd); hd->reference_temperature_ = 10; hd->m_ = 1.3558; m->hardening_ = hd;
    } else if (h == "balken") {  // balken_plasticity of the reference's tests/test_nonlinear_solid.py:16-33 (C_ stays value-initialised)
      auto hd = std::make_shared<mm::JohnsonCookTemperatureAndRateDependentHardening>();
      hd->A_ = 70; hd->B_ = 140; hd->n_ = 0.2835; hd->m_ = 1.3558; hd->effective_plastic_strain_rate_ = 0.004;
      hd->reference_temperature_ = 20; m->hardening_ = hd;
    } else { std::fprintf(stderr, "unknown hardening %s\n", h.c_str()); std::exit(2); }
  }
  else { std::fprintf(stderr, "unknown material %s\n", material.c_str()); std::exit(2); }
  // defaults = the reference's test values (tests/test_nonlinear_solid.py:59-62); the examples use others
  // (examples/nl_contact.py:15-19: density 7e4, E = 1e10): MIMI_REF_DENSITY / MIMI_REF_YOUNG / MIMI_REF_POISSON override
  const char* env_rho = std::getenv("MIMI_REF_DENSITY");
  const char* env_E = std::getenv("MIMI_REF_YOUNG");
  const char* env_nu = std::getenv("MIMI_REF_POISSON");
  s.mat->density_ = env_rho ? std::atof(env_rho) : 1.0;
  s.mat->viscosity_ = -1.0;
  s.mat->SetYoungPoisson(env_E ? std::atof(env_E) : 2100.0, env_nu ? std::atof(env_nu) : 0.3);

  auto t0 = std::chrono::steady_clock::now();
  s.pre = std::make_shared<mimi::utils::PrecomputedData>();
  s.pre->PrepareThreadSafety(*s.fes, nthreads);
  s.pre->PrepareElementData();
  s.pre->PrepareSparsity();
  s.pre->CreateElementQuadData("domain", nullptr);
  if (with_mass) {
    s.mass = std::make_unique<mfem::BilinearForm>(s.fes.get());
    mfem::ConstantCoefficient rho(s.mat->density_);
    s.mass->AddDomainIntegrator(new mfem::VectorMassIntegrator(rho));
    s.mass->Assemble(0);
    s.mass->Finalize(0);
    s.mass->SpMat().SortColumnIndices();  // operators/nonlinear_solid.cpp:86-122 expects sorted CSR
  }
  s.integ = std::make_shared<DomainIntegrator>(s.mat->Name() + "-NonlinearSolid", s.mat, s.pre);
  s.pre->PrecomputeElementQuadData("domain", -1, true);
  s.integ->Prepare();
  s.integ->dt_ = 0.5;  // pushed into the material by AddDomainResidual[AndGrad] (nonlinear_solid.cpp:154,167)
  s.integ->first_effective_dt_ = 0.5;
  s.integ->second_effective_dt_ = 0.5;
  s.t_precompute = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return s;
}

// ---- analytic rigid bodies (closed-form closest point) filling NearestDistanceBase::Results exactly as
// MortarContact reads it (nearest_distance.hpp:139-193): physical_minus_query_, distance_, first_derivatives_
// [para_dim x dim, row-major]. The operation order below is the contract the C oracle (oracle/mimi_oracle.c,
// orc_rigid_nearest) and the CUDA path (mb2_contact.cu, rigid_nearest) restate.
struct RigidPlane : mimi::coefficients::NearestDistanceBase MIMI_REF_B200_BODY {
  int dim = 0;
  double o[3] = {0, 0, 0}, t[2][3] = {{0, 0, 0}, {0, 0, 0}}, lo = -1e300, hi = 1e300;
  void NearestDistance(const Query& query, Results& r) const override {
    const double* q = query.query.data();
    double diff[3];
    for (int d = 0; d < dim; ++d) diff[d] = q[d] - o[d];
    double s[2] = {0, 0};
    for (int k = 0; k < dim - 1; ++k) {
      double num = 0.0, den = 0.0;
      for (int d = 0; d < dim; ++d) num += diff[d] * t[k][d];
      for (int d = 0; d < dim; ++d) den += t[k][d] * t[k][d];
      double v = num / den;
      if (v < lo) v = lo;
      if (v > hi) v = hi;
      s[k] = v;
      r.parametric_[k] = v;
    }
    double sum = 0.0;
    for (int d = 0; d < dim; ++d) {
      double x = o[d] + s[0] * t[0][d];
      if (dim == 3) x = x + s[1] * t[1][d];
      r.physical_[d] = x;
      const double pmq = x - q[d];
      r.physical_minus_query_[d] = pmq;
      sum += pmq * pmq;
    }
    r.distance_ = std::sqrt(sum);
    for (int k = 0; k < dim - 1; ++k)
      for (int d = 0; d < dim; ++d) r.first_derivatives_[k * dim + d] = t[k][d];
  }
  int Size() const override { return 1; }
#ifdef MIMI_REF_USE_B200
  int B200Kind() const override { return MB2_RIGID_PLANE; }
  std::vector<double> B200Params() const override {
    return {o[0], o[1], o[2], t[0][0], t[0][1], t[0][2], t[1][0], t[1][1], t[1][2], lo, hi};
  }
#endif
};

struct RigidSphere : mimi::coefficients::NearestDistanceBase MIMI_REF_B200_BODY {
  int dim = 0;
  double c[3] = {0, 0, 0}, R = 1.0;
  void NearestDistance(const Query& query, Results& r) const override {
    const double* q = query.query.data();
    double diff[3] = {0, 0, 0}, e[3] = {0, 0, 0};
    double n2 = 0.0;
    for (int d = 0; d < dim; ++d) {
      diff[d] = q[d] - c[d];
      n2 += diff[d] * diff[d];
    }
    const double dn = std::sqrt(n2);
    double sum = 0.0;
    for (int d = 0; d < dim; ++d) {
      e[d] = diff[d] / dn;
      const double x = c[d] + R * e[d];
      r.physical_[d] = x;
      const double pmq = x - q[d];
      r.physical_minus_query_[d] = pmq;
      sum += pmq * pmq;
    }
    r.distance_ = std::sqrt(sum);
    double* fd = r.first_derivatives_.data();
    if (dim == 2) {  // ComputeNormal: n = (d1, -d0)/|d|  ->  tangent (-e1, e0) gives n = e (outward)
      fd[0] = -e[1];
      fd[1] = e[0];
    } else {  // t0 = a x e (a = z unless e is near z, then x), t1 = e x t0  ->  t0 x t1 = |t0|^2 e
      if (std::abs(e[2]) < 0.9) {
        fd[0] = -e[1]; fd[1] = e[0]; fd[2] = 0.0;
      } else {
        fd[0] = 0.0; fd[1] = -e[2]; fd[2] = e[1];
      }
      fd[3] = e[1] * fd[2] - e[2] * fd[1];
      fd[4] = e[2] * fd[0] - e[0] * fd[2];
      fd[5] = e[0] * fd[1] - e[1] * fd[0];
    }
  }
  int Size() const override { return 1; }
#ifdef MIMI_REF_USE_B200
  int B200Kind() const override { return MB2_RIGID_SPHERE; }
  std::vector<double> B200Params() const override { return {c[0], c[1], c[2], R}; }
#endif
};

// ---- the reference's own scene: NearestDistanceToSplines holding a real splinepy spline, planted exactly as
// examples/nl_contact.py:31-33 does (scene.add_spline(curv); scene.plant_kd_tree(resolution, nthreads)).
// Flat description (all numbers, the harness' command line): resolution, rational, has_knots, deg_u, deg_v, ncp_u, ncp_v,
// [knots_u (ncp_u + deg_u + 1), knots_v (3D only)] when has_knots, control points [ncp_v][ncp_u][dim (+ weight when rational)]
// (u fastest: splinepy's and BSplineLib's control-point order). 2D scenes: deg_v = 0, ncp_v = 1.
#ifdef MIMI_REF_USE_B200
using SplineScene = mimi::integrators::B200NearestDistanceToSplines;
#else
using SplineScene = mimi::coefficients::NearestDistanceToSplines;
#endif
struct SplineDesc {
  int resolution = 0, rational = 0, has_knots = 0, deg[2] = {0, 0}, ncp[2] = {1, 1}, para_dim = 1, dim = 2;
  std::vector<std::vector<double>> knots;
  std::vector<double> cps, weights;
  // parses bp[k...]; returns the index after the description
  size_t Parse(const std::vector<double>& bp, size_t k, int dim_) {
    dim = dim_;
    para_dim = dim - 1;
    resolution = (int)bp.at(k++);
    rational = (int)bp.at(k++);
    has_knots = (int)bp.at(k++);
    deg[0] = (int)bp.at(k++); deg[1] = (int)bp.at(k++);
    ncp[0] = (int)bp.at(k++); ncp[1] = (int)bp.at(k++);
    knots.clear();
    if (has_knots)
      for (int d = 0; d < para_dim; ++d) {
        knots.emplace_back();
        for (int i = 0; i < ncp[d] + deg[d] + 1; ++i) knots.back().push_back(bp.at(k++));
      }
    const int n = ncp[0] * (para_dim > 1 ? ncp[1] : 1);
    cps.assign((size_t)n * dim, 0.0);
    weights.assign(rational ? n : 0, 1.0);
    for (int i = 0; i < n; ++i) {
      for (int c = 0; c < dim; ++c) cps[(size_t)i * dim + c] = bp.at(k++);
      if (rational) weights[i] = bp.at(k++);
    }
    return k;
  }
  std::shared_ptr<splinepy::splines::SplinepyBase> Core() {
    // what PySpline::NewCore does with the kwargs of sp.Bezier / sp.RationalBezier / sp.BSpline / sp.NURBS
    return splinepy::splines::SplinepyBase::SplinepyCreate(para_dim, dim, deg, has_knots ? &knots : nullptr, cps.data(),
                                                           rational ? weights.data() : nullptr);
  }
};
std::shared_ptr<SplineScene> make_spline_scene(SplineDesc& d, int nthreads) {
  if (!Py_IsInitialized()) Py_Initialize();  // PySpline carries a py::dict member
  auto spline = std::make_shared<splinepy::py::PySpline>(d.Core());
  auto scene = std::make_shared<SplineScene>();
  scene->AddSpline(spline);
  scene->PlantKdTree(d.resolution, nthreads);
  return scene;
}

// exposes MortarContact's protected nodal vectors for the dump
struct ContactProbe : mimi::integrators::MortarContact {
  using mimi::integrators::MortarContact::MortarContact;
  const mfem::Vector& Gap() const { return average_gap_; }
  const mfem::Vector& Pressure() const { return average_pressure_; }
  const mfem::Vector& Area() const { return area_; }
  const mimi::utils::Vector<int>& MarkedVDofs() const { return marked_boundary_v_dofs_; }
};

#ifdef MIMI_REF_USE_B200
using ContactIntegrator = mimi::integrators::B200MortarContact;
#else
using ContactIntegrator = ContactProbe;
#endif

void fill_u(mfem::Vector& u, const char* u_file) {
  if (std::strcmp(u_file, "-") == 0) {
    for (int i = 0; i < u.Size(); ++i) u[i] = synth_u(12345, (uint64_t)i);
  } else {
    std::ifstream f(u_file, std::ios::binary);
    f.read(reinterpret_cast<char*>(u.GetData()), sizeof(double) * u.Size());
    if (!f) { std::fprintf(stderr, "short read of %s\n", u_file); std::exit(2); }
  }
}

}  // namespace

int main(int argc, char** argv) {
  if (argc < 8) { std::fprintf(stderr, "usage: see header of oracle/ref_harness.cpp\n"); return 2; }
  const std::string mode = argv[1];
  const char* mesh_file = argv[2];
  const int elevate = std::atoi(argv[3]), refine = std::atoi(argv[4]);
  const std::string material = argv[5];
  const int nthreads = std::atoi(argv[6]);
  std::cout.setstate(std::ios_base::failbit);  // silence the reference's PrintInfo chatter

  if (mode == "eig") {
    // eig <matrices.f64> <dim> <count> - - <out dir> -: mfem::DenseMatrix::CalcEigenvalues on column-major dim x dim
    // symmetric matrices (the log / exp maps of J2Log go through it) -> lambda.f64 [count][dim], vec.f64 [count][dim*dim]
    const int dim = elevate, count = refine;
    const std::string out = argv[7];
    std::vector<double> in((size_t)count * dim * dim), lam((size_t)count * dim), vec((size_t)count * dim * dim);
    FILE* f = std::fopen(mesh_file, "rb");
    if (!f || std::fread(in.data(), sizeof(double), in.size(), f) != in.size()) { std::fprintf(stderr, "eig: bad input\n"); return 2; }
    std::fclose(f);
    for (int i = 0; i < count; ++i) {
      mfem::DenseMatrix A(dim, dim);
      std::memcpy(A.GetData(), &in[(size_t)i * dim * dim], sizeof(double) * dim * dim);
      A.CalcEigenvalues(&lam[(size_t)i * dim], &vec[(size_t)i * dim * dim]);
    }
    dump(out, "lambda.f64", lam.data(), lam.size());
    dump(out, "vec.f64", vec.data(), vec.size());
    return 0;
  }

  if (mode == "gmres") {
    // gmres <dir with I.i32, J.i32, A.f64, b.f64> <n> <nnz> - - <out dir> -: the reference's iterative linear solver
    // (py_nonlinear_solid.cpp:330-339): GMRESSolver + DSmoother, rel 1e-8, abs 1e-12, 300 iterations -> x.f64, info.f64
    const int n = elevate, nnz = refine;
    const std::string in = mesh_file, out = argv[7];
    std::vector<int> I(n + 1), J(nnz);
    std::vector<double> A(nnz), b(n);
    auto rd = [&](const char* name, void* p, size_t bytes) {
      std::ifstream f(in + "/" + name, std::ios::binary);
      f.read(reinterpret_cast<char*>(p), bytes);
      if (!f) { std::fprintf(stderr, "gmres: short read of %s\n", name); std::exit(2); }
    };
    rd("I.i32", I.data(), sizeof(int) * (n + 1));
    rd("J.i32", J.data(), sizeof(int) * nnz);
    rd("A.f64", A.data(), sizeof(double) * nnz);
    rd("b.f64", b.data(), sizeof(double) * n);
    mfem::SparseMatrix M(I.data(), J.data(), A.data(), n, n, false, false, true);
    mfem::GMRESSolver lin;
    mfem::DSmoother prec;
    lin.SetRelTol(1e-8);
    lin.SetAbsTol(1e-12);
    lin.SetMaxIter(300);
    lin.SetPrintLevel(-1);
    lin.SetPreconditioner(prec);
    lin.SetOperator(M);
    mfem::Vector bv(b.data(), n), x(n);
    x = 0.0;
    lin.Mult(bv, x);
    double info[3] = {(double)lin.GetNumIterations(), lin.GetFinalNorm(), lin.GetConverged() ? 1.0 : 0.0};
    dump(out, "x.f64", x.GetData(), (size_t)n);
    dump(out, "info.f64", info, 3);
    return 0;
  }

  if (mode == "proximity") {
    // proximity <points.f64> <dim> <npoints> - <nthreads> <out dir> <spline description ...>: the reference's closest-point
    // query NearestDistanceToSplines::NearestDistance (coefficients/nearest_distance.hpp:259-281 -> splinepy
    // VerboseProximity) on the given points, exactly as MortarContact::ElementGapAndArea drives it
    // (mortar_contact.cpp:157-181): -> prox.f64 [npoints][para_dim + 2 dim + 2 + para_dim*dim + dim + 1]
    // = parametric, physical, physical - query, distance, convergence, first derivatives, unit normal, normal gap
    const int dim = elevate, np = refine;
    const std::string out = argv[7];
    std::vector<double> bp, pts((size_t)np * dim);
    for (int i = 8; i < argc; ++i) bp.push_back(std::atof(argv[i]));
    FILE* f = std::fopen(mesh_file, "rb");
    if (!f || std::fread(pts.data(), sizeof(double), pts.size(), f) != pts.size()) { std::fprintf(stderr, "proximity: bad input\n"); return 2; }
    std::fclose(f);
    SplineDesc d;
    d.Parse(bp, 0, dim);
    auto scene = make_spline_scene(d, nthreads);
    const int pd = dim - 1, rec = pd + 2 * dim + 2 + pd * dim + dim + 1;
    std::vector<double> res((size_t)np * rec);
    mimi::coefficients::NearestDistanceBase::Query q;
    mimi::coefficients::NearestDistanceBase::Results r;
    q.SetSize(dim);
    r.SetSize(pd, dim);
    for (int i = 0; i < np; ++i) {
      for (int c = 0; c < dim; ++c) q.query[c] = pts[(size_t)i * dim + c];
      scene->NearestDistance(q, r);
      r.ComputeNormal<true>();
      double* o = &res[(size_t)i * rec];
      for (int k = 0; k < pd; ++k) *o++ = r.parametric_[k];
      for (int c = 0; c < dim; ++c) *o++ = r.physical_[c];
      for (int c = 0; c < dim; ++c) *o++ = r.physical_minus_query_[c];
      *o++ = r.distance_;
      *o++ = r.convergence_;
      for (int k = 0; k < pd * dim; ++k) *o++ = r.first_derivatives_[k];
      for (int c = 0; c < dim; ++c) *o++ = r.normal_[c];
      *o++ = r.NormalGap();
    }
    dump(out, "prox.f64", res.data(), res.size());
    return 0;
  }

  if (mode == "dump") {
    const double gf = std::atof(argv[7]);
    const char* u_file = argv[8];
    const std::string out = argv[9];
    const bool lite = argc > 10 && std::strcmp(argv[10], "lite") == 0;
    Setup s = build(mesh_file, elevate, refine, material, nthreads, true);
    auto& pre = *s.pre;
    const int NV = pre.n_v_dofs_, NE = pre.n_elements_;
    mfem::SparseMatrix& sp = *pre.sparsity_pattern_;
    const int nnz = sp.NumNonZeroElems();
    dump(out, "I.i32", sp.GetI(), (size_t)NV + 1);
    dump(out, "J.i32", sp.GetJ(), (size_t)nnz);
    auto& eqd = pre.GetElementQuadData("domain");
    const int tdof = eqd[0].GetElementData().n_tdof, nd = eqd[0].GetElementData().n_dof, nq = eqd[0].n_quad;
    const int dim = pre.dim_;
    std::vector<int> vd((size_t)NE * tdof), aid;
    std::vector<double> w((size_t)NE * nq), dj((size_t)NE * nq), N, dxi, dX((size_t)NE * nq * nd * dim);
    if (!lite) { aid.resize((size_t)NE * tdof * tdof); N.resize((size_t)NE * nq * nd); dxi.resize((size_t)NE * nq * nd * dim); }
    for (int e = 0; e < NE; ++e) {
      const auto& ed = eqd[e].GetElementData();
      std::memcpy(&vd[(size_t)e * tdof], ed.v_dofs.GetData(), sizeof(int) * tdof);
      if (!lite) std::memcpy(&aid[(size_t)e * tdof * tdof], ed.A_ids.data(), sizeof(int) * tdof * tdof);
      for (int q = 0; q < nq; ++q) {
        const auto& qd = eqd[e].GetQuadData()[q];
        const size_t eq = (size_t)e * nq + q;
        w[eq] = qd.integration_weight; dj[eq] = qd.det_dX_dxi;
        std::memcpy(&dX[eq * nd * dim], qd.dN_dX.GetData(), sizeof(double) * nd * dim);
        if (!lite) {
          std::memcpy(&N[eq * nd], qd.N.GetData(), sizeof(double) * nd);
          std::memcpy(&dxi[eq * nd * dim], qd.dN_dxi.GetData(), sizeof(double) * nd * dim);
        }
      }
    }
    dump(out, "v_dofs.i32", vd.data(), vd.size());
    dump(out, "w.f64", w.data(), w.size());
    dump(out, "detJ.f64", dj.data(), dj.size());
    dump(out, "dN_dX.f64", dX.data(), dX.size());
    if (!lite) {
      dump(out, "A_ids.i32", aid.data(), aid.size());
      dump(out, "N.f64", N.data(), N.size());
      dump(out, "dN_dxi.f64", dxi.data(), dxi.size());
    }
    // nodes (x_ref) in the FE space's byVDIM ordering
    mfem::GridFunction& nodes = *s.mesh->GetNodes();
    dump(out, "x_ref.f64", nodes.GetData(), (size_t)nodes.Size());
    // the patch as MFEM's NURBSExtension holds it after elevation / refinement (single-patch meshes): degree, control
    // points per direction, knot vectors, knot-span index of every element, NURBS weight of every node
    if (s.mesh->NURBSext) {
      // p_meta = {degree, dim, number of patches, then per patch: ncp[dim], knot counts[dim]}; knots patch by patch;
      // per element: its patch (el_patch) and its knot-span index inside that patch (el_ijk)
      mfem::NURBSExtension* ext = s.mesh->NURBSext;
      const int NP = ext->GetNP();
      mfem::Array<const mfem::KnotVector*> kvs;
      ext->GetPatchKnotVectors(0, kvs);
      std::vector<int> pm = {kvs[0]->GetOrder(), dim, NP};
      std::vector<double> kn;
      for (int pch = 0; pch < NP; ++pch) {
        ext->GetPatchKnotVectors(pch, kvs);
        for (int d = 0; d < dim; ++d) pm.push_back(kvs[d]->GetNCP());
        for (int d = 0; d < dim; ++d) {
          pm.push_back(kvs[d]->Size());
          for (int i = 0; i < kvs[d]->Size(); ++i) kn.push_back((*kvs[d])[i]);
        }
      }
      std::vector<int> ijk, epatch;
      mfem::Array<int> e_ijk(dim);
      for (int e = 0; e < NE; ++e) {
        ext->GetElementIJK(e, e_ijk);
        for (int d = 0; d < dim; ++d) ijk.push_back(e_ijk[d]);
        epatch.push_back(ext->GetElementPatch(e));
      }
      dump(out, "p_meta.i32", pm.data(), pm.size());
      dump(out, "p_knots.f64", kn.data(), kn.size());
      dump(out, "el_ijk.i32", ijk.data(), ijk.size());
      dump(out, "el_patch.i32", epatch.data(), epatch.size());
      dump(out, "node_w.f64", ext->GetWeights().GetData(), (size_t)ext->GetWeights().Size());
    }

    mfem::Vector u(NV), r(NV), r2(NV);
    fill_u(u, u_file);
    dump(out, "u.f64", u.GetData(), (size_t)NV);
    r = 0.0;
    s.integ->AddDomainResidual(u, r);
    dump(out, "r.f64", r.GetData(), (size_t)NV);
    // Jacobian = copy of the mass CSR (operators/nonlinear_solid.cpp:114-121,257) + gf * K
    mfem::SparseMatrix A(s.mass->SpMat());
    if (A.NumNonZeroElems() != nnz) { std::fprintf(stderr, "mass pattern != sparsity pattern\n"); return 3; }
    dump(out, "mass.f64", A.GetData(), (size_t)nnz);
    {  // body-force linear form as PyNonlinearSolid::Setup builds it (py_nonlinear_solid.cpp:221-241), f = (0.3, -1.7, 0.9)
      mfem::LinearForm lf(s.fes.get());
      mfem::VectorArrayCoefficient bf(dim);
      const double fv[3] = {0.3, -1.7, 0.9};
      for (int c = 0; c < dim; ++c) bf.Set(c, new mfem::ConstantCoefficient(fv[c]));
      lf.AddDomainIntegrator(new mfem::VectorDomainLFIntegrator(bf));
      lf.Assemble();
      dump(out, "bodyforce.f64", lf.GetData(), (size_t)NV);
    }
    {  // viscosity form as PyNonlinearSolid::Setup builds it (py_nonlinear_solid.cpp:174-191), coefficient 0.37
      mfem::BilinearForm visc(s.fes.get());
      mfem::ConstantCoefficient v_coef(0.37);
      visc.AddDomainIntegrator(new mfem::VectorDiffusionIntegrator(v_coef));
      visc.Assemble(0);
      visc.Finalize(0);
      visc.SpMat().SortColumnIndices();
      if (visc.SpMat().NumNonZeroElems() != nnz) { std::fprintf(stderr, "viscosity pattern != sparsity pattern\n"); return 3; }
      dump(out, "viscform.f64", visc.SpMat().GetData(), (size_t)nnz);
    }
    {  // body force + traction in ONE linear form, as PyNonlinearSolid::Setup assembles rhs (py_nonlinear_solid.cpp:221-282):
       // traction (0.5, -0.25, 1.5) on the boundary elements of the highest attribute; and the tables of those faces the
       // way the reference precomputes boundary data, at VectorBoundaryLFIntegrator's rule (order 2p)
      const int attr = s.mesh->bdr_attributes.Max();
      const double tv[3] = {0.5, -0.25, 1.5}, fv[3] = {0.3, -1.7, 0.9};
      mfem::LinearForm lf(s.fes.get());
      mfem::VectorArrayCoefficient bf(dim), tc(dim);
      for (int c = 0; c < dim; ++c) {
        bf.Set(c, new mfem::ConstantCoefficient(fv[c]));
        mfem::Vector pw(attr);
        pw = 0.0;
        pw(attr - 1) = tv[c];
        tc.Set(c, new mfem::PWConstCoefficient(pw));
      }
      lf.AddDomainIntegrator(new mfem::VectorDomainLFIntegrator(bf));
      lf.AddBoundaryIntegrator(new mfem::VectorBoundaryLFIntegrator(tc));
      lf.Assemble();
      dump(out, "rhs_bf_traction.f64", lf.GetData(), (size_t)NV);
      std::vector<int> marked;
      for (int b = 0; b < s.mesh->GetNBE(); ++b)
        if (s.mesh->GetBdrAttribute(b) == attr) marked.push_back(b);
      const int p_order = s.fes->GetFE(0)->GetOrder();
      pre.CreateBoundaryElementQuadData("traction", marked.data(), (int)marked.size());
      pre.PrecomputeElementQuadData("traction", 2 * p_order, false);
      auto& tq = pre.GetElementQuadData("traction");
      const int NF = (int)tq.size(), ndf = tq[0].GetElementData().n_dof, tdf = tq[0].GetElementData().n_tdof, nqf = tq[0].n_quad;
      std::vector<int> tvd((size_t)NF * tdf), tmeta = {NF, ndf, nqf};
      std::vector<double> tw((size_t)NF * nqf), tdet((size_t)NF * nqf), tN((size_t)NF * nqf * ndf);
      for (int f = 0; f < NF; ++f) {
        std::memcpy(&tvd[(size_t)f * tdf], tq[f].GetElementData().v_dofs.GetData(), sizeof(int) * tdf);
        for (int q = 0; q < nqf; ++q) {
          const auto& qd = tq[f].GetQuadData()[q];
          tw[(size_t)f * nqf + q] = qd.integration_weight;
          tdet[(size_t)f * nqf + q] = qd.det_dX_dxi;
          std::memcpy(&tN[((size_t)f * nqf + q) * ndf], qd.N.GetData(), sizeof(double) * ndf);
        }
      }
      dump(out, "t_meta.i32", tmeta.data(), tmeta.size());
      dump(out, "t_v_dofs.i32", tvd.data(), tvd.size());
      dump(out, "t_w.f64", tw.data(), tw.size());
      dump(out, "t_detW.f64", tdet.data(), tdet.size());
      dump(out, "t_N.f64", tN.data(), tN.size());
    }
#ifdef MIMI_REF_USE_B200  // the same three forms through the adapter (device assembly), dumped beside MFEM's
    if (s.mesh->NURBSext->GetNP() == 1) {  // the device forms take one tensor-product patch per call
      mfem::SparseMatrix Mb(sp), Sb(sp);
      s.integ->AssembleMass(*s.mesh, s.mat->density_, Mb);
      dump(out, "mass_b200.f64", Mb.GetData(), (size_t)nnz);
      s.integ->AssembleViscosity(*s.mesh, 0.37, Sb);
      dump(out, "viscform_b200.f64", Sb.GetData(), (size_t)nnz);
      mfem::Vector fv(dim), bb(NV);
      const double fc[3] = {0.3, -1.7, 0.9};
      for (int c = 0; c < dim; ++c) fv[c] = fc[c];
      s.integ->AssembleBodyForce(*s.mesh, fv, bb);
      dump(out, "bodyforce_b200.f64", bb.GetData(), (size_t)NV);
      mfem::Vector tvv(dim);
      const double tcv[3] = {0.5, -0.25, 1.5};
      for (int c = 0; c < dim; ++c) tvv[c] = tcv[c];
      s.integ->AddTraction("traction", tvv, bb);  // the "traction" boundary data were created above at the rule 2p
      dump(out, "rhs_b200.f64", bb.GetData(), (size_t)NV);
    }
#endif
    r2 = 0.0;
    s.integ->AddDomainResidualAndGrad(u, gf, r2, A);
    dump(out, "r2.f64", r2.GetData(), (size_t)NV);
    dump(out, "A.f64", A.GetData(), (size_t)nnz);
#ifdef MIMI_REF_USE_B200
    if (const char* ng = std::getenv("MIMI_REF_B200_GPUS")) {
      // the same two calls through adapter/b200_multigpu.hpp: N contexts in this process (devices round-robin over the
      // visible GPUs), partition planned by the library, element matrices exchanged GPU to GPU
      const char* nd_env = std::getenv("MIMI_REF_B200_NDEV");   // visible GPUs (the harness does not link the CUDA runtime)
      const int ndev = nd_env ? std::atoi(nd_env) : 1;
      std::vector<int> devs;
      for (int i = 0; i < std::atoi(ng); ++i) devs.push_back(i % std::max(ndev, 1));
      auto mg = std::make_shared<mimi::integrators::B200MultiGpuNonlinearSolid>("mg", s.mat, s.pre, devs);
      mg->dt_ = s.integ->dt_;
      mg->Prepare();
      mfem::Vector rm(NV), rm2(NV);
      rm = 0.0;
      mg->AddDomainResidual(u, rm);
      dump(out, "r_mg.f64", rm.GetData(), (size_t)NV);
      mfem::SparseMatrix Am(s.mass->SpMat());
      rm2 = 0.0;
      mg->AddDomainResidualAndGrad(u, gf, rm2, Am);
      mg->AddDomainResidualAndGrad(u, gf, rm2, Am);   // adds into: twice
      dump(out, "r2_mg.f64", rm2.GetData(), (size_t)NV);
      dump(out, "A_mg.f64", Am.GetData(), (size_t)nnz);
    }
#endif
    if (material == "j2linear" || material.rfind("j2", 0) == 0) {
      s.integ->DomainPostTimeAdvance(u);
      // history layout: every state matrix in a 9-double slot (column-major dim x dim in front), then the scalars;
      // J2Linear has k_state_scalars = 2 but only uses the first (19 doubles), J2 = 9 + 2
      const bool lin = material == "j2linear";
#ifdef MIMI_REF_USE_B200
      (void)lin;
      std::vector<double> st = s.integ->State();  // the history lives on the device, same [NE][nq][ns] layout
#else
      const int ns = lin ? 19 : 9 * (int)eqd[0].GetQuadData()[0].material_state->matrices.size() + 2;
      std::vector<double> st((size_t)NE * nq * ns, 0.0);
      for (int e = 0; e < NE; ++e)
        for (int q = 0; q < nq; ++q) {
          const auto& ms = *eqd[e].GetQuadData()[q].material_state;
          double* o = &st[((size_t)e * nq + q) * ns];
          for (size_t k = 0; k < ms.matrices.size(); ++k)
            std::memcpy(o + 9 * k, ms.matrices[k].GetData(), sizeof(double) * dim * dim);
          o[9 * ms.matrices.size()] = ms.scalars[0];
          if (!lin) o[9 * ms.matrices.size() + 1] = ms.scalars[1];
        }
#endif
      dump(out, "state.f64", st.data(), st.size());
      mfem::Vector ub(u), rb(NV);
      ub *= 1.5;
      rb = 0.0;
      mfem::SparseMatrix Ab(s.mass->SpMat());
      s.integ->AddDomainResidualAndGrad(ub, gf, rb, Ab);
      dump(out, "r_b.f64", rb.GetData(), (size_t)NV);
      dump(out, "A_b.f64", Ab.GetData(), (size_t)nnz);
    }
    std::ofstream m(out + "/meta.json");
    m << "{\"NV\": " << NV << ", \"NE\": " << NE << ", \"nnz\": " << nnz << ", \"nd\": " << nd << ", \"nq\": " << nq
      << ", \"dim\": " << dim << ", \"tdof\": " << tdof << ", \"gf\": " << gf << ", \"nthreads\": " << nthreads
      << ", \"material\": \"" << material << "\", \"lambda\": \"" << std::hexfloat << s.mat->lambda_ << "\", \"mu\": \""
      << s.mat->mu_ << std::defaultfloat << "\", \"lite\": " << (lite ? 1 : 0) << "}\n";
    return 0;
  }
  if (mode == "contact") {
    if (argc < 13) { std::fprintf(stderr, "contact: too few arguments\n"); return 2; }
    const double gf = std::atof(argv[7]);
    const char* u_file = argv[8];
    const std::string out = argv[9];
    const int bdr_attr = std::atoi(argv[10]);
    const double penalty = std::atof(argv[11]);
    const std::string body = argv[12];
    Setup s = build(mesh_file, elevate, refine, material, nthreads, true);
    auto& pre = *s.pre;
    const int dim = pre.dim_, NV = pre.n_v_dofs_;
    std::vector<double> bp;
    for (int i = 13; i < argc; ++i) bp.push_back(std::atof(argv[i]));
    std::shared_ptr<mimi::coefficients::NearestDistanceBase> nd;
    if (body == "plane") {
      auto b = std::make_shared<RigidPlane>();
      b->dim = dim;
      size_t k = 0;
      for (int d = 0; d < dim; ++d) b->o[d] = bp.at(k++);
      for (int j = 0; j < dim - 1; ++j)
        for (int d = 0; d < dim; ++d) b->t[j][d] = bp.at(k++);
      b->lo = bp.at(k++);
      b->hi = bp.at(k++);
      nd = b;
    } else if (body == "sphere") {
      auto b = std::make_shared<RigidSphere>();
      b->dim = dim;
      size_t k = 0;
      for (int d = 0; d < dim; ++d) b->c[d] = bp.at(k++);
      b->R = bp.at(k++);
      nd = b;
    } else if (body == "spline") {
      SplineDesc& d = *new SplineDesc();  // outlives the scene: splinepy's BSpline / NURBS only VIEW the control points
      d.Parse(bp, 0, dim);
      nd = make_spline_scene(d, nthreads);
    } else { std::fprintf(stderr, "unknown rigid body %s\n", body.c_str()); return 2; }
    nd->coefficient_ = penalty;
    nd->dim_ = dim;
    nd->para_dim_ = dim - 1;

    // wiring of py_nonlinear_solid.cpp:292-325
    mfem::Array<int> marker(s.mesh->bdr_attributes.Max());
    marker = 0;
    marker[bdr_attr - 1] = 1;
#ifdef MIMI_REF_USE_B200
    auto integ = std::make_shared<ContactIntegrator>(nd, "contact" + std::to_string(bdr_attr), s.pre, s.integ->Context());
#else
    auto integ = std::make_shared<ContactIntegrator>(nd, "contact" + std::to_string(bdr_attr), s.pre);
#endif
    integ->runtime_communication_ = std::make_shared<mimi::utils::RuntimeCommunication>();
    integ->SetBoundaryMarker(&marker);
    auto& marked = integ->MarkedBoundaryElements();
    pre.CreateBoundaryElementQuadData(integ->Name(), marked.data(), (int)marked.size());
    pre.PrecomputeElementQuadData(integ->Name(), -1, false);
    integ->Prepare();

    auto& beqd = pre.GetElementQuadData(integ->Name());
    const int NF = (int)beqd.size();
    const int ndf = beqd[0].GetElementData().n_dof, tdf = beqd[0].GetElementData().n_tdof, nqf = beqd[0].n_quad;
    std::vector<int> vd((size_t)NF * tdf), aid((size_t)NF * tdf * tdf);
    std::vector<double> w((size_t)NF * nqf), N((size_t)NF * nqf * ndf), dxi((size_t)NF * nqf * ndf * (dim - 1)),
        xr((size_t)NF * ndf * dim);
    for (int f = 0; f < NF; ++f) {
      const auto& ed = beqd[f].GetElementData();
      std::memcpy(&vd[(size_t)f * tdf], ed.v_dofs.GetData(), sizeof(int) * tdf);
      std::memcpy(&aid[(size_t)f * tdf * tdf], ed.A_ids.data(), sizeof(int) * tdf * tdf);
#ifdef MIMI_REF_USE_B200  // MortarContact::Prepare's per-face x_ref matrices do not exist: same values from their source
      for (int a = 0; a < ndf; ++a)
        for (int c = 0; c < dim; ++c) xr[(size_t)f * ndf * dim + a + (size_t)c * ndf] = ed.element_trans->GetPointMat()(c, a);
#else
      std::memcpy(&xr[(size_t)f * ndf * dim], beqd[f].GetMatrix(0).GetData(), sizeof(double) * ndf * dim);  // kXRef = 0
#endif
      for (int q = 0; q < nqf; ++q) {
        const auto& qd = beqd[f].GetQuadData()[q];
        const size_t fq = (size_t)f * nqf + q;
        w[fq] = qd.integration_weight;
        std::memcpy(&N[fq * ndf], qd.N.GetData(), sizeof(double) * ndf);
        std::memcpy(&dxi[fq * ndf * (dim - 1)], qd.dN_dxi.GetData(), sizeof(double) * ndf * (dim - 1));
      }
    }
    {  // domain connectivity (the device builds the CSR pattern from it)
      auto& eqd = pre.GetElementQuadData("domain");
      const int tdof = eqd[0].GetElementData().n_tdof;
      std::vector<int> dvd((size_t)pre.n_elements_ * tdof);
      for (int e = 0; e < pre.n_elements_; ++e)
        std::memcpy(&dvd[(size_t)e * tdof], eqd[e].GetElementData().v_dofs.GetData(), sizeof(int) * tdof);
      dump(out, "v_dofs.i32", dvd.data(), dvd.size());
    }
    dump(out, "c_marked.i32", marked.data(), marked.size());
    dump(out, "c_v_dofs.i32", vd.data(), vd.size());
    dump(out, "c_A_ids.i32", aid.data(), aid.size());
    dump(out, "c_w.f64", w.data(), w.size());
    dump(out, "c_N.f64", N.data(), N.size());
    dump(out, "c_dN_dxi.f64", dxi.data(), dxi.size());
    dump(out, "c_x_ref.f64", xr.data(), xr.size());
    dump(out, "c_marked_v_dofs.i32", integ->MarkedVDofs().data(), integ->MarkedVDofs().size());
    mfem::SparseMatrix& sp = *pre.sparsity_pattern_;
    const int nnz = sp.NumNonZeroElems();
    dump(out, "I.i32", sp.GetI(), (size_t)NV + 1);
    dump(out, "J.i32", sp.GetJ(), (size_t)nnz);
    mfem::GridFunction& nodes = *s.mesh->GetNodes();
    dump(out, "x_ref.f64", nodes.GetData(), (size_t)nodes.Size());

    mfem::Vector u(NV), r(NV), r2(NV);
    fill_u(u, u_file);
    dump(out, "u.f64", u.GetData(), (size_t)NV);
    // AddBoundaryResidual on top of a non-zero residual (the domain residual): the faces add into it in order
    r = 0.0;
    s.integ->AddDomainResidual(u, r);
    dump(out, "c_r0.f64", r.GetData(), (size_t)NV);
    integ->AddBoundaryResidual(u, r);
    dump(out, "c_r.f64", r.GetData(), (size_t)NV);
    dump(out, "c_gap.f64", integ->Gap().GetData(), (size_t)integ->Gap().Size());
    dump(out, "c_area.f64", integ->Area().GetData(), (size_t)integ->Area().Size());
    dump(out, "c_pressure.f64", integ->Pressure().GetData(), (size_t)integ->Pressure().Size());
    double scal[8] = {integ->last_area_, integ->last_pressure_, integ->last_force_[0], integ->last_force_[1],
                      dim > 2 ? integ->last_force_[2] : 0.0, 0, 0, 0};
    mfem::SparseMatrix A(s.mass->SpMat());
    dump(out, "mass.f64", A.GetData(), (size_t)nnz);
    r2 = 0.0;
    integ->AddBoundaryResidualAndGrad(u, gf, r2, A);
    dump(out, "c_r2.f64", r2.GetData(), (size_t)NV);
    dump(out, "c_A.f64", A.GetData(), (size_t)nnz);
    scal[5] = integ->GapNorm(u, 1);
    scal[6] = integ->last_area_;
    scal[7] = integ->last_pressure_;
    dump(out, "c_scalars.f64", scal, 8);
    std::ofstream m(out + "/meta.json");
    m << "{\"NV\": " << NV << ", \"NE\": " << pre.n_elements_ << ", \"nnz\": " << nnz << ", \"dim\": " << dim
      << ", \"NF\": " << NF << ", \"ndf\": " << ndf << ", \"nqf\": " << nqf << ", \"gf\": " << gf
      << ", \"penalty\": " << penalty << ", \"n_marked_nodes\": " << integ->Area().Size()
      << ", \"nd\": " << pre.GetElementQuadData("domain")[0].GetElementData().n_dof
      << ", \"nq\": " << pre.GetElementQuadData("domain")[0].n_quad << "}\n";
    return 0;
  }
  if (mode == "steps") {
    // steps <mesh> <elevate> <refine> <material> <nthreads> <body force on component 1> <dt> <n steps> <out dir>
    // The reference's own time loop (tests/test_nonlinear_solid.py:54-112): PyNonlinearSolid::Setup
    // (py_nonlinear_solid.cpp:155-386) restated on the reference's classes -- forms::Nonlinear, operators::NonlinearSolid,
    // LineSearchNewton, GeneralizedAlpha2 -- with essential dofs on boundary id 2, rho_inf = 0.5 and Newton
    // (1e-12, 1e-8, 10, false). UMFPACK is absent from this image: the 50-dof systems are solved by a dense LU.
    const double force = std::atof(argv[7]);
    double dt = std::atof(argv[8]);
    const int nsteps = std::atoi(argv[9]);
    const std::string out = argv[10];
    Setup s = build(mesh_file, elevate, refine, material, nthreads, false);
    auto& fes = *s.fes;
    const int dim = s.pre->dim_, NV = s.pre->n_v_dofs_;
    mfem::Array<int> zero_dofs;
    for (int c = 0; c < dim; ++c) {
      mfem::Array<int> marker(s.mesh->bdr_attributes.Max()), list;
      marker = 0;
      marker[2] = 1;  // bc.initial.dirichlet(2, c): PySolid::FindBoundaryDofIds, py_solid.cpp:184-234
      fes.GetEssentialTrueDofs(marker, list, c);
      zero_dofs.Append(list);
    }
    zero_dofs.Sort();
    zero_dofs.Unique();
    auto nl_oper = std::make_unique<mimi::operators::NonlinearSolid>(fes);
    mfem::SparseMatrix tmp;
    auto mass = std::make_shared<mfem::BilinearForm>(&fes);
    nl_oper->AddBilinearForm("mass", mass);
    mfem::ConstantCoefficient rho(s.mat->density_);
    mass->AddDomainIntegrator(new mfem::VectorMassIntegrator(rho));
    mass->Assemble(0);
    mass->FormSystemMatrix(zero_dofs, tmp);
    auto stiffness = std::make_shared<mimi::forms::Nonlinear>(&fes);
    nl_oper->AddNonlinearForm("nonlinear_stiffness", stiffness);
    stiffness->AddDomainIntegrator(s.integ);
    stiffness->SetEssentialTrueDofs(zero_dofs);
    auto rhs = std::make_shared<mfem::LinearForm>(&fes);
    mfem::VectorArrayCoefficient bf(dim);
    bf.Set(1, new mfem::ConstantCoefficient(force));
    rhs->AddDomainIntegrator(new mfem::VectorDomainLFIntegrator(bf));
    rhs->Assemble();
    rhs->SetSubVector(zero_dofs, 0.0);
    nl_oper->AddLinearForm("rhs", rhs);
    struct DenseLU : mfem::Solver {
      std::unique_ptr<mfem::DenseMatrix> d;
      std::unique_ptr<mfem::DenseMatrixInverse> inv;
      void SetOperator(const mfem::Operator& op) override {
        d.reset(dynamic_cast<const mfem::SparseMatrix&>(op).ToDenseMatrix());
        inv = std::make_unique<mfem::DenseMatrixInverse>(*d);
        height = width = d->Height();
      }
      void Mult(const mfem::Vector& b, mfem::Vector& x) const override { inv->Mult(b, x); }
    } lin;
    // MIMI_REF_LIN=gmres: the reference's iterative option (`use_iterative_solver`, py_nonlinear_solid.cpp:330-339) -- the
    // CPU binary builds mfem::GMRESSolver + mfem::DSmoother with the reference's constants, the B200 binary puts
    // adapter/b200_gmres_solver.hpp in their place (device GMRES on the library's resident CSR). Either way the linear
    // iteration count of every Newton iteration is recorded -> lin_its.i32.
    struct Counting : mfem::Solver {
      mfem::Solver* inner = nullptr;
      std::function<int()> its;
      mutable std::vector<int> log;
      void SetOperator(const mfem::Operator& op) override {
        inner->SetOperator(op);
        height = width = op.Height();
      }
      void Mult(const mfem::Vector& b, mfem::Vector& x) const override {
        // LineSearchNewton::Mult sets prec->iterative_mode = false (solvers/newton.cpp:74) on the solver it was given: hand
        // that on, as if the inner solver had been registered directly (an IterativeSolver defaults to true)
        inner->iterative_mode = iterative_mode;
        inner->Mult(b, x);
        log.push_back(its());
      }
    } counting;
    mfem::GMRESSolver gm;
    mfem::DSmoother dprec;
    const char* env_lin = std::getenv("MIMI_REF_LIN");
    const bool use_gmres = env_lin && std::string(env_lin) == "gmres";
#ifdef MIMI_REF_USE_B200
    std::unique_ptr<mimi::solvers::B200GMRESSolver> bg;
    // MIMI_REF_LIN=direct (B200 binary only): adapter/b200_gmres_solver.hpp's B200DirectSolver (cuSOLVER sparse QR on the
    // device CSR) in the place of the direct solver (UMFPACK in the reference, the dense LU above in this harness)
    std::unique_ptr<mimi::solvers::B200DirectSolver> bd;
    if (env_lin && std::string(env_lin) == "direct") bd = std::make_unique<mimi::solvers::B200DirectSolver>(s.integ->Context());
#endif
    if (use_gmres) {
#ifdef MIMI_REF_USE_B200
      bg = std::make_unique<mimi::solvers::B200GMRESSolver>(s.integ->Context());
      counting.inner = bg.get();
      counting.its = [&bg]() { return bg->GetNumIterations(); };
#else
      gm.SetRelTol(1e-8);
      gm.SetAbsTol(1e-12);
      gm.SetMaxIter(300);
      gm.SetPrintLevel(-1);
      gm.SetPreconditioner(dprec);
      counting.inner = &gm;
      counting.its = [&gm]() { return gm.GetNumIterations(); };
#endif
    }
    auto newton = std::make_shared<mimi::solvers::LineSearchNewton>();
    newton->nl_oper_ = nl_oper.get();
    newton->iterative_mode = false;
    newton->SetOperator(*nl_oper);
    if (use_gmres) newton->SetSolver(counting);
#ifdef MIMI_REF_USE_B200
    else if (bd) newton->SetSolver(*bd);
#endif
    else newton->SetSolver(lin);
    newton->SetPrintLevel(-1);
    newton->SetRelTol(1e-12);
    newton->SetAbsTol(1e-8);
    newton->SetMaxIter(10);
    nl_oper->SetNewtonSolver(newton);
    // rho_inf: the reference's tests set ode_coefficient = 0.5; PyNonlinearSolid::Setup's default, which examples/*.py run
    // with, is 0.25 (py_nonlinear_solid.cpp:367-368): MIMI_REF_RHO_INF overrides
    const char* env_rho_inf = std::getenv("MIMI_REF_RHO_INF");
    mimi::solvers::GeneralizedAlpha2 ode(*nl_oper, env_rho_inf ? std::atof(env_rho_inf) : 0.5);
    nl_oper->Setup();
    ode.SetupDirichletDofs(nl_oper->dirichlet_dofs_);
    mfem::Vector x(NV), v(NV);
    x = 0.0;
    v = 0.0;
    double t = 0.0;
    std::vector<double> hist;
    std::vector<int> its;
    for (int i = 0; i < nsteps; ++i) {
      ode.StepTime2(x, v, t, dt);
      hist.insert(hist.end(), x.GetData(), x.GetData() + NV);
      its.push_back(newton->GetNumIterations());
    }
    dump(out, "x_hist.f64", hist.data(), hist.size());
    dump(out, "newton_its.i32", its.data(), its.size());
    if (use_gmres) dump(out, "lin_its.i32", counting.log.data(), counting.log.size());
    return 0;
  }

  if (mode == "steps_contact") {
    // steps_contact <mesh> <elevate> <refine> <material> <nthreads> <dt> <n steps> <out dir> <bdr attr of the contact
    //   boundary> <bdr attr of the clamped boundary> <penalty> <dx per step> <dy per step> [<dz per step>] <spline description ...>
    // The contact time loop of examples/nl_contact.py: PyNonlinearSolid::Setup (py_nonlinear_solid.cpp:155-386) restated on
    // the reference's own classes -- forms::Nonlinear for the stiffness and for "contact" with MortarContact registered as in
    // py_nonlinear_solid.cpp:286-326, operators::NonlinearSolid, LineSearchNewton (nl_contact.py: 1e-10, 1e-8, 100
    // iterations), GeneralizedAlpha2(0.25) -- with the rigid spline moved before every step and its kd-tree re-planted
    // (nl_contact.py:58-64). UMFPACK is absent from this image: a dense LU solves the linear systems.
    // -> x_hist.f64 [steps][NV], newton_its.i32 [steps], contact_hist.f64 [steps][6] = area, pressure, force x/y/z, gap norm
    double dt = std::atof(argv[7]);
    const int nsteps = std::atoi(argv[8]);
    const std::string out = argv[9];
    const int bdr_contact = std::atoi(argv[10]), bdr_clamp = std::atoi(argv[11]);
    const double penalty = std::atof(argv[12]);
    Setup s = build(mesh_file, elevate, refine, material, nthreads, false);
    auto& fes = *s.fes;
    auto& pre = *s.pre;
    const int dim = pre.dim_, NV = pre.n_v_dofs_;
    double shift[3] = {0, 0, 0};
    int k0 = 13;
    for (int c = 0; c < dim; ++c) shift[c] = std::atof(argv[k0++]);
    std::vector<double> bp;
    for (int i = k0; i < argc; ++i) bp.push_back(std::atof(argv[i]));
    SplineDesc d;
    d.Parse(bp, 0, dim);
    auto scene = make_spline_scene(d, nthreads);
    scene->coefficient_ = penalty;
    scene->dim_ = dim;
    scene->para_dim_ = dim - 1;

    mfem::Array<int> zero_dofs;
    for (int c = 0; c < dim; ++c) {
      mfem::Array<int> marker(s.mesh->bdr_attributes.Max()), list;
      marker = 0;
      marker[bdr_clamp - 1] = 1;   // bc.initial.dirichlet(bid, c): boundary ids are 0-based there, attributes 1-based
      fes.GetEssentialTrueDofs(marker, list, c);
      zero_dofs.Append(list);
    }
    zero_dofs.Sort();
    zero_dofs.Unique();
    auto nl_oper = std::make_unique<mimi::operators::NonlinearSolid>(fes);
    mfem::SparseMatrix tmp;
    auto mass = std::make_shared<mfem::BilinearForm>(&fes);
    nl_oper->AddBilinearForm("mass", mass);
    mfem::ConstantCoefficient rho(s.mat->density_);
    mass->AddDomainIntegrator(new mfem::VectorMassIntegrator(rho));
    mass->Assemble(0);
    mass->FormSystemMatrix(zero_dofs, tmp);
    auto stiffness = std::make_shared<mimi::forms::Nonlinear>(&fes);
    nl_oper->AddNonlinearForm("nonlinear_stiffness", stiffness);
    stiffness->AddDomainIntegrator(s.integ);
    stiffness->SetEssentialTrueDofs(zero_dofs);
    // contact form (py_nonlinear_solid.cpp:286-326)
    mfem::Array<int> cmarker(s.mesh->bdr_attributes.Max());
    cmarker = 0;
    cmarker[bdr_contact - 1] = 1;
    auto contact_form = std::make_shared<mimi::forms::Nonlinear>(&fes);
    nl_oper->AddNonlinearForm("contact", contact_form);
#ifdef MIMI_REF_USE_B200
    auto cinteg = std::make_shared<ContactIntegrator>(scene, "contact" + std::to_string(bdr_contact), s.pre, s.integ->Context());
#else
    auto cinteg = std::make_shared<ContactIntegrator>(scene, "contact" + std::to_string(bdr_contact), s.pre);
#endif
    cinteg->runtime_communication_ = std::make_shared<mimi::utils::RuntimeCommunication>();
    cinteg->SetBoundaryMarker(&cmarker);
    auto& marked = cinteg->MarkedBoundaryElements();
    pre.CreateBoundaryElementQuadData(cinteg->Name(), marked.data(), (int)marked.size());
    pre.PrecomputeElementQuadData(cinteg->Name(), -1, false);
    cinteg->Prepare();
    contact_form->AddBdrFaceIntegrator(cinteg, &cmarker);

    struct DenseLU : mfem::Solver {
      std::unique_ptr<mfem::DenseMatrix> d;
      std::unique_ptr<mfem::DenseMatrixInverse> inv;
      void SetOperator(const mfem::Operator& op) override {
        d.reset(dynamic_cast<const mfem::SparseMatrix&>(op).ToDenseMatrix());
        inv = std::make_unique<mfem::DenseMatrixInverse>(*d);
        height = width = d->Height();
      }
      void Mult(const mfem::Vector& b, mfem::Vector& x) const override { inv->Mult(b, x); }
    } lin;
    auto newton = std::make_shared<mimi::solvers::LineSearchNewton>();
    newton->nl_oper_ = nl_oper.get();
    newton->iterative_mode = false;
    newton->SetOperator(*nl_oper);
    newton->SetSolver(lin);
    newton->SetPrintLevel(-1);
    newton->SetRelTol(1e-10);   // nl.configure_newton("nonlinear_solid", 1e-10, 1e-8, 100, False)
    newton->SetAbsTol(1e-8);
    newton->SetMaxIter(100);
    nl_oper->SetNewtonSolver(newton);
    mimi::solvers::GeneralizedAlpha2 ode(*nl_oper, 0.25);
    nl_oper->Setup();
    ode.SetupDirichletDofs(nl_oper->dirichlet_dofs_);
    mfem::Vector x(NV), v(NV);
    x = 0.0;
    v = 0.0;
    double t = 0.0;
    std::vector<double> hist, chist;
    std::vector<int> its;
    for (int i = 0; i < nsteps; ++i) {
      // move(): curv.cps[:] -= shift; scene.plant_kd_tree(resolution, nthreads)   (nl_contact.py:58-64)
      const int n = (int)d.cps.size() / dim;
      for (int a = 0; a < n; ++a)
        for (int c = 0; c < dim; ++c) d.cps[(size_t)a * dim + c] -= shift[c];
      scene->Clear();
      auto spline = std::make_shared<splinepy::py::PySpline>(d.Core());
      scene->AddSpline(spline);
      scene->PlantKdTree(d.resolution, nthreads);
      ode.StepTime2(x, v, t, dt);
      hist.insert(hist.end(), x.GetData(), x.GetData() + NV);
      its.push_back(newton->GetNumIterations());
      const double gn = cinteg->GapNorm(x, nthreads);
      chist.insert(chist.end(), {cinteg->last_area_, cinteg->last_pressure_, cinteg->last_force_[0], cinteg->last_force_[1],
                                 dim > 2 ? cinteg->last_force_[2] : 0.0, gn});
    }
    dump(out, "x_hist.f64", hist.data(), hist.size());
    dump(out, "newton_its.i32", its.data(), its.size());
    dump(out, "contact_hist.f64", chist.data(), chist.size());
    return 0;
  }

  if (mode == "sweeps") {
    // The O(nnz) sweeps around the integrators in one Newton iteration (operators/nonlinear_solid.cpp:240-283,
    // forms/nonlinear.hpp:53-116): y = M a; J = M; J += fac0 K(x + fac0 a) (+ r); r[ess] = 0; EliminateRowCol(ess);
    // J += fac1 S; y -= rhs.   mimi_ref sweeps <mesh> <elevate> <refine> <material> <nthreads> <fac0> <fac1> <outdir>
    const double fac0 = std::atof(argv[7]), fac1 = std::atof(argv[8]);
    const std::string out = argv[9];
    Setup s = build(mesh_file, elevate, refine, material, nthreads, true);
    auto& pre = *s.pre;
    const int NV = pre.n_v_dofs_, dim = pre.dim_;
    mfem::SparseMatrix& M = s.mass->SpMat();
    const int nnz = M.NumNonZeroElems();
    dump(out, "I.i32", M.GetI(), (size_t)NV + 1);
    dump(out, "J.i32", M.GetJ(), (size_t)nnz);
    dump(out, "mass.f64", M.GetData(), (size_t)nnz);
    // a second matrix on the same pattern standing in for the viscosity form: S = 0.37 M (+ index-dependent wobble)
    mfem::SparseMatrix S(M);
    for (int i = 0; i < nnz; ++i) S.GetData()[i] = M.GetData()[i] * (0.37 + 1.0e-3 * (i % 7));
    dump(out, "visc.f64", S.GetData(), (size_t)nnz);
    mfem::Vector x(NV), v(NV), a(NV), rhs(NV), tx(NV), tv(NV), y(NV);
    for (int i = 0; i < NV; ++i) {
      x[i] = synth_u(12345, (uint64_t)i);
      v[i] = 10.0 * synth_u(777, (uint64_t)i);
      a[i] = 100.0 * synth_u(4242, (uint64_t)i);
      rhs[i] = synth_u(99, (uint64_t)i);
    }
    dump(out, "x.f64", x.GetData(), (size_t)NV); dump(out, "v.f64", v.GetData(), (size_t)NV);
    dump(out, "a.f64", a.GetData(), (size_t)NV); dump(out, "rhs.f64", rhs.GetData(), (size_t)NV);
    // essential dofs: every vdof of the nodes on x = 0 (clamped face)
    mfem::GridFunction& nodes = *s.mesh->GetNodes();
    std::vector<int> ess;
    for (int n = 0; n < NV / dim; ++n)
      if (nodes[dim * n] == 0.0)
        for (int c = 0; c < dim; ++c) ess.push_back(dim * n + c);
    dump(out, "ess.i32", ess.data(), ess.size());
    // NonlinearSolid::ResidualAndGrad (operators/nonlinear_solid.cpp:240-283)
    add(x, fac0, a, tx);
    M.Mult(a, y);
    add(v, fac1, a, tv);
    S.AddMult(tv, y);
    mfem::SparseMatrix Jm(M);
    s.integ->AddDomainResidualAndGrad(tx, fac0, y, Jm);
    for (const int t : ess) { y[t] = 0.0; Jm.EliminateRowCol(t); }   // forms/nonlinear.hpp:111-115
    Jm.Add(fac1, S);
    y.Add(-1.0, rhs);
    for (const int t : ess) y[t] = 0.0;
    dump(out, "y.f64", y.GetData(), (size_t)NV);
    dump(out, "Jac.f64", Jm.GetData(), (size_t)nnz);
    dump(out, "tx.f64", tx.GetData(), (size_t)NV);
    std::ofstream m(out + "/meta.json");
    m << "{\"NV\": " << NV << ", \"nnz\": " << nnz << ", \"n_ess\": " << ess.size() << ", \"fac0\": " << fac0
      << ", \"fac1\": " << fac1 << "}\n";
    return 0;
  }
  if (mode == "time") {
    const int reps = std::atoi(argv[7]);
    const int warm = argc > 8 ? std::atoi(argv[8]) : 0;
    Setup s = build(mesh_file, elevate, refine, material, nthreads, false);
    auto& pre = *s.pre;
    const int NV = pre.n_v_dofs_, NE = pre.n_elements_;
    const int nq = pre.GetElementQuadData("domain")[0].n_quad;
    mfem::Vector u(NV), r(NV);
    fill_u(u, "-");
    mfem::SparseMatrix A(*pre.sparsity_pattern_);
    double best_rg = 1e300, best_r = 1e300, sum_rg = 0;
    int omp_team = 0;  // threads an OpenMP parallel region really gets here (what NThreadExe runs on)
#pragma omp parallel
    {
#pragma omp single
      omp_team = omp_get_num_threads();
    }
    for (int i = 0; i < warm; ++i) {
      r = 0.0; A = 0.0;
      s.integ->AddDomainResidualAndGrad(u, 1.0, r, A);
    }
    for (int i = 0; i < reps; ++i) {
      r = 0.0; A = 0.0;
      auto t0 = std::chrono::steady_clock::now();
      s.integ->AddDomainResidualAndGrad(u, 1.0, r, A);
      double t = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      best_rg = std::min(best_rg, t); sum_rg += t;
      r = 0.0;
      t0 = std::chrono::steady_clock::now();
      s.integ->AddDomainResidual(u, r);
      t = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      best_r = std::min(best_r, t);
    }
    std::fprintf(stderr,
                 "{\"NE\": %d, \"nq\": %d, \"NV\": %d, \"nnz\": %d, \"nthreads\": %d, \"reps\": %d, "
                 "\"t_res_grad_best\": %.6f, \"t_res_grad_mean\": %.6f, \"t_res_best\": %.6f, \"t_precompute\": %.3f, "
                 "\"omp_max_threads\": %d, \"omp_team\": %d}\n",
                 NE, nq, NV, A.NumNonZeroElems(), nthreads, reps, best_rg, sum_rg / reps, best_r, s.t_precompute,
                 omp_get_max_threads(), omp_team);
    return 0;
  }
  std::fprintf(stderr, "unknown mode\n");
  return 2;
}
