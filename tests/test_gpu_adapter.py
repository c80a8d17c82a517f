"""Drop-in proof (GPU): the reference's OWN call chain -- mfem::Mesh / FiniteElementSpace, PrecomputedData, the material
objects, the mass BilinearForm, AddDomainResidual / AddDomainResidualAndGrad / DomainPostTimeAdvance called through the
NonlinearBase interface -- with the domain integrator replaced by adapter/b200_nonlinear_solid.hpp, the NonlinearBase
subclass INTEGRATION.md describes, which forwards to the C ABI of libmimi_b200.so.

oracle/_ref/mimi_ref_b200 is that program: oracle/ref_harness.cpp compiled with -DMIMI_REF_USE_B200 against the
unmodified reference sources (oracle/Makefile.ref, built where /root/reference exists; the binary travels to the GPU
box). Its dumps must equal the golden vectors the reference's CPU integrator produced for the same inputs."""
import os
import tempfile

import numpy as np
import pytest

from mimi_b200 import meshgen
from oracle import oracle_py as O
from tests.cases import load_golden, load_j2_golden

pytestmark = pytest.mark.gpu

# (fixture, mesh writer, elevate, refine): the recipes of oracle/make_golden.py CASES
RECIPES = {
    "cube_p2_n2": (lambda f: meshgen.write_box_p1(f, (1.0, 1.0, 1.0)), 1, 1),
    "beam2d_p3_n2": (lambda f: meshgen.write_box_p1(f, (5.0, 1.0)), 2, 1),
    "quad_p2_n4": (lambda f: meshgen.write_box_p1(f, (2.0, 1.0)), 1, 2),
    "cube_p3_n2": (lambda f: meshgen.write_box_p1(f, (1.0, 1.0, 1.0)), 2, 1),
    # rational / curved / non-affine geometry (oracle/make_golden.py CURVED_CASES)
    "annulus_p2_n2": (lambda f: meshgen.write_quarter_annulus(f), 0, 1),
    "cylinder_p2_n2": (lambda f: meshgen.write_quarter_annulus(f, height=1.0), 0, 1),
    "square_nurbs_p2_n4": (lambda f: meshgen.write_cell_p1(f, meshgen.SQUARE_NURBS_VERTS), 1, 2),
}


# multi-patch beams (oracle/make_golden.py MULTIPATCH_CASES): only the integrator chain (the device forms are per patch)
MP_RECIPES = {
    "beam4_p3": (lambda f: meshgen.write_beam_patches(f, 3, (2, 1, 1), 4, (1.0, 0.5, 0.5)), 0, 0),
    "beam3_p2": (lambda f: meshgen.write_beam_patches(f, 2, (2, 2, 2), 3, (1.0, 1.0, 1.0)), 0, 0),
}


def run_adapter_env(name, material, gf, env):
    if not O.have_ref_b200():
        pytest.skip("oracle/_ref/mimi_ref_b200 not built (make -f oracle/Makefile.ref b200, needs /root/reference)")
    writer, elevate, refine = RECIPES[name] if name in RECIPES else MP_RECIPES[name]
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        writer(mesh)
        return O.run_ref_dump(mesh, elevate, refine, material, os.path.join(tmp, "d"), gf=gf, binary=O.REF_B200_BIN, env=env)


def run_adapter(name, material, gf):
    if not O.have_ref_b200():
        pytest.skip("oracle/_ref/mimi_ref_b200 not built (make -f oracle/Makefile.ref b200, needs /root/reference)")
    writer, elevate, refine = RECIPES[name] if name in RECIPES else MP_RECIPES[name]
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        writer(mesh)
        return O.run_ref_dump(mesh, elevate, refine, material, os.path.join(tmp, "d"), gf=gf, binary=O.REF_B200_BIN)


@pytest.mark.parametrize("material", ["neohookean", "stvk", "j2linear"])
@pytest.mark.parametrize("name", list(RECIPES) + list(MP_RECIPES))
def test_reference_chain_through_the_adapter_bit_exact(name, material):
    G = load_golden(name)
    D = run_adapter(name, material, G["gf"])
    assert np.array_equal(D["u"], G["u"]) and np.array_equal(D["mass"], G["mass"])     # same inputs
    assert np.array_equal(D["r"], G["r_" + material])
    assert np.array_equal(D["r2"], G["r2_" + material])
    assert np.array_equal(D["A"], G["A_" + material])          # mass + gf * K, added into the reference's SparseMatrix
    if material == "j2linear":
        assert np.array_equal(D["state"].reshape(G["state_j2linear"].shape), G["state_j2linear"])
        assert np.array_equal(D["r_b"], G["rb_j2linear"]) and np.array_equal(D["A_b"], G["Ab_j2linear"])


@pytest.mark.parametrize("name", list(RECIPES))
def test_setup_forms_through_the_adapter_bit_exact(name):
    """B200NonlinearSolid::AssembleMass / AssembleViscosity / AssembleBodyForce / AddTraction (patch description taken from the
    reference's own NURBSExtension) against MFEM's VectorMassIntegrator / VectorDiffusionIntegrator /
    VectorDomainLFIntegrator in the same process, on the same mesh."""
    G = load_golden(name)
    D = run_adapter(name, "neohookean", G["gf"])
    assert np.array_equal(D["mass"], G["mass"])
    assert np.array_equal(D["mass_b200"], D["mass"])
    assert np.array_equal(D["viscform_b200"], D["viscform"]) and np.abs(D["viscform"]).max() > 0
    assert np.array_equal(D["bodyforce_b200"], D["bodyforce"]) and np.abs(D["bodyforce"]).max() > 0
    assert np.array_equal(D["rhs_b200"], D["rhs_bf_traction"])       # body force + traction = the reference's rhs form
    assert not np.array_equal(D["rhs_b200"], D["bodyforce"])


def _rel(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


@pytest.mark.parametrize("material,key,ktol", [("j2_jcthermo", "_jcthermo", 1e-4), ("j2_voce", "_voce", 1e-4),
                                               ("j2simo_voce", "_simo_voce", 2e-3), ("j2log_jcthermo", "_log_jcthermo", 1e-2)])
def test_reference_chain_through_the_adapter_j2_family(material, key, ktol):
    """The adapter's material dispatch (MaterialBase::Name + dynamic_cast on the hardening law) for the J2 family;
    tolerances as in tests/test_gpu_j2.py."""
    name = "j2_cube_p2_n2"
    G = load_j2_golden(name)
    D = run_adapter("cube_p2_n2", material, G["gf"])
    assert _rel(D["r"], G["r" + key]) <= 1e-12 and _rel(D["A"], G["A" + key]) <= ktol
    assert _rel(D["state"].reshape(G["state" + key].shape), G["state" + key]) <= 1e-12
    assert _rel(D["r_b"], G["rb" + key]) <= 1e-12 and _rel(D["A_b"], G["Ab" + key]) <= ktol


HISTORY_RUNS = {"neohook": ("neohookean", -5.0, 0.05), "j2": ("j2_balken", -3.0, 0.5),
                "j2_simo": ("j2simo_balken", -3.0, 0.5), "j2_log": ("j2log_balken", -3.0, 0.5)}


@pytest.mark.parametrize("case", list(HISTORY_RUNS))
def test_reference_time_loop_through_the_adapter(case):
    """The reference's own test (tests/test_nonlinear_solid.py:54-112 there), end to end on the reference's own classes:
    forms::Nonlinear, operators::NonlinearSolid, LineSearchNewton, GeneralizedAlpha2 (harness mode 'steps' = the Setup of
    py_nonlinear_solid.cpp:155-386) with B200NonlinearSolid registered by AddDomainIntegrator. Ten steps; np.allclose
    against the reference's committed x_<i>.txt as in its test, Newton iteration counts identical to the CPU
    reference's, and the trajectory within 1e-9 of the one the compiled CPU reference produces."""
    if not O.have_ref_b200():
        pytest.skip("oracle/_ref/mimi_ref_b200 not built (make -f oracle/Makefile.ref b200, needs /root/reference)")
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "balken_histories.npz"))
    mat, force, dt = HISTORY_RUNS[case]
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        meshgen.write_box_p1(mesh, (5.0, 1.0))        # same discretisation as the reference's balken.mesh (fixtures equal)
        x, its = O.run_ref_steps(mesh, 2, 1, mat, force, dt, 10, os.path.join(tmp, "o"), binary=O.REF_B200_BIN)
    ref_its = [int(v) for v in z["ref_its_" + case]]
    its = [int(v) for v in its]
    print("\n[%s] adapter: Newton iterations %s (CPU reference %s), max |x - x_cpu_reference| = %.2e" % (
        case, its, ref_its, float(np.abs(x - z["ref_x_" + case]).max())))
    for i in range(10):
        assert np.allclose(x[i], z[case][i]), (case, i)          # the reference's own assertion
    if case == "neohook":
        # bit-exact integrator => the whole nonlinear trajectory is bit-identical to the CPU reference's
        assert its == ref_its and np.array_equal(x, z["ref_x_" + case])
    else:
        # J2 family: pow / exp / log come from libdevice here and from glibc in the reference. How much that may move is
        # MEASURED, not argued: tests/golden/j2_libm_spread.npz (oracle/libm_spread.py) holds the same compiled reference run
        # with glibc's FMA ifunc variants and with its generic ones -- bit-equal residuals, FD tangents 3e-9 .. 3e-5 apart,
        # trajectories 1e-12 apart, equal iteration counts. The device: trajectory within 1e-11 (absolute; |x| = O(1)) of the
        # CPU reference's, every step whose Newton count differs reported, none by more than one iteration.
        S = np.load(os.path.join(os.path.dirname(__file__), "golden", "j2_libm_spread.npz"))
        diff = [(i, its[i], ref_its[i]) for i in range(10) if its[i] != ref_its[i]]
        dev = float(np.abs(x - z["ref_x_" + case]).max())
        print("[%s] steps whose Newton count differs from the CPU reference's (step, device, reference): %s; trajectory deviation %.2e;"
              " reference-vs-reference (glibc FMA variants on / off): counts %s, trajectory %.2e" % (
                  case, diff, dev, "equal" if np.array_equal(S["its_default_" + case], S["its_generic_" + case]) else "differ",
                  float(S["x_spread_" + case])))
        assert S["its_default_" + case].tolist() == ref_its
        assert dev <= 1e-10
        assert all(abs(a - b) <= 1 for _, a, b in diff) and len(diff) <= 1
        if case != "j2":
            assert not diff          # J2Simo / J2Log: identical counts (J2: one step of ten takes 4 instead of 5)


from oracle.make_golden import CONTACT_CASES as CONTACT_RECIPES  # noqa: E402  (mesh recipe, boundary, penalty, body)
from tests.cases import load_contact_golden  # noqa: E402


@pytest.mark.parametrize("name", list(CONTACT_RECIPES))
def test_reference_contact_chain_through_the_adapter(name):
    """adapter/b200_mortar_contact.hpp in the place of mimi::integrators::MortarContact, wired as
    py_nonlinear_solid.cpp:292-325 does (boundary marker, CreateBoundaryElementQuadData, PrecomputeElementQuadData,
    Prepare), sharing the domain adapter's context: AddDomainResidual then AddBoundaryResidual on the same r,
    AddBoundaryResidualAndGrad into the reference's SparseMatrix (pre-loaded with mass), GapNorm, and the recordings
    (nodal gap / area / pressure, marked vdofs, area / pressure / force). Golden = the reference's own MortarContact."""
    if not O.have_ref_b200():
        pytest.skip("oracle/_ref/mimi_ref_b200 not built (make -f oracle/Makefile.ref b200, needs /root/reference)")
    writer, elevate, refine, gf, attr, penalty, spec = CONTACT_RECIPES[name]
    G = load_contact_golden(name)
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        writer(mesh)
        D = O.run_ref_contact(mesh, elevate, refine, os.path.join(tmp, "c"), attr, penalty, spec, gf=gf, binary=O.REF_B200_BIN)
    assert np.array_equal(D["c_marked_v_dofs"], G["c_marked_v_dofs"])
    for k in ("c_r0", "c_r", "c_gap", "c_area", "c_pressure", "c_r2", "c_A"):
        assert np.array_equal(D[k], G[k]), k
    assert np.any(G["c_pressure"] != 0)
    assert np.allclose(D["c_scalars"], G["c_scalars"], rtol=1e-12, atol=0)      # whole-boundary sums: DESIGN.md section 2


from tests.cases import SPLINE_CONTACT_CASES  # noqa: E402


@pytest.mark.parametrize("name", SPLINE_CONTACT_CASES)
def test_reference_splinepy_scene_through_the_adapter(name):
    """The reference's OWN contact scene -- NearestDistanceToSplines holding a real splinepy spline, add_spline +
    plant_kd_tree as examples/nl_contact.py:31-33 -- handed to B200MortarContact: the adapter's
    B200NearestDistanceToSplines reads degrees / knots / control points / weights from the splinepy core and describes
    them to the device (MB2_RIGID_BEZIER / RATIONAL_BEZIER / BSPLINE). Golden = the unmodified MortarContact with
    splinepy's own closest-point search; tolerances as in tests/test_gpu_contact.py (projection: 1e-12; FD tangent 1e-4)."""
    if not O.have_ref_b200():
        pytest.skip("oracle/_ref/mimi_ref_b200 not built (make -f oracle/Makefile.ref b200, needs /root/reference)")
    from oracle.make_golden import _spline_bodies
    writer, elevate, refine, gf, attr, penalty, spec = _spline_bodies()[name]
    G = load_contact_golden(name)
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        writer(mesh)
        D = O.run_ref_contact(mesh, elevate, refine, os.path.join(tmp, "c"), attr, penalty, spec, gf=gf, binary=O.REF_B200_BIN)
    assert np.array_equal(D["c_marked_v_dofs"], G["c_marked_v_dofs"]) and np.array_equal(D["c_area"], G["c_area"])
    assert np.array_equal(D["c_pressure"] != 0, G["c_pressure"] != 0) and np.any(G["c_pressure"] != 0)
    for k in ("c_gap", "c_pressure", "c_r2"):
        assert _rel(D[k], G[k]) <= 1e-12, k
    assert _rel(D["c_r"] - D["c_r0"], G["c_r"] - G["c_r0"]) <= 1e-12
    assert _rel(D["c_A"] - D["mass"], G["c_A"] - G["mass"]) <= 1e-4
    assert np.allclose(D["c_scalars"], G["c_scalars"], rtol=1e-11, atol=0)


def test_nl_contact_time_loop_through_the_adapter():
    """Config 2 of BASELINE.json: examples/nl_contact.py's time loop (square-nurbs mesh, elevate 1, subdivide 3,
    NeoHookean E = 1e10 / rho = 7e4, cubic Bezier rigid curve moved 0.005 down per step with its kd-tree re-planted,
    penalty 1e11, dt = 1e-3, Newton 1e-10 / 1e-8 / 100) on the reference's own classes with B200NonlinearSolid and
    B200MortarContact registered instead of NonlinearSolid and MortarContact. Golden = the same harness mode on the CPU
    reference with real splinepy (tests/golden/contact_steps_square.npz), incl. the reference's own variants (nthreads,
    kd-tree resolution) that bound what "identical Newton iteration counts" can mean for this example."""
    if not O.have_ref_b200():
        pytest.skip("oracle/_ref/mimi_ref_b200 not built (make -f oracle/Makefile.ref b200, needs /root/reference)")
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "contact_steps_square.npz"))
    n = z["x"].shape[0]
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        meshgen.write_cell_p1(mesh, O.NL_CONTACT["verts"])
        x, its, ch = O.run_ref_steps_contact(mesh, os.path.join(tmp, "o"), n, binary=O.REF_B200_BIN)
    V = z["its_variants"]                       # the reference's own runs: nthreads 1..4, kd-tree resolutions 5000 / 20000
    lo, hi = V.min(axis=0), V.max(axis=0)
    stable = lo == hi
    print("\n[nl_contact] adapter Newton iterations %s" % its.tolist())
    for lab, iv, sp in zip(z["variant_labels"], V, z["x_rel_spread"]):
        print("             CPU reference %-16s %s  (trajectory vs nthreads=1: %.1e)" % (str(lab), iv.tolist(), float(sp)))
    print("             steps on which the reference agrees with itself: %s; adapter trajectory vs nthreads=1: %.2e" % (
        np.nonzero(stable)[0].tolist(), _rel(x, z["x"])))
    assert (z["contact_hist"][:, 1] != 0).sum() >= 8               # the body is in contact for most of the steps
    # Newton iteration counts: this example's counts are NOT a function of the problem alone -- the reference itself
    # takes 43 / 32 / 46 / 39 iterations on step 3 at nthreads = 1 / 2 / 3 / 4 (penalty 1e11: an active-set change is one
    # rounding error away). Where the reference agrees with itself the adapter must take exactly that count; elsewhere
    # it must stay inside the reference's own band widened by its width.
    assert stable.sum() >= 4 and np.array_equal(its[stable], lo[stable])
    w = hi - lo
    assert np.all(its >= lo - w) and np.all(its <= hi + w)
    # the converged trajectory (each step solved to 1e-10 relative): the reference's own spread is 3e-12 .. 2e-11
    assert _rel(x, z["x"]) <= 1e-9
    assert np.allclose(ch[:, :5], z["contact_hist"][:, :5], rtol=1e-7, atol=1e-9 * np.abs(z["contact_hist"][:, :5]).max())


def test_config1_example_as_shipped_through_the_adapter():
    """BASELINE config 1 = examples/nonlinear_solid.py as shipped (balken.mesh, elevate 1, subdivide 2, NeoHookean, body
    force -5, dt 0.05, Newton 1e-12 / 1e-8 / 10, default ode_coefficient 0.25): 20 generalized-alpha steps on the reference's
    own classes with B200NonlinearSolid as the domain integrator. Identical Newton iteration counts and a bit-identical
    trajectory against the compiled CPU reference (tests/golden/config1_steps.npz, nthreads 1 and 2 agree there)."""
    if not O.have_ref_b200():
        pytest.skip("oracle/_ref/mimi_ref_b200 not built (make -f oracle/Makefile.ref b200, needs /root/reference)")
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "config1_steps.npz"))
    n = z["x_nt1"].shape[0]
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        meshgen.write_box_p1(mesh, (5.0, 1.0))
        x, its = O.run_ref_steps(mesh, 1, 2, "neohookean", -5.0, 0.05, n, os.path.join(tmp, "o"), binary=O.REF_B200_BIN, rho_inf=0.25)
    print("\n[config 1] adapter Newton iterations %s; CPU reference %s; max |x - x_ref| = %.2e" % (
        its.tolist(), z["its_nt1"].tolist(), float(np.abs(x - z["x_nt1"]).max())))
    assert np.array_equal(z["its_nt1"], z["its_nt2"])
    assert its.tolist() == z["its_nt1"].tolist()
    assert np.array_equal(x, z["x_nt1"])          # bit-exact integrator => bit-identical nonlinear trajectory


def test_config1_with_the_direct_solver_adapter():
    """adapter/b200_gmres_solver.hpp::B200DirectSolver (mb2_newton_solve_direct_device: cuSOLVER sparse QR on the device CSR)
    in the place of the reference's default direct solver, driven by the reference's LineSearchNewton over the 20 steps of
    config 1. Golden = the CPU reference with its direct solve (UMFPACK is absent here: the harness' dense LU). Equal
    Newton iteration counts; trajectory within 1e-9 relative (two direct solvers agree to rounding, the Newton iterates
    inherit it)."""
    if not O.have_ref_b200():
        pytest.skip("oracle/_ref/mimi_ref_b200 not built (make -f oracle/Makefile.ref b200, needs /root/reference)")
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "config1_steps.npz"))
    n = z["x_nt1"].shape[0]
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        meshgen.write_box_p1(mesh, (5.0, 1.0))
        x, its = O.run_ref_steps(mesh, 1, 2, "neohookean", -5.0, 0.05, n, os.path.join(tmp, "o"), binary=O.REF_B200_BIN,
                                 rho_inf=0.25, lin="direct")
    dev = float(np.abs(x - z["x_nt1"]).max() / np.abs(z["x_nt1"]).max())
    print("\n[config 1, direct solver adapter] Newton iterations %s (reference %s), max |x - x_ref| / max |x_ref| = %.2e" % (
        its.tolist(), z["its_nt1"].tolist(), dev))
    assert its.tolist() == z["its_nt1"].tolist()
    assert dev <= 1e-9


def test_config1_with_the_iterative_solver_adapter():
    """`use_iterative_solver` (py_nonlinear_solid.cpp:330-339) through the adapter: the same 20 steps of config 1 with
    adapter/b200_gmres_solver.hpp (device GMRES(50) + Jacobi on the CSR the integrator adapter filled) driven by the
    reference's LineSearchNewton in place of mfem::GMRESSolver + mfem::DSmoother. Golden = the CPU reference with its own
    GMRES (60 Newton iterations, 37 .. 97 linear iterations each; Newton hands the solver iterative_mode = false,
    solvers/newton.cpp:74). Newton iteration counts must be equal. The first linear system of a time step is the
    reference's own up to the trajectory's 1e-11, so its GMRES count must agree within 1 (the device sums its dot products
    in a tree: iterates agree to rounding, not bitwise); every later system of a step carries the previous solve's own
    truncation error (rel 1e-8) in its right-hand side, so those counts are only comparable loosely: within 10 per solve
    (observed <= 7) and 2 % in total. Trajectory within 1e-8 relative (observed 9e-12)."""
    if not O.have_ref_b200():
        pytest.skip("oracle/_ref/mimi_ref_b200 not built (make -f oracle/Makefile.ref b200, needs /root/reference)")
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "config1_steps.npz"))
    n = z["x_gmres"].shape[0]
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        meshgen.write_box_p1(mesh, (5.0, 1.0))
        x, its, lin = O.run_ref_steps(mesh, 1, 2, "neohookean", -5.0, 0.05, n, os.path.join(tmp, "o"), binary=O.REF_B200_BIN,
                                      rho_inf=0.25, lin="gmres")
    dev = float(np.abs(x - z["x_gmres"]).max() / np.abs(z["x_gmres"]).max())
    print("\n[config 1, GMRES adapter] Newton iterations %s (reference %s)\n linear iterations %s\n reference         %s\n"
          " max |x - x_ref| / max |x_ref| = %.2e" % (its.tolist(), z["its_gmres"].tolist(), lin.tolist(), z["lin_its_gmres"].tolist(), dev))
    assert its.tolist() == z["its_gmres"].tolist()
    ref_lin = z["lin_its_gmres"]
    assert lin.size == ref_lin.size == int(its.sum())
    first = np.concatenate([[0], np.cumsum(its)[:-1]])
    assert np.abs(lin - ref_lin)[first].max() <= 1
    assert np.abs(lin - ref_lin).max() <= 10 and abs(int(lin.sum()) - int(ref_lin.sum())) <= 0.02 * ref_lin.sum()
    assert dev <= 1e-8


@pytest.mark.parametrize("name,material,ngpu", [("cube_p2_n2", "neohookean", 2), ("cube_p2_n2", "j2linear", 3), ("beam4_p3", "stvk", 4),
                                                ("quad_p2_n4", "neohookean", 2)])
def test_multi_gpu_adapter_in_one_process(name, material, ngpu):
    """adapter/b200_multigpu.hpp: a NonlinearBase subclass owning N contexts -- the partition planned by the library
    (mb2_plan_partition), element matrices exchanged by mb2_mg_group_assemble -- called through the reference's virtuals
    inside mimi_ref_b200 (contexts round-robin over the visible GPUs; on a one-GPU box they share device 0). r and the
    CSR values (adds-into, called twice) equal the reference's single-thread assembly bit for bit."""
    G = load_golden(name)
    ndev = 1
    try:
        import torch
        ndev = max(1, torch.cuda.device_count())
    except Exception:
        pass
    D = run_adapter_env(name, material, G["gf"], {"MIMI_REF_B200_GPUS": str(ngpu), "MIMI_REF_B200_NDEV": str(ndev)})
    assert np.array_equal(D["r_mg"], G["r_" + material])
    assert np.array_equal(D["r2_mg"], 2.0 * G["r2_" + material])
    K = G["A_" + material] - G["mass"]
    want = G["A_" + material].copy()
    want += (G["gf"] * 0 + 1.0) * K            # second call adds gf * K once more onto mass + gf * K
    assert np.allclose(D["A_mg"], want, rtol=1e-13, atol=1e-13 * np.abs(K).max())
    assert np.array_equal(D["A"], G["A_" + material])          # the single-context adapter in the same run
