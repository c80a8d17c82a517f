"""GPU: MB2_TANGENT_ANALYTIC, the non-parity fast tangent mode (mimi_b200/csrc/mb2_fast.cu), through the C ABI.
Bars: (a) equal to the numpy statement of the same closed form to rounding; (b) the reference's
forward-difference Jacobian reproduced at ITS noise level (tests/test_fast_reference.py measures that floor on the CPU);
(c) the reference's own np.allclose assertion on its committed beam histories with Newton iteration counts never above the reference's;
(d) the two contraction engines for p = 2 hexahedra (FP64 tensor cores / CUDA cores) agree; (e) parity mode untouched."""
import numpy as np
import pytest

from mimi_b200 import capi
from oracle import oracle_py as O
from tests import analytic_ref as AR
from tests.cases import MATS, load_golden, setup_ctx

pytestmark = pytest.mark.gpu

ANALYTIC = 1


def fast_ctx(ctx, D, material, gemm=0):
    setup_ctx(ctx, D, material)
    ctx.set_option("tangent_mode", ANALYTIC)
    ctx.set_option("fast_gemm", gemm)
    return ctx


@pytest.mark.parametrize("material", ["neohookean", "stvk"])
@pytest.mark.parametrize("name", ["cube_p2_n2", "cube_p2_n3_patch", "cube_p1_n2", "cube_p3_n2", "quad_p2_n4", "balken_p3_n2",
                                  "beam2d_p3_n2"])
def test_fast_mode_equals_the_closed_form_and_tracks_the_reference(ctx_factory, name, material):
    D = load_golden(name)
    lam, mu = (float(x) for x in D["lambda_mu"])
    ctx = fast_ctx(ctx_factory(), D, material)
    r = np.zeros(D["NV"])
    A = D["mass"].copy()
    ctx.add_domain_residual_and_grad(D["u"], D["gf"], r, A)
    r_n, A_n = AR.assemble(material, lam, mu, D, D["u"], D["gf"], D["mass"])
    K, K_n, K_ref = A - D["mass"], A_n - D["mass"], D["A_" + material] - D["mass"]
    # (a) the same closed form: rounding only (FMA contraction, summation order)
    assert np.abs(r - r_n).max() <= 1e-12 * np.abs(r_n).max()
    assert np.abs(K - K_n).max() <= 1e-11 * np.abs(K_n).max()
    # residual vs the reference's: the same quadrature
    assert np.abs(r - D["r2_" + material]).max() <= 1e-12 * np.abs(D["r2_" + material]).max()
    # (b) vs the reference's forward-difference Jacobian: its own noise floor (see tests/test_fast_reference.py)
    fro = np.linalg.norm(K - K_ref) / np.linalg.norm(K_ref)
    print("\n[%s %s] fast mode vs reference FD Jacobian: Frobenius-relative %.2e" % (name, material, fro))
    assert fro <= 1e-3
    # adds-into semantics and the residual-only entry are those of the parity mode
    r2 = r.copy()
    ctx.add_domain_residual(D["u"], r2)
    assert np.array_equal(r2 - r, D["r_" + material]) or np.abs((r2 - r) - D["r_" + material]).max() <= 1e-15 * np.abs(r).max()


@pytest.mark.parametrize("material", ["neohookean", "stvk"])
def test_tensor_core_and_cuda_core_contractions_agree(ctx_factory, material):
    D = load_golden("cube_p2_n3_patch")
    out = []
    for gemm in (0, 1, 2):
        ctx = fast_ctx(ctx_factory(), D, material, gemm)
        r, A = np.zeros(D["NV"]), np.zeros(D["nnz"])
        ctx.add_domain_residual_and_grad(D["u"], 1.0, r, A)
        out.append((r, A))
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][0], out[2][0])
    assert np.abs(out[0][1] - out[1][1]).max() <= 1e-13 * np.abs(out[1][1]).max()
    assert np.array_equal(out[2][1], out[1][1])           # T-buffer tensor-core form == DFMA tiles, bit for bit
    # and the generic kernel (elem_variant 1) as a third opinion
    ctx = fast_ctx(ctx_factory(), D, material)
    ctx.set_option("elem_variant", 1)
    r, A = np.zeros(D["NV"]), np.zeros(D["nnz"])
    ctx.add_domain_residual_and_grad(D["u"], 1.0, r, A)
    assert np.abs(A - out[0][1]).max() <= 1e-12 * np.abs(A).max()


def test_fast_mode_is_refused_for_materials_without_a_closed_form(ctx_factory):
    D = load_golden("cube_p2_n2")
    ctx = setup_ctx(ctx_factory(), D, "j2linear")
    ctx.set_option("tangent_mode", ANALYTIC)
    with pytest.raises(capi.Mb2Error, match="closed-form"):
        ctx.add_domain_residual_and_grad(D["u"], 1.0, np.zeros(D["NV"]), np.zeros(D["nnz"]))
    ctx.set_option("tangent_mode", 0)                       # parity mode still there, still bit-exact
    r, A = np.zeros(D["NV"]), D["mass"].copy()
    ctx.add_domain_residual_and_grad(D["u"], D["gf"], r, A)
    assert np.array_equal(A, D["A_j2linear"])
    with pytest.raises(capi.Mb2Error):
        ctx.set_option("tangent_mode", 7)


def test_dmma_cannot_serve_the_parity_mode(ctx_factory):
    """Why the FP64 tensor cores are confined to the fast mode. Measured on the B200: mma.sync.m8n8k4.f64 is a chain of
    FUSED multiply-adds over k in ascending order -- the tensor-core and the CUDA-core (DFMA) contraction give the same bits.
    The reference adds separately rounded products (densemat.cpp:3002-3015: `R += dN * bjk`, built without FMA), and the
    parity contract is 1e-12 on a difference quotient that amplifies every last bit by 1/h ~ 1e11:
    a fused accumulation cannot reproduce it. On one element fused and unfused evaluation of the SAME closed form already
    differ in the last bits."""
    D = load_golden("cube_p2_n2")
    lam, mu = (float(x) for x in D["lambda_mu"])
    A = []
    for gemm in (2, 1, 0):          # 2 = tensor cores fed from T buffers, 1 = DFMA tiles on the same buffers, 0 = default
        ctx = fast_ctx(ctx_factory(), D, "neohookean", gemm)
        a = np.zeros(D["nnz"])
        ctx.add_domain_residual_and_grad(D["u"], 1.0, np.zeros(D["NV"]), a)
        A.append(a)
    assert np.array_equal(A[0], A[1])                     # DMMA == chained DFMA (same summation order), bit for bit
    # the default form sums the same products in another order ((n, q) instead of (q, n)): rounding only
    assert np.abs(A[2] - A[0]).max() <= 1e-13 * np.abs(A[0]).max()
    _, A_n = AR.assemble("neohookean", lam, mu, D, D["u"], 1.0, np.zeros(D["nnz"]))      # unfused float64 (numpy)
    d = np.abs(A[0] - A_n).max() / np.abs(A_n).max()
    print("\ntensor-core == CUDA-core DFMA bit for bit; fused vs unfused closed form: max relative difference %.2e, "
          "%d of %d entries differ in their bits" % (d, int((A[0] != A_n).sum()), A_n.size))
    assert 0 < d <= 1e-12


@pytest.mark.parametrize("gemm", [0])
def test_newton_iteration_counts_never_exceed_the_reference_s(ctx_factory, gemm):
    """(c) The reference's own test problem (2D p = 3 clamped beam, 10 generalized-alpha steps, tests/data/ref/neohook_h1_p2)
    with the fast tangent: np.allclose against the committed vectors as the reference's test asserts; per-step Newton
    iteration counts never above the parity path's (= the CPU reference's), same converged trajectory."""
    from tests.test_reference_histories import DeviceBackend, OracleBackend, SETUP, histories
    from tests.beam_driver import Beam
    D = load_golden("balken_p3_n2")
    force, dt = SETUP["neohook"]
    ctx = ctx_factory()
    be = DeviceBackend(D, "neohook", ctx)
    ctx.set_option("tangent_mode", ANALYTIC)
    beam = Beam(D, be, {1: force}, dt)
    ref = histories()["neohook"]
    for i in range(10):
        u = beam.step()
        assert np.allclose(u, ref[i]), i
    ob = Beam(D, OracleBackend(D, "neohook"), {1: force}, dt)
    for i in range(10):
        ob.step()
    z = np.load(__import__("os").path.join(__import__("os").path.dirname(__file__), "golden", "balken_histories.npz"))
    print("\nfast tangent: Newton iterations %s; parity / oracle %s; compiled CPU reference %s; max |x - x_parity| = %.2e" % (
        beam.newton_iterations, ob.newton_iterations, [int(v) for v in z["ref_its_neohook"]], float(np.abs(beam.x - ob.x).max())))
    # The reference's forward-difference Jacobian carries 1e-4 of noise (tests/test_fast_reference.py); Newton with it
    # needs 5 / 4 iterations to reach 1e-12 relative on the first steps where the closed-form tangent needs 2: the fast mode
    # never takes MORE iterations than the reference, and lands on the same trajectory.
    assert ob.newton_iterations == [int(v) for v in z["ref_its_neohook"]]
    assert all(a <= b for a, b in zip(beam.newton_iterations, ob.newton_iterations))
    assert np.allclose(beam.x, ob.x, rtol=1e-8, atol=1e-11)


def test_3d_newton_solve_needs_no_more_iterations_with_the_fast_tangent(ctx_factory):
    """A 3D p = 2 static Newton solve (clamped face x = 0, body force) on the 27-element patch with the parity and the fast
    tangent: the fast tangent needs no more iterations (measured 4 vs 5), same solution."""
    D = load_golden("cube_p2_n3_patch")
    NV, dim = D["NV"], 3
    xr = D["x_ref"].reshape(-1, dim)
    ess = np.sort(np.concatenate([dim * np.nonzero(xr[:, 0] == 0.0)[0] + c for c in range(dim)]))
    free = np.setdiff1d(np.arange(NV), ess)
    I, J = D["I"], D["J"]
    rows = np.repeat(np.arange(NV), np.diff(I))
    f = np.zeros(NV)
    f[1::3] = -2.0 * np.bincount(D["v_dofs"].reshape(D["NE"], 3, D["nd"])[:, 1].ravel(), minlength=NV)[1::3] / D["nq"]
    res = []
    for mode in (0, ANALYTIC):
        ctx = setup_ctx(ctx_factory(), D, "neohookean")
        ctx.set_option("tangent_mode", mode)
        u = np.zeros(NV)
        its = 0
        for its in range(1, 30):
            r, A = np.zeros(NV), np.zeros(D["nnz"])
            ctx.add_domain_residual_and_grad(u, 1.0, r, A)
            g = (r - f)[free]
            if np.linalg.norm(g) <= 1e-9 * max(np.linalg.norm(f[free]), 1.0):
                break
            Kd = np.zeros((NV, NV))
            Kd[rows, J] = A
            u[free] -= np.linalg.solve(Kd[np.ix_(free, free)], g)
        res.append((its, u.copy()))
    print("\n3D Newton: parity tangent %d iterations, fast tangent %d iterations, |du| = %.2e" % (
        res[0][0], res[1][0], float(np.abs(res[0][1] - res[1][1]).max())))
    assert res[1][0] <= res[0][0] and 3 <= res[0][0] < 29
    assert np.abs(res[0][1] - res[1][1]).max() <= 1e-9 * np.abs(res[0][1]).max()
