"""CPU: the plain-C oracle against the golden vectors dumped from the unmodified reference (bit-exact)."""
import os
import numpy as np
import pytest

from oracle import oracle_py as O
from tests.cases import GOLDEN_CASES, MATS, load_golden


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_pattern_and_index_map(name):
    D = load_golden(name)
    I, J = O.build_pattern(D["tdof"], D["v_dofs"], D["NV"])
    assert np.array_equal(I, D["I"]) and np.array_equal(J, D["J"])
    assert np.array_equal(O.build_A_ids(D["tdof"], D["v_dofs"], I, J), D["A_ids"])


@pytest.mark.parametrize("material", ["neohookean", "stvk"])
@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_residual_and_tangent_bit_exact(name, material):
    D = load_golden(name)
    m = O.mat(MATS[material])
    assert [m.lam, m.mu] == list(D["lambda_mu"])
    r = O.assemble_residual(D["dim"], D["nd"], D["nq"], m, D["v_dofs"], D["dN_dX"], D["w"], D["detJ"], D["u"])
    assert np.array_equal(r, D["r_" + material])
    r2, A = O.assemble_residual_and_grad(D["dim"], D["nd"], D["nq"], m, D["v_dofs"], D["A_ids"], D["dN_dX"], D["w"],
                                         D["detJ"], D["u"], D["gf"], D["mass"])
    assert np.array_equal(r2, D["r2_" + material])
    assert np.array_equal(A, D["A_" + material])


@pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref not built")
def test_oracle_vs_live_reference_4096_free():
    """A mesh no fixture covers (n=4 via DegreeElevate+2 refinements), straight from the compiled reference."""
    from tests.cases import ref_case
    D = ref_case("refine", 2, 4, "neohookean", gf=0.7)
    m = O.mat(0)
    I, J = O.build_pattern(D["tdof"], D["v_dofs"], D["NV"])
    assert np.array_equal(I, D["I"]) and np.array_equal(J, D["J"])
    aid = O.build_A_ids(D["tdof"], D["v_dofs"], I, J)
    r2, A = O.assemble_residual_and_grad(3, 27, 64, m, D["v_dofs"], aid, D["dN_dX"], D["w"], D["detJ"], D["u"], 0.7,
                                         D["mass"])
    assert np.array_equal(r2, D["r2"]) and np.array_equal(A, D["A"])


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_j2linear_with_history_bit_exact(name):
    """J2Linear (materials.hpp:138-249): virgin-state residual/tangent, the history variables committed by
    DomainPostTimeAdvance (nonlinear_solid.cpp:179-199), and the assembly at 1.5 u on that history."""
    D = load_golden(name)
    S = np.zeros((D["NE"], D["nq"], O.J2LINEAR_NS))
    m = O.mat_j2linear(S)
    args = (D["dim"], D["nd"], D["nq"], m, D["v_dofs"])
    r = O.assemble_residual(*args, D["dN_dX"], D["w"], D["detJ"], D["u"])
    assert np.array_equal(r, D["r_j2linear"])
    r2, A = O.assemble_residual_and_grad(*args, D["A_ids"], D["dN_dX"], D["w"], D["detJ"], D["u"], D["gf"], D["mass"])
    assert np.array_equal(r2, D["r2_j2linear"]) and np.array_equal(A, D["A_j2linear"])
    assert not S.any()                                   # evaluation never touches the history
    O.post_time_advance(D["dim"], D["nd"], D["nq"], m, D["v_dofs"], D["dN_dX"], D["u"])
    assert np.array_equal(S, D["state_j2linear"])
    y = (S[:, :, 18] > 0).mean()
    assert 0.05 < y < 0.999, y                           # both branches of the yield check are exercised
    rb, Ab = O.assemble_residual_and_grad(*args, D["A_ids"], D["dN_dX"], D["w"], D["detJ"], 1.5 * D["u"], D["gf"], D["mass"])
    assert np.array_equal(rb, D["rb_j2linear"]) and np.array_equal(Ab, D["Ab_j2linear"])


from tests.cases import J2_CASES, J2_LAWS, load_j2_golden  # noqa: E402


@pytest.mark.parametrize("law", J2_LAWS)
@pytest.mark.parametrize("name", J2_CASES)
def test_j2_hardening_laws_bit_exact(name, law):
    """J2 (materials.hpp:259-404) + the hardening laws of material_hardening.hpp + ScalarSolve (newton.hpp:53-169):
    the oracle runs on the same libm as the reference, so it is bit-exact."""
    D = load_j2_golden(name)
    S = O.j2_initial_state(D["NE"], D["nq"])
    m = O.mat_j2(S, law)
    args = (D["dim"], D["nd"], D["nq"], m, D["v_dofs"])
    assert np.array_equal(O.assemble_residual(*args, D["dN_dX"], D["w"], D["detJ"], D["u"]), D["r_" + law])
    r2, A = O.assemble_residual_and_grad(*args, D["A_ids"], D["dN_dX"], D["w"], D["detJ"], D["u"], D["gf"], D["mass"])
    assert np.array_equal(r2, D["r2_" + law]) and np.array_equal(A, D["A_" + law])
    O.post_time_advance(D["dim"], D["nd"], D["nq"], m, D["v_dofs"], D["dN_dX"], D["u"])
    assert np.array_equal(S, D["state_" + law])
    assert 0.05 < (S[:, :, 9] > 0).mean() < 0.999
    rb, Ab = O.assemble_residual_and_grad(*args, D["A_ids"], D["dN_dX"], D["w"], D["detJ"], 1.5 * D["u"], D["gf"], D["mass"])
    assert np.array_equal(rb, D["rb_" + law]) and np.array_equal(Ab, D["Ab_" + law])
    assert m._keep[1].value == 0


from tests.cases import SIMO_LAWS, LOG_LAWS  # noqa: E402


@pytest.mark.parametrize("law", SIMO_LAWS)
@pytest.mark.parametrize("name", J2_CASES)
def test_j2simo_bit_exact(name, law):
    """J2Simo (finite strain, materials.hpp:406-579): be_old / F_old history, cbrt, s_effective residual."""
    D = load_j2_golden(name)
    S = O.j2simo_initial_state(D["NE"], D["nq"], D["dim"])
    m = O.mat_j2(S, law, simo=True)
    args = (D["dim"], D["nd"], D["nq"], m, D["v_dofs"])
    k = "_simo_" + law
    assert np.array_equal(O.assemble_residual(*args, D["dN_dX"], D["w"], D["detJ"], D["u"]), D["r" + k])
    r2, A = O.assemble_residual_and_grad(*args, D["A_ids"], D["dN_dX"], D["w"], D["detJ"], D["u"], D["gf"], D["mass"])
    assert np.array_equal(r2, D["r2" + k]) and np.array_equal(A, D["A" + k])
    O.post_time_advance(D["dim"], D["nd"], D["nq"], m, D["v_dofs"], D["dN_dX"], D["u"])
    assert np.array_equal(S, D["state" + k])
    rb, Ab = O.assemble_residual_and_grad(*args, D["A_ids"], D["dN_dX"], D["w"], D["detJ"], 1.5 * D["u"], D["gf"], D["mass"])
    assert np.array_equal(rb, D["rb" + k]) and np.array_equal(Ab, D["Ab" + k])


@pytest.mark.parametrize("law", LOG_LAWS)
@pytest.mark.parametrize("name", J2_CASES)
def test_j2log_bit_exact(name, law):
    """J2Log (finite strain, logarithmic, materials.hpp:580-756): Fp_inv history; the log / exp maps go through the
    restated mfem CalcEigenvalues (oracle/mimi_oracle_eig.c), which these fixtures pin to the bit."""
    D = load_j2_golden(name)
    S = O.j2log_initial_state(D["NE"], D["nq"], D["dim"])
    m = O.mat_j2(S, law, log=True)
    args = (D["dim"], D["nd"], D["nq"], m, D["v_dofs"])
    k = "_log_" + law
    assert np.array_equal(O.assemble_residual(*args, D["dN_dX"], D["w"], D["detJ"], D["u"]), D["r" + k])
    r2, A = O.assemble_residual_and_grad(*args, D["A_ids"], D["dN_dX"], D["w"], D["detJ"], D["u"], D["gf"], D["mass"])
    assert np.array_equal(r2, D["r2" + k]) and np.array_equal(A, D["A" + k])
    O.post_time_advance(D["dim"], D["nd"], D["nq"], m, D["v_dofs"], D["dN_dX"], D["u"])
    assert np.array_equal(S, D["state" + k])
    assert np.any(S[:, :, 9] > 0) and not np.all(S[:, :, 9] > 0)   # both the elastic and the plastic branch ran
    rb, Ab = O.assemble_residual_and_grad(*args, D["A_ids"], D["dN_dX"], D["w"], D["detJ"], 1.5 * D["u"], D["gf"], D["mass"])
    assert np.array_equal(rb, D["rb" + k]) and np.array_equal(Ab, D["Ab" + k])


@pytest.mark.parametrize("dim", [2, 3])
def test_eig_sym_properties(dim):
    """The restated symmetric eigen-solver: ascending eigenvalues, orthonormal vectors, A V = V Lambda, vs numpy eigh;
    degenerate inputs (multiple of I, zero, a repeated eigenvalue) take the reference's special branches."""
    rng = np.random.default_rng(5)
    mats = []
    for _ in range(200):
        B = rng.standard_normal((dim, dim))
        mats.append(B + B.T)
    for _ in range(50):      # near the identity, as C_e = F_e^T F_e is
        B = 1e-3 * rng.standard_normal((dim, dim))
        mats.append(np.eye(dim) + B + B.T)
    mats += [np.eye(dim) * 2.5, np.zeros((dim, dim)), np.diag([1.0, 1.0, 3.0][:dim]), 1e300 * np.eye(dim), np.diag([3.0, 1.0, 2.0][:dim])]
    for A in mats:
        lam, V = O.eig_sym(A)
        assert np.all(np.diff(lam) >= 0)
        scale = max(np.abs(A).max(), 1e-300)
        assert np.allclose(lam / scale, np.linalg.eigvalsh(A / scale), rtol=0, atol=1e-14)
        assert np.allclose(V @ V.T, np.eye(dim), atol=1e-14)
        assert np.allclose((A / scale) @ V.T, V.T * (lam / scale), atol=2e-14)


@pytest.mark.parametrize("dim", [2, 3])
def test_eig_sym_bit_exact_vs_reference(dim):
    """oracle/mimi_oracle_eig.c against mfem::DenseMatrix::CalcEigenvalues of the unmodified reference (fixture
    eig_sym.npz from harness mode 'eig'): generic, near-identity, tiny deviatoric, repeated-eigenvalue and degenerate
    matrices -- every pivoting branch -- to the bit."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "eig_sym.npz"))
    A, lam_ref, vec_ref = z["A%d" % dim], z["lambda%d" % dim], z["vec%d" % dim]
    for i in range(A.shape[0]):
        lam, V = O.eig_sym(A[i])
        assert np.array_equal(lam, lam_ref[i]), i
        assert np.array_equal(V, vec_ref[i]), i
