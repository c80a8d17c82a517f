"""GPU: materials with per-quadrature-point history (J2Linear, src/mimi/materials/materials.hpp:142-249) through the
C ABI against the golden vectors dumped from the unmodified reference: virgin-state residual + FD tangent, the history
committed by mb2_post_time_advance (NonlinearSolid::DomainPostTimeAdvance, nonlinear_solid.cpp:179-199), and the
assembly at 1.5 u on that history. Bar: bit-exact FP64 (np.array_equal); north_star's tolerance is 1e-12 relative."""
import numpy as np
import pytest

from oracle import oracle_py as O
from tests.cases import GOLDEN_CASES, load_golden, setup_ctx

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_j2linear_golden(ctx_factory, name):
    D = load_golden(name)
    ctx = setup_ctx(ctx_factory(), D, "j2linear")
    assert ctx.state_size() == O.J2LINEAR_NS
    assert not ctx.get_state().any()                      # J2Linear::CreateState zero-initialises (materials.cpp:120-133)
    r = np.zeros(D["NV"])
    ctx.add_domain_residual(D["u"], r)
    assert np.array_equal(r, D["r_j2linear"])
    r2 = np.zeros(D["NV"])
    A = D["mass"].copy()
    ctx.add_domain_residual_and_grad(D["u"], D["gf"], r2, A)
    assert np.array_equal(r2, D["r2_j2linear"])
    assert np.abs(A - D["A_j2linear"]).max() <= 1e-12 * np.abs(D["A_j2linear"]).max()
    assert np.array_equal(A, D["A_j2linear"])
    assert not ctx.get_state().any()                      # evaluation never touches the history
    ctx.post_time_advance(D["u"])
    S = ctx.get_state()
    assert np.array_equal(S, D["state_j2linear"])
    rb = np.zeros(D["NV"])
    Ab = D["mass"].copy()
    ctx.add_domain_residual_and_grad(np.ascontiguousarray(1.5 * D["u"]), D["gf"], rb, Ab)
    assert np.array_equal(rb, D["rb_j2linear"]) and np.array_equal(Ab, D["Ab_j2linear"])
    # checkpoint / restore of the history, and re-initialisation
    ctx.set_state(None)
    assert not ctx.get_state().any()
    ctx.set_state(S)
    rb2 = np.zeros(D["NV"])
    ctx.add_domain_residual(np.ascontiguousarray(1.5 * D["u"]), rb2)
    assert np.array_equal(rb2, rb)


def test_j2linear_two_commits_vs_oracle(ctx_factory):
    """Two successive DomainPostTimeAdvance commits (history on history), checked against the C oracle."""
    D = load_golden("cube_p2_n2")
    ctx = setup_ctx(ctx_factory(), D, "j2linear")
    S = np.zeros((D["NE"], D["nq"], O.J2LINEAR_NS))
    m = O.mat_j2linear(S)
    for f in (1.0, 2.5):
        u = np.ascontiguousarray(f * D["u"])
        O.post_time_advance(3, 27, 64, m, D["v_dofs"], D["dN_dX"], u)
        ctx.post_time_advance(u)
        assert np.array_equal(ctx.get_state(), S)
    u = np.ascontiguousarray(3.0 * D["u"])
    r_o, A_o = O.assemble_residual_and_grad(3, 27, 64, m, D["v_dofs"], D["A_ids"], D["dN_dX"], D["w"], D["detJ"], u,
                                            1.0, D["mass"])
    r = np.zeros(D["NV"])
    A = D["mass"].copy()
    ctx.add_domain_residual_and_grad(u, 1.0, r, A)
    assert np.array_equal(r, r_o) and np.array_equal(A, A_o)


def test_stateless_post_time_advance_is_a_noop(ctx_factory):
    D = load_golden("cube_p1_n2")
    ctx = setup_ctx(ctx_factory(), D, "neohookean")
    assert ctx.state_size() == 0
    ctx.post_time_advance(D["u"])


# ---- J2 with nonlinear isotropic hardening (materials.hpp:259-404), hardening laws (material_hardening.hpp), bracketed
# Newton return mapping (newton.hpp:53-169). The laws call pow / exp / log: CUDA's libdevice (<= 1 ulp) where the
# reference calls glibc, so this family is NOT bit-exact. Stated tolerances (vs golden vectors of the unmodified
# reference):
#   residual, committed history:  1e-12 relative to the largest entry  (north_star's tolerance)
#   FD tangent:                   1e-4  relative to the largest entry  -- the return mapping stops at |f| < sigma_y *
#                                 1e-10 (materials.hpp:353-358), so every yielded stress carries solver noise of that
#                                 size and a last-bit libm difference can change the iterate it stops at; the forward
#                                 difference divides that by h = 1e-8 |u| ~ 1e-11. The reference's own K carries the
#                                 same noise (observed deviations are printed: 1e-9 .. 5e-6).
from tests.cases import J2_CASES, J2_LAWS, j2_params, load_j2_golden  # noqa: E402


def _rel(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


def _print_spread(mat):
    """The reference against itself (oracle/libm_spread.py: glibc's FMA ifunc variants of pow / exp / log on and off, 3D
    fixture): the floor for the FD tangent of this material, printed beside the device's deviation."""
    import os
    f = os.path.join(os.path.dirname(__file__), "golden", "j2_libm_spread.npz")
    if os.path.exists(f):
        z = np.load(f)
        if "spread_" + mat in z.files:
            print("    reference vs reference (libm variants), cube_p2_n2: r %.1e  A %.1e  state %.1e  rb %.1e  Ab %.1e" % tuple(z["spread_" + mat]))


@pytest.mark.parametrize("law", J2_LAWS)
@pytest.mark.parametrize("name", J2_CASES)
def test_j2_hardening_golden(ctx_factory, name, law):
    D = load_j2_golden(name)
    ctx = ctx_factory()
    ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
    ctx.upload_quad_tables(D["dN_dX"], D["w"], D["detJ"])
    ctx.build_pattern()
    ctx.set_material(3, j2_params(law))
    assert ctx.state_size() == O.J2_NS
    assert np.array_equal(ctx.get_state(), O.j2_initial_state(D["NE"], D["nq"]))   # J2::CreateState
    r = np.zeros(D["NV"])
    ctx.add_domain_residual(D["u"], r)
    assert _rel(r, D["r_" + law]) <= 1e-12
    r2 = np.zeros(D["NV"])
    A = D["mass"].copy()
    ctx.add_domain_residual_and_grad(D["u"], D["gf"], r2, A)
    assert np.array_equal(r2, r)                                   # same kernel arithmetic in both entry points
    assert _rel(A, D["A_" + law]) <= 1e-4
    ctx.post_time_advance(D["u"])
    S = ctx.get_state()
    assert _rel(S, D["state_" + law]) <= 1e-12
    assert np.array_equal(S[:, :, 9] > 0, D["state_" + law][:, :, 9] > 0)   # the same points yield
    rb = np.zeros(D["NV"])
    Ab = D["mass"].copy()
    ctx.add_domain_residual_and_grad(np.ascontiguousarray(1.5 * D["u"]), D["gf"], rb, Ab)
    assert _rel(rb, D["rb_" + law]) <= 1e-12 and _rel(Ab, D["Ab_" + law]) <= 1e-4
    print("\n[j2 %s %s] rel dev: r %.1e  A %.1e  state %.1e  rb %.1e  Ab %.1e  (bit-equal r: %s)" % (
        name, law, _rel(r, D["r_" + law]), _rel(A, D["A_" + law]), _rel(S, D["state_" + law]), _rel(rb, D["rb_" + law]),
        _rel(Ab, D["Ab_" + law]), np.array_equal(r, D["r_" + law])))
    _print_spread("j2_" + law)


def test_j2_scalar_solve_failure_is_reported(ctx_factory):
    """A root that cannot be bracketed makes the reference throw (newton.hpp:81-93); here the device flag turns into
    MB2_ERR_NUMERIC. A rate-dependent law with an absurd rate coefficient pushes residual(upper) > 0."""
    from mimi_b200 import capi
    D = load_j2_golden("j2_quad_p2_n4")
    ctx = ctx_factory()
    ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
    ctx.upload_quad_tables(D["dN_dX"], D["w"], D["detJ"])
    ctx.build_pattern()
    p = j2_params("jcrate")
    p[5 + 3] = -50.0          # C < 0: the rate term lowers the yield stress at the upper bound -> both ends positive
    ctx.set_material(3, p)
    r = np.zeros(D["NV"])
    with pytest.raises(capi.Mb2Error, match="ScalarSolve"):
        ctx.add_domain_residual(D["u"], r)
    ctx.set_material(3, j2_params("jcrate"))      # recovers
    ctx.add_domain_residual(D["u"], r)


from tests.cases import SIMO_LAWS  # noqa: E402


@pytest.mark.parametrize("law", SIMO_LAWS)
@pytest.mark.parametrize("name", J2_CASES)
def test_j2simo_golden(ctx_factory, name, law):
    """J2Simo (materials.hpp:406-579): libdevice cbrt / pow / exp / log. Residual and history 1e-12 as for J2; FD tangent
    2e-3 of the largest entry: the formulation takes dev(b_e) of b_e = I + O(strain), so a last-bit difference in
    cbrt(det f) (1e-16 of b_e) is a 1e-13 relative difference of the deviatoric stress at strains of 1e-3, which the
    forward difference divides by h = 1e-8 |u| ~ 1e-11. The reference's own K carries rounding noise of that size
    (observed deviations are printed: 2e-5 .. 4e-4)."""
    D = load_j2_golden(name)
    k = "_simo_" + law
    ctx = ctx_factory()
    ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
    ctx.upload_quad_tables(D["dN_dX"], D["w"], D["detJ"])
    ctx.build_pattern()
    ctx.set_material(4, j2_params(law))
    assert ctx.state_size() == O.J2SIMO_NS
    assert np.array_equal(ctx.get_state(), O.j2simo_initial_state(D["NE"], D["nq"], D["dim"]))
    r = np.zeros(D["NV"])
    ctx.add_domain_residual(D["u"], r)
    assert _rel(r, D["r" + k]) <= 1e-12
    r2 = np.zeros(D["NV"])
    A = D["mass"].copy()
    ctx.add_domain_residual_and_grad(D["u"], D["gf"], r2, A)
    assert np.array_equal(r2, r) and _rel(A, D["A" + k]) <= 2e-3
    ctx.post_time_advance(D["u"])
    S = ctx.get_state()
    assert _rel(S, D["state" + k]) <= 1e-12
    rb = np.zeros(D["NV"])
    Ab = D["mass"].copy()
    ctx.add_domain_residual_and_grad(np.ascontiguousarray(1.5 * D["u"]), D["gf"], rb, Ab)
    assert _rel(rb, D["rb" + k]) <= 1e-12 and _rel(Ab, D["Ab" + k]) <= 2e-3
    print("\n[j2simo %s %s] rel dev: r %.1e  A %.1e  state %.1e  rb %.1e  Ab %.1e" % (
        name, law, _rel(r, D["r" + k]), _rel(A, D["A" + k]), _rel(S, D["state" + k]), _rel(rb, D["rb" + k]),
        _rel(Ab, D["Ab" + k])))
    _print_spread("j2simo_" + law)


from tests.cases import LOG_LAWS  # noqa: E402


@pytest.mark.parametrize("law", LOG_LAWS)
@pytest.mark.parametrize("name", J2_CASES)
def test_j2log_golden(ctx_factory, name, law):
    """J2Log (materials.hpp:580-756): logarithmic strain through the device restatement of mfem's CalcEigenvalues
    (mb2_eig.cuh); libdevice hypot / acos / cos / log / exp against glibc in the reference. Residual and history 1e-12
    (history: Fp_inv, eqps). FD tangent: the eigenvalues of C_e = I + O(strain) are known to 1e-16 absolute in either
    implementation, so log(lambda) = O(1e-3) carries 1e-13 relative rounding noise (the residual deviates by 4e-14),
    which the forward difference divides by h = 1e-8 |u| ~ 1e-11: stated bound 1e-2 of the largest entry (the
    reference's own K has noise of that order); observed deviations are printed."""
    D = load_j2_golden(name)
    k = "_log_" + law
    ctx = ctx_factory()
    ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
    ctx.upload_quad_tables(D["dN_dX"], D["w"], D["detJ"])
    ctx.build_pattern()
    ctx.set_material(5, j2_params(law))
    assert ctx.state_size() == O.J2LOG_NS
    assert np.array_equal(ctx.get_state(), O.j2log_initial_state(D["NE"], D["nq"], D["dim"]))
    r = np.zeros(D["NV"])
    ctx.add_domain_residual(D["u"], r)
    r2 = np.zeros(D["NV"])
    A = D["mass"].copy()
    ctx.add_domain_residual_and_grad(D["u"], D["gf"], r2, A)
    ctx.post_time_advance(D["u"])
    S = ctx.get_state()
    rb = np.zeros(D["NV"])
    Ab = D["mass"].copy()
    ctx.add_domain_residual_and_grad(np.ascontiguousarray(1.5 * D["u"]), D["gf"], rb, Ab)
    print("\n[j2log %s %s] rel dev: r %.1e  A %.1e  state %.1e  rb %.1e  Ab %.1e" % (
        name, law, _rel(r, D["r" + k]), _rel(A, D["A" + k]), _rel(S, D["state" + k]), _rel(rb, D["rb" + k]),
        _rel(Ab, D["Ab" + k])))
    _print_spread("j2log_" + law)
    assert _rel(r, D["r" + k]) <= 1e-12
    assert np.array_equal(r2, r) and _rel(A, D["A" + k]) <= 1e-2
    assert _rel(S, D["state" + k]) <= 1e-12
    assert _rel(rb, D["rb" + k]) <= 1e-12 and _rel(Ab, D["Ab" + k]) <= 1e-2


@pytest.mark.parametrize("mat,law", [(2, None), (3, "jcthermo"), (3, "powerlaw"), (4, "jcthermo"), (5, "jcthermo")])
def test_accumulators_parked_in_tensor_memory_give_the_same_bits(ctx_factory, mat, law):
    """element_fd_kernel parks its 28 FP64 accumulators per thread in tensor memory (tcgen05.st / tcgen05.ld) while the
    return mapping runs (option park_acc, default 1) instead of letting them spill to local memory: a move, no
    arithmetic -- residual and Jacobian are bit-equal with the option off, on a yielded state (after one commit)."""
    name = "cube_p2_n2"
    D = load_golden(name) if mat == 2 else load_j2_golden(J2_CASES[0])
    assert (D["dim"], D["nd"], D["nq"]) == (3, 27, 64)       # the two-phase kernel's element type
    out = []
    for park in (1, 0):
        ctx = ctx_factory()
        ctx.set_option("park_acc", park)
        if mat == 2:
            setup_ctx(ctx, D, "j2linear")
        else:
            ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
            ctx.upload_quad_tables(D["dN_dX"], D["w"], D["detJ"])
            ctx.build_pattern()
            ctx.set_material(mat, j2_params(law))
        ctx.post_time_advance(D["u"])
        r = np.zeros(D["NV"])
        A = D["mass"].copy()
        ctx.add_domain_residual_and_grad(np.ascontiguousarray(1.5 * D["u"]), D["gf"], r, A)
        out.append((r, A, ctx.get_state()))
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])
    assert np.array_equal(out[0][2], out[1][2]) and np.abs(out[0][1] - D["mass"]).max() > 0


@pytest.mark.parametrize("material", ["neohookean", "stvk", "j2linear"])
def test_split_kernel_parks_its_accumulators_with_the_same_bits(ctx_factory, material):
    """p = 3 (element_fd_split_kernel): the 40 accumulators per thread are parked in tensor memory for every material
    (2 x 256 columns = the whole TMEM of an SM with two resident CTAs); bit-equal to the reference's dump with the option
    on (default) and off."""
    D = load_golden("beam4_p3")
    assert (D["dim"], D["nd"], D["nq"]) == (3, 64, 125)
    for park in (1, 0):
        ctx = ctx_factory()
        ctx.set_option("park_acc", park)
        setup_ctx(ctx, D, material)
        r = np.zeros(D["NV"])
        A = D["mass"].copy()
        ctx.add_domain_residual_and_grad(D["u"], D["gf"], r, A)
        assert np.array_equal(r, D["r2_" + material]) and np.array_equal(A, D["A_" + material])
