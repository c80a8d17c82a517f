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
