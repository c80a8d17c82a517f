"""GPU: rigid-body mortar contact through the C ABI against golden vectors dumped
from the unmodified reference's MortarContact (src/mimi/integrators/mortar_contact.cpp) and against the C oracle.
Bar: index maps bit-exact; nodal gap / area / pressure, residual and Jacobian bit-exact in FP64 (north_star: 1e-12
relative); whole-boundary scalars (area, force, pressure integral, gap norm) within 1e-12 relative -- they are summed
per face and then across faces, not as the reference's single running sum."""
import numpy as np
import pytest

from mimi_b200 import capi
from oracle import oracle_py as O
from tests.cases import CONTACT_CASES, SPLINE_CONTACT_CASES, load_contact_golden

pytestmark = pytest.mark.gpu

KIND = {"plane": capi.MB2_RIGID_PLANE, "sphere": capi.MB2_RIGID_SPHERE}


def setup_contact(ctx, D, body=None, penalty=None):
    body = body or D["body"]
    ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
    ctx.set_nodes(D["x_ref"])
    ctx.build_pattern()
    ctx.contact_set_faces(D["c_v_dofs"], D["c_N"], D["c_dN_dxi"], D["c_w"])
    ctx.contact_set_rigid_body(KIND[body[0]], O.rigid_params(D["dim"], body), penalty or D["penalty"])
    return ctx


def close(a, b, tol=1e-12):
    return np.abs(np.asarray(a) - np.asarray(b)).max() <= tol * max(np.abs(np.asarray(b)).max(), 1e-300)


@pytest.mark.parametrize("name", CONTACT_CASES)
def test_contact_golden(ctx_factory, name):
    D = load_contact_golden(name)
    ctx = setup_contact(ctx_factory(), D)
    # A11 Prepare: index maps
    I, J = ctx.get_pattern()
    assert np.array_equal(I, D["I"]) and np.array_equal(J, D["J"])
    assert ctx.n_marked == D["n_marked_nodes"]
    assert np.array_equal(ctx.contact_marked_v_dofs(), D["c_marked_v_dofs"])
    assert np.array_equal(ctx.contact_A_ids(), D["c_A_ids"])
    # A12 + A13 residual, added on top of the domain residual
    r = D["c_r0"].copy()
    ctx.add_boundary_residual(D["u"], r)
    gap, area, pr = ctx.contact_nodal()
    assert np.array_equal(gap, D["c_gap"]) and np.array_equal(area, D["c_area"]) and np.array_equal(pr, D["c_pressure"])
    assert np.array_equal(r, D["c_r"])
    s = ctx.contact_last_scalars()
    assert close(s[0], D["c_scalars"][0]) and close(s[1], D["c_scalars"][1]) and close(s[2:5], D["c_scalars"][2:5])
    # A13 residual + FD tangent, added onto the mass values
    r2 = np.zeros(D["NV"])
    A = D["mass"].copy()
    ctx.add_boundary_residual_and_grad(D["u"], D["gf"], r2, A)
    assert np.array_equal(r2, D["c_r2"])
    assert np.abs(A - D["c_A"]).max() <= 1e-12 * np.abs(D["c_A"]).max()   # north_star tolerance
    assert np.array_equal(A, D["c_A"])                                     # and in fact bit-exact
    assert np.abs(A - D["mass"]).max() > 0
    # A14 GapNorm
    assert close(ctx.contact_gap_norm(D["u"]), D["c_scalars"][5])


@pytest.mark.parametrize("name", SPLINE_CONTACT_CASES)
def test_spline_contact_against_the_reference_on_real_splinepy(ctx_factory, name):
    """A12 pinned on the device: golden vectors = the unmodified MortarContact driven by the reference's own
    NearestDistanceToSplines on real splinepy splines (Bezier curve of examples/nl_contact.py, Bezier surface, rational
    arc, B-spline with interior knots). The device's closest-point search converges to the same points: nodal gap /
    pressure / residual within 1e-12 relative, FD tangent within 1e-4 of its largest entry (it divides the pressures'
    1e-13 by h; fed the reference's pressures the oracle's tangent is bit-exact, tests/test_oracle_contact.py), and the
    device equals the C oracle of the same search bit for bit."""
    D = load_contact_golden(name)
    dim = D["dim"]
    kind, params = O.mb2_body_from_desc(dim, D["body"][1])
    ctx = ctx_factory()
    ctx.set_connectivity(dim, D["nd"], D["nq"], D["NV"], D["v_dofs"])
    ctx.set_nodes(D["x_ref"])
    ctx.build_pattern()
    ctx.contact_set_faces(D["c_v_dofs"], D["c_N"], D["c_dN_dxi"], D["c_w"])
    ctx.contact_set_rigid_body(kind, params, D["penalty"])
    assert np.array_equal(ctx.contact_A_ids(), D["c_A_ids"])
    r = D["c_r0"].copy()
    ctx.add_boundary_residual(D["u"], r)
    gap, area, pr = ctx.contact_nodal()
    assert np.array_equal(area, D["c_area"]) and np.array_equal(pr != 0, D["c_pressure"] != 0) and (pr != 0).sum() >= 4
    assert close(gap, D["c_gap"]) and close(pr, D["c_pressure"])
    assert close(r - D["c_r0"], D["c_r"] - D["c_r0"])
    r2 = np.zeros(D["NV"])
    A = D["mass"].copy()
    ctx.add_boundary_residual_and_grad(D["u"], D["gf"], r2, A)
    assert close(r2, D["c_r2"])
    assert close(A - D["mass"], D["c_A"] - D["mass"], 1e-4)
    assert close(ctx.contact_gap_norm(D["u"]), D["c_scalars"][5])
    # device == oracle of the same search, bit for bit
    body = O.rigid_from_desc(dim, D["body"][1])
    local, nm, _ = O.contact_prepare(dim, D["c_v_dofs"])
    g_o, a_o, p_o, _ = O.contact_pressure(dim, D["c_v_dofs"], local, nm, D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"],
                                          D["u"], body, D["penalty"])
    r_o, A_o = np.zeros(D["NV"]), D["mass"].copy()
    O.contact_assemble(dim, D["c_v_dofs"], local, D["c_A_ids"], D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"], D["u"], p_o,
                       True, D["gf"], r_o, A_o)
    assert np.array_equal(gap, g_o) and np.array_equal(pr, p_o) and np.array_equal(r2, r_o) and np.array_equal(A, A_o)


def test_contact_moving_body_vs_oracle(ctx_factory):
    """The rigid body moves between steps (examples/nl_contact.py moves the spline every step): no contact at all
    (every face skipped), then deeper penetration with another penalty; checked against the C oracle."""
    D = load_contact_golden("contact_cube_p2_n4_sphere")
    ctx = setup_contact(ctx_factory(), D)
    dim = 3
    local, nm, _ = O.contact_prepare(dim, D["c_v_dofs"])
    for zc, penalty in ((2.5, 1e5), (1.998, 3e6)):
        body = ("sphere", [0.5, 0.5, zc, 1.0])
        ctx.contact_set_rigid_body(KIND["sphere"], O.rigid_params(dim, body), penalty)
        gap, area, pr, tot = O.contact_pressure(dim, D["c_v_dofs"], local, nm, D["c_N"], D["c_dN_dxi"], D["c_w"],
                                                D["c_x_ref"], D["u"], O.rigid_from_spec(dim, body), penalty)
        r_o = D["c_r0"].copy()
        A_o = D["mass"].copy()
        O.contact_assemble(dim, D["c_v_dofs"], local, D["c_A_ids"], D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"],
                           D["u"], pr, True, 0.25, r_o, A_o)
        r = D["c_r0"].copy()
        A = D["mass"].copy()
        ctx.add_boundary_residual_and_grad(D["u"], 0.25, r, A)
        assert np.array_equal(ctx.contact_nodal()[2], pr)
        assert np.array_equal(r, r_o) and np.array_equal(A, A_o)
        if zc > 2.4:
            assert not pr.any() and np.array_equal(A, D["mass"])     # no contact: nothing is touched
        else:
            assert (pr != 0).sum() > 8


def test_contact_after_domain_on_device(ctx_factory):
    """Domain + contact assembled back to back on the resident r / A (the order forms::Nonlinear::AddMultGrad uses,
    src/mimi/forms/nonlinear.hpp:53-116), vs oracle domain + oracle contact."""
    D = load_contact_golden("contact_cube_p2_n2_plane")
    ctx = setup_contact(ctx_factory(), D)
    from tests.cases import load_golden
    G = load_golden("cube_p2_n2")
    assert np.array_equal(G["v_dofs"], D["v_dofs"])
    ctx.upload_quad_tables(G["dN_dX"], G["w"], G["detJ"])
    lam, mu = O.lame()
    ctx.set_material(0, [lam, mu])
    ctx.set_u(D["u"]); ctx.set_r(None); ctx.set_A(D["mass"])
    ctx.assemble_residual_and_grad_device(0.7)
    ctx.contact_residual_and_grad_device(0.7)
    r, A = ctx.get_r(), ctx.get_A()
    r_o, A_o = O.assemble_residual_and_grad(3, 27, 64, O.mat(0), G["v_dofs"], G["A_ids"], G["dN_dX"], G["w"], G["detJ"],
                                            D["u"], 0.7, D["mass"])
    local, nm, _ = O.contact_prepare(3, D["c_v_dofs"])
    pr = D["c_pressure"]
    O.contact_assemble(3, D["c_v_dofs"], local, D["c_A_ids"], D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"], D["u"], pr,
                       True, 0.7, r_o, A_o)
    assert np.array_equal(r, r_o) and np.array_equal(A, A_o)


def test_contact_call_order_errors(ctx_factory):
    D = load_contact_golden("contact_quad_p2_n4_plane_clamped")
    ctx = ctx_factory()
    ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
    with pytest.raises(capi.Mb2Error):
        ctx.contact_set_faces(D["c_v_dofs"], D["c_N"], D["c_dN_dxi"], D["c_w"])   # pattern first
    ctx.build_pattern()
    ctx.contact_set_faces(D["c_v_dofs"], D["c_N"], D["c_dN_dxi"], D["c_w"])
    r = np.zeros(D["NV"])
    with pytest.raises(capi.Mb2Error):
        ctx.add_boundary_residual(D["u"], r)                                         # no rigid body yet


# ---- Bezier rigid bodies: the device repeats the C oracle's closest-point search bit for bit; against the reference
# (splinepy's search) this body kind is parity-unpinned -- see oracle/mimi_oracle_contact.c and DESIGN.md section 5.
def _bezier_cases():
    # the curve of examples/nl_contact.py (its normal (d1, -d0) points down onto the body's top edge), moved into light
    # penetration of the 2 x 1 quad fixture: the curve's lowest point is 0.85 above its control polygon's y origin
    curve = (np.array([[-2.5, 1.3], [0.3, 0.7], [0.7, 0.7], [1.5, 1.3]]) + [0.6, 0.1495])
    g = np.linspace(0.0, 1.0, 3)
    dome = np.zeros((3, 3, 3))
    for j in range(3):
        for i in range(3):
            # u runs along y and v along x so that S_u x S_v points down (-z), towards the cube below
            dome[j, i] = [1.4 * g[j] - 0.2, 1.4 * g[i] - 0.2, 0.6996 + 1.2 * ((g[i] - 0.5) ** 2 + (g[j] - 0.5) ** 2)]
    return {"contact_quad_p3_n4_sphere": curve, "contact_cube_p2_n4_sphere": dome}


@pytest.mark.parametrize("name", ["contact_quad_p3_n4_sphere", "contact_cube_p2_n4_sphere"])
def test_contact_bezier_body_vs_oracle(ctx_factory, name):
    D = load_contact_golden(name)
    dim = D["dim"]
    cps = _bezier_cases()[name]
    ctx = setup_contact(ctx_factory(), D)
    ctx.contact_set_rigid_body(capi.MB2_RIGID_BEZIER, O.bezier_params(dim, cps), D["penalty"])
    body = O.rigid_bezier(dim, cps)
    local, nm, _ = O.contact_prepare(dim, D["c_v_dofs"])
    gap, area, pr, tot = O.contact_pressure(dim, D["c_v_dofs"], local, nm, D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"],
                                            D["u"], body, D["penalty"])
    assert 2 <= (pr != 0).sum()                        # the body really touches
    r_o, A_o = D["c_r0"].copy(), D["mass"].copy()
    O.contact_assemble(dim, D["c_v_dofs"], local, D["c_A_ids"], D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"], D["u"], pr,
                       True, D["gf"], r_o, A_o)
    r, A = D["c_r0"].copy(), D["mass"].copy()
    ctx.add_boundary_residual_and_grad(D["u"], D["gf"], r, A)
    g2, a2, p2 = ctx.contact_nodal()
    assert np.array_equal(g2, gap) and np.array_equal(a2, area) and np.array_equal(p2, pr)
    assert np.array_equal(r, r_o) and np.array_equal(A, A_o)
    assert close(ctx.contact_gap_norm(D["u"]),
                 O.contact_gap_norm(dim, D["c_v_dofs"], D["c_N"], D["c_w"], D["c_x_ref"], D["u"], body))


@pytest.mark.parametrize("name,world", [("contact_cube_p2_n4_sphere", 2), ("contact_cube_p2_n4_sphere", 3),
                                        ("contact_quad_p3_n4_sphere", 2)])
def test_contact_on_emulated_ranks(ctx_factory, name, world):
    """Partitioned run: every rank evaluates ALL marked faces (faces << elements, so pass 1 is
    replicated instead of all-reduced: no collective, identical nodal pressures everywhere) and gathers the face
    contributions only into the CSR rows / residual entries of the nodes it owns. The union of the owned rows must equal
    the reference's assembly bit for bit."""
    from mimi_b200 import partition as P
    D = load_contact_golden(name)
    dim = D["dim"]
    ranges = P.element_ranges(D["v_dofs"].shape[0], world)
    gI = D["I"].astype(np.int64)
    r_glob, A_glob = np.full(D["NV"], np.nan), np.full(D["nnz"], np.nan)
    touched = 0
    for rk in range(world):
        R = P.rank_problem(D["v_dofs"], dim, D["NV"], ranges, rk)
        ctx = ctx_factory()
        ctx.set_connectivity(dim, D["nd"], D["nq"], D["NV"], R["v_dofs"])
        ctx.mg_setup(R, manual=True)
        ctx.set_nodes(D["x_ref"])
        ctx.build_pattern()
        ctx.contact_set_faces(D["c_v_dofs"], D["c_N"], D["c_dN_dxi"], D["c_w"])
        ctx.contact_set_rigid_body(KIND[D["body"][0]], O.rigid_params(dim, D["body"]), D["penalty"])
        rows = np.repeat(R["owned"], dim)
        idx = np.concatenate([np.arange(gI[v], gI[v + 1]) for v in np.nonzero(rows)[0]])
        ctx.set_u(D["u"]); ctx.set_r(D["c_r0"]); ctx.set_A(np.ascontiguousarray(D["mass"][idx]))
        ctx.contact_residual_and_grad_device(D["gf"])
        assert np.array_equal(ctx.contact_nodal()[2], D["c_pressure"])       # replicated pass 1
        r_loc, A_loc = ctx.get_r(), ctx.get_A()
        r_glob[rows] = r_loc[rows]
        A_glob[idx] = A_loc
        touched += int((A_loc != D["mass"][idx]).any())
    # AddBoundaryResidualAndGrad started from r = 0 in the fixture (c_r2); here from the domain residual c_r0
    assert np.array_equal(r_glob, D["c_r0"] + (D["c_r"] - D["c_r0"])) or np.array_equal(r_glob, D["c_r"])
    assert np.array_equal(A_glob, D["c_A"])
    assert touched >= 1


# ---- B-spline rigid bodies (interior knots): split into Bezier patches, every patch queried, the nearest wins
def _bspline_cases():
    """The Bezier bodies above with knots inserted (same geometry, now 3 / 2x2 patches) -- elevating the knot vector does
    not move the curve, so everything must agree with the single-patch body to rounding."""
    from scipy.interpolate import BSpline  # noqa: F401  (only to make the intent explicit; the insertion is our own)
    curve, dome = _bezier_cases()["contact_quad_p3_n4_sphere"], _bezier_cases()["contact_cube_p2_n4_sphere"]
    return {"contact_quad_p3_n4_sphere": (3, 0, curve), "contact_cube_p2_n4_sphere": (2, 2, dome)}


def _insert_knots(p, U, P, new):
    """Boehm insertion, one knot at a time (the B-spline a user would hand over: same curve, interior knots)."""
    U = list(U)
    P = [np.array(x, float) for x in P]
    for u in new:
        k = max(i for i in range(len(U) - 1) if U[i] <= u < U[i + 1])
        Q = []
        for i in range(len(P) + 1):
            if i <= k - p:
                Q.append(P[i])
            elif i > k:
                Q.append(P[i - 1])
            else:
                a = (u - U[i]) / (U[i + p] - U[i])
                Q.append(a * P[i] + (1.0 - a) * P[i - 1])
        P = Q
        U.insert(k + 1, u)
    return np.array(U), np.array(P)


@pytest.mark.parametrize("name", ["contact_quad_p3_n4_sphere", "contact_cube_p2_n4_sphere"])
def test_contact_bspline_body(ctx_factory, name):
    D = load_contact_golden(name)
    dim = D["dim"]
    pu, pv, cps = _bspline_cases()[name]
    if dim == 2:
        ku, P = _insert_knots(pu, [0.0] * (pu + 1) + [1.0] * (pu + 1), cps, [0.3, 0.55])
        kv = []
        patches = O.bspline_to_bezier(pu, ku, P)
    else:
        ku, rows = None, []
        for j in range(cps.shape[0]):                                  # u direction, every row
            ku, Pj = _insert_knots(pu, [0.0] * (pu + 1) + [1.0] * (pu + 1), cps[j], [0.4])
            rows.append(Pj)
        rows = np.array(rows)
        cols = []
        for i in range(rows.shape[1]):                                 # v direction, every column
            kv, Pi = _insert_knots(pv, [0.0] * (pv + 1) + [1.0] * (pv + 1), rows[:, i], [0.6])
            cols.append(Pi)
        P = np.transpose(np.array(cols), (1, 0, 2))                    # [ncp_v][ncp_u][3]
        patches = O.bspline_surface_to_bezier(pu, pv, ku, kv, P)
    assert patches.shape[0] == (3 if dim == 2 else 4)

    def run(kind, params):
        ctx = setup_contact(ctx_factory(), D)
        ctx.contact_set_rigid_body(kind, params, D["penalty"])
        r, A = D["c_r0"].copy(), D["mass"].copy()
        ctx.add_boundary_residual_and_grad(D["u"], D["gf"], r, A)
        return r, A, ctx.contact_nodal(), ctx.contact_gap_norm(D["u"])

    # (1) the library's own extraction == the patches handed over explicitly, bit for bit
    r_s, A_s, nod_s, gn_s = run(capi.MB2_RIGID_BSPLINE, O.bspline_params(dim, pu, pv, ku, kv, P))
    r_p, A_p, nod_p, gn_p = run(capi.MB2_RIGID_BEZIER, O.bezier_patches_params(dim, patches))
    assert np.array_equal(r_s, r_p) and np.array_equal(A_s, A_p) and gn_s == gn_p
    # (2) device == C oracle on the multi-patch body, bit for bit
    body = O.rigid_bezier_patches(dim, patches)
    local, nm, _ = O.contact_prepare(dim, D["c_v_dofs"])
    gap, area, pr, tot = O.contact_pressure(dim, D["c_v_dofs"], local, nm, D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"],
                                            D["u"], body, D["penalty"])
    assert 2 <= (pr != 0).sum()
    r_o, A_o = D["c_r0"].copy(), D["mass"].copy()
    O.contact_assemble(dim, D["c_v_dofs"], local, D["c_A_ids"], D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"], D["u"], pr,
                       True, D["gf"], r_o, A_o)
    assert np.array_equal(nod_p[0], gap) and np.array_equal(nod_p[1], area) and np.array_equal(nod_p[2], pr)
    assert np.array_equal(r_p, r_o) and np.array_equal(A_p, A_o)
    # (3) same geometry as the single Bezier patch: gaps / pressures agree to rounding
    r_1, A_1, nod_1, gn_1 = run(capi.MB2_RIGID_BEZIER, O.bezier_params(dim, cps))
    assert np.allclose(nod_p[0], nod_1[0], rtol=0, atol=1e-12 * max(1.0, np.abs(nod_1[0]).max()))
    assert np.allclose(nod_p[2], nod_1[2], rtol=1e-6, atol=1e-6 * np.abs(nod_1[2]).max())
    assert close(gn_p, gn_1, 1e-8)


# ---- rational patches / NURBS rigid bodies (MB2_RIGID_RATIONAL_BEZIER, MB2_RIGID_NURBS)
def _arc(c, R, a0, a1):
    c = np.asarray(c, np.float64)
    h, m = 0.5 * (a1 - a0), 0.5 * (a0 + a1)
    P = np.array([c + R * np.array([np.cos(a0), np.sin(a0)]),
                  c + R / np.cos(h) * np.array([np.cos(m), np.sin(m)]),
                  c + R * np.array([np.cos(a1), np.sin(a1)])])
    return P, np.array([1.0, np.cos(h), 1.0])


def _rational_cases():
    # 2D: the lower 120 degrees of the fixture's own rigid circle as a rational quadratic arc (left to right, so that the
    # normal (d1, -d0) points down onto the body): the same geometry as the reference's RigidSphere fixture
    D = load_contact_golden("contact_quad_p3_n4_sphere")
    arc, w = _arc(D["body"][1][:2], D["body"][1][2], np.deg2rad(210.0), np.deg2rad(330.0))
    # 3D: the Bezier dome with non-uniform weights (pulls the surface towards the middle control point)
    dome = _bezier_cases()["contact_cube_p2_n4_sphere"]
    wd = np.array([[1.0, 0.8, 1.0], [0.8, 1.7, 0.8], [1.0, 0.8, 1.0]])
    return {"contact_quad_p3_n4_sphere": (arc, w), "contact_cube_p2_n4_sphere": (dome, wd)}


def _run_body(ctx_factory, D, kind, params):
    ctx = setup_contact(ctx_factory(), D)
    ctx.contact_set_rigid_body(kind, params, D["penalty"])
    r, A = D["c_r0"].copy(), D["mass"].copy()
    ctx.add_boundary_residual_and_grad(D["u"], D["gf"], r, A)
    return r, A, ctx.contact_nodal(), ctx.contact_gap_norm(D["u"])


def _oracle_body(D, body):
    dim = D["dim"]
    local, nm, _ = O.contact_prepare(dim, D["c_v_dofs"])
    gap, area, pr, tot = O.contact_pressure(dim, D["c_v_dofs"], local, nm, D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"],
                                            D["u"], body, D["penalty"])
    r_o, A_o = D["c_r0"].copy(), D["mass"].copy()
    O.contact_assemble(dim, D["c_v_dofs"], local, D["c_A_ids"], D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"], D["u"], pr,
                       True, D["gf"], r_o, A_o)
    return r_o, A_o, (gap, area, pr)


@pytest.mark.parametrize("name", ["contact_quad_p3_n4_sphere", "contact_cube_p2_n4_sphere"])
def test_contact_rational_body_vs_oracle(ctx_factory, name):
    D = load_contact_golden(name)
    dim = D["dim"]
    cps, w = _rational_cases()[name]
    r, A, nod, gn = _run_body(ctx_factory, D, capi.MB2_RIGID_RATIONAL_BEZIER, O.rational_bezier_params(dim, cps[None], w[None]))
    body = O.rigid_rational_patches(dim, O.homogeneous(cps, w)[None])
    r_o, A_o, nod_o = _oracle_body(D, body)
    assert 2 <= (nod_o[2] != 0).sum()
    for a, b in zip(nod, nod_o):
        assert np.array_equal(a, b)                                    # device == oracle, bit for bit
    assert np.array_equal(r, r_o) and np.array_equal(A, A_o)
    assert close(gn, O.contact_gap_norm(dim, D["c_v_dofs"], D["c_N"], D["c_w"], D["c_x_ref"], D["u"], body))
    if dim == 2:
        # the arc IS the fixture's circle: the reference's own MortarContact dumps (RigidSphere scene) are reproduced
        assert np.allclose(nod[2], D["c_pressure"], rtol=0, atol=1e-10 * np.abs(D["c_pressure"]).max())
        assert np.allclose(nod[0], D["c_gap"], rtol=0, atol=1e-13)
        assert np.allclose(r, D["c_r"], rtol=0, atol=1e-10 * np.abs(D["c_r"]).max())
        assert np.allclose(A, D["c_A"], rtol=0, atol=1e-7 * np.abs(D["c_A"] - D["mass"]).max())
    else:
        # weights matter: the unweighted dome gives different pressures
        _, _, nod_1, _ = _run_body(ctx_factory, D, capi.MB2_RIGID_BEZIER, O.bezier_params(dim, cps))
        assert not np.allclose(nod[2], nod_1[2], rtol=1e-3, atol=0)


@pytest.mark.parametrize("name", ["contact_quad_p3_n4_sphere", "contact_cube_p2_n4_sphere"])
def test_contact_nurbs_body(ctx_factory, name):
    """MB2_RIGID_NURBS: knot insertion on the homogeneous points inside the library == the rational patches handed over
    explicitly, bit for bit; and knot refinement does not change the geometry (results equal the single patch's to
    rounding)."""
    D = load_contact_golden(name)
    dim = D["dim"]
    cps, w = _rational_cases()[name]
    H = O.homogeneous(cps, w)
    pu = 2
    if dim == 2:
        pv, kv = 0, []
        ku, Hn = _insert_knots(pu, [0.0] * 3 + [1.0] * 3, H, [0.3, 0.55])
        patches = O.bspline_to_bezier(pu, ku, Hn)
    else:
        pv = 2
        rows = []
        for j in range(3):
            ku, Hj = _insert_knots(pu, [0.0] * 3 + [1.0] * 3, H[j], [0.4])
            rows.append(Hj)
        rows = np.array(rows)
        cols = []
        for i in range(rows.shape[1]):
            kv, Hi = _insert_knots(pv, [0.0] * 3 + [1.0] * 3, rows[:, i], [0.6])
            cols.append(Hi)
        Hn = np.transpose(np.array(cols), (1, 0, 2))
        patches = O.bspline_surface_to_bezier(pu, pv, ku, kv, Hn)
    assert patches.shape[0] == (3 if dim == 2 else 4) and patches.shape[-1] == dim + 1
    # the refined NURBS in the caller's format {x, w}; the library multiplies back: exact only if x * w round-trips,
    # so hand the library dyadic-safe data by re-deriving the homogeneous points the way it does
    Pn, wn = Hn[..., :dim] / Hn[..., dim:], Hn[..., dim]
    Hlib = O.homogeneous(Pn, wn)
    patches_lib = (O.bspline_to_bezier(pu, ku, Hlib) if dim == 2 else O.bspline_surface_to_bezier(pu, pv, ku, kv, Hlib))
    r_s, A_s, nod_s, gn_s = _run_body(ctx_factory, D, capi.MB2_RIGID_NURBS, O.nurbs_params(dim, pu, pv, ku, kv, Pn, wn))
    body = O.rigid_rational_patches(dim, patches_lib)
    r_o, A_o, nod_o = _oracle_body(D, body)
    for a, b in zip(nod_s, nod_o):
        assert np.array_equal(a, b)
    assert np.array_equal(r_s, r_o) and np.array_equal(A_s, A_o)
    r_1, A_1, nod_1, gn_1 = _run_body(ctx_factory, D, capi.MB2_RIGID_RATIONAL_BEZIER,
                                      O.rational_bezier_params(dim, cps[None], w[None]))
    assert np.allclose(nod_s[0], nod_1[0], rtol=0, atol=1e-12 * max(1.0, np.abs(nod_1[0]).max()))
    assert np.allclose(nod_s[2], nod_1[2], rtol=1e-6, atol=1e-6 * np.abs(nod_1[2]).max())
    assert close(gn_s, gn_1, 1e-8)
    with pytest.raises(capi.Mb2Error):                                   # non-positive weight
        bad = O.rational_bezier_params(dim, cps[None], (w * 0.0)[None])
        setup_contact(ctx_factory(), D).contact_set_rigid_body(capi.MB2_RIGID_RATIONAL_BEZIER, bad, D["penalty"])
