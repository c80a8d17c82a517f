"""CPU: the plain-C contact oracle (oracle/mimi_oracle_contact.c) against golden vectors dumped from the unmodified
reference's MortarContact (src/mimi/integrators/mortar_contact.cpp) driven by analytic rigid bodies
(oracle/ref_harness.cpp, mode "contact"). Bit-exact."""
import numpy as np
import pytest

from oracle import oracle_py as O
from tests.cases import CONTACT_CASES, SPLINE_CONTACT_CASES, load_contact_golden


@pytest.mark.parametrize("name", CONTACT_CASES)
def test_contact_oracle_bit_exact(name):
    D = load_contact_golden(name)
    dim = D["dim"]
    body = O.rigid_from_spec(dim, D["body"])
    # A11 Prepare: sorted unique marked vdofs and the reduced nodal numbering
    local, nm, marked = O.contact_prepare(dim, D["c_v_dofs"])
    assert nm == D["n_marked_nodes"] and np.array_equal(marked, D["c_marked_v_dofs"])
    # A12 ComputePressure
    gap, area, pr, tot = O.contact_pressure(dim, D["c_v_dofs"], local, nm, D["c_N"], D["c_dN_dxi"], D["c_w"],
                                            D["c_x_ref"], D["u"], body, D["penalty"])
    assert np.array_equal(gap, D["c_gap"]) and np.array_equal(area, D["c_area"])
    assert np.array_equal(pr, D["c_pressure"]) and tot == D["c_scalars"][0]
    assert 0 < (pr != 0).sum()
    # A13 AddBoundaryResidual on top of the domain residual, then AddBoundaryResidualAndGrad on the mass values
    r = D["c_r0"].copy()
    force, pint = O.contact_assemble(dim, D["c_v_dofs"], local, None, D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"],
                                     D["u"], pr, False, 0.0, r, None)
    assert np.array_equal(r, D["c_r"])
    assert np.array_equal(force, D["c_scalars"][2:5]) and pint == D["c_scalars"][1]
    r2 = np.zeros_like(r)
    A = D["mass"].copy()
    O.contact_assemble(dim, D["c_v_dofs"], local, D["c_A_ids"], D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"], D["u"],
                       pr, True, D["gf"], r2, A)
    assert np.array_equal(r2, D["c_r2"]) and np.array_equal(A, D["c_A"])
    # A14 GapNorm
    assert O.contact_gap_norm(dim, D["c_v_dofs"], D["c_N"], D["c_w"], D["c_x_ref"], D["u"], body) == D["c_scalars"][5]


def test_clamped_plane_exercises_the_angle_filter():
    """Closest points on the patch edge are not perpendicular: ElementGapAndArea zeroes those gaps
    (mortar_contact.cpp:174-181)."""
    D = load_contact_golden("contact_quad_p2_n4_plane_clamped")
    unclamped = ("plane", list(D["body"][1][:4]) + [-1e300, 1e300])
    local, nm, _ = O.contact_prepare(2, D["c_v_dofs"])
    a = O.contact_pressure(2, D["c_v_dofs"], local, nm, D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"], D["u"],
                           O.rigid_from_spec(2, D["body"]), D["penalty"])
    b = O.contact_pressure(2, D["c_v_dofs"], local, nm, D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"], D["u"],
                           O.rigid_from_spec(2, unclamped), D["penalty"])
    assert np.abs(a[0]).sum() < 0.1 * np.abs(b[0]).sum()      # most penetrating gaps are filtered out


def test_bezier_closest_point_properties():
    """Bezier rigid bodies are parity-unpinned against splinepy's search; the restated search is checked by what a
    closest point must satisfy: no sampled point is closer, and the residual is orthogonal to the tangents unless the
    point sits on the parametric boundary."""
    rng = np.random.default_rng(0)
    cps = np.array([[-2.5, 1.3], [0.3, 0.7], [0.7, 0.7], [1.5, 1.3]]) + [0.05, 1.0]     # examples/nl_contact.py
    b = O.rigid_bezier(2, cps)
    ts = np.linspace(0, 1, 20001)
    B = np.stack([(1 - ts) ** 3, 3 * ts * (1 - ts) ** 2, 3 * ts ** 2 * (1 - ts), ts ** 3], 1) @ cps
    for _ in range(100):
        q = rng.uniform([-2, 0.5], [1.5, 2.5])
        xi, pmq, d, fd = O.rigid_query(b, q)
        assert d <= np.sqrt(((B - q) ** 2).sum(1).min()) + 1e-12
        assert abs(pmq @ fd[0]) <= 1e-10 * np.linalg.norm(fd[0]) * d or xi[0] in (0.0, 1.0)
    g = np.linspace(0.0, 1.0, 3)
    cp = np.zeros((3, 3, 3))
    for j in range(3):
        for i in range(3):
            cp[j, i] = [1.4 * g[i] - 0.2, 1.4 * g[j] - 0.2, 1.0 + 1.2 * ((g[i] - 0.5) ** 2 + (g[j] - 0.5) ** 2)]
    b = O.rigid_bezier(3, cp)
    s = np.linspace(0, 1, 201)
    Bs = np.stack([(1 - s) ** 2, 2 * s * (1 - s), s ** 2], -1)
    S = np.einsum("ai,bj,jid->bad", Bs, Bs, cp).reshape(-1, 3)
    for _ in range(50):
        q = rng.uniform([0, 0, 0.8], [1, 1, 1.3])
        xi, pmq, d, fd = O.rigid_query(b, q)
        assert d <= np.sqrt(((S - q) ** 2).sum(1).min()) + 1e-12
        assert max(abs(pmq @ fd[0]), abs(pmq @ fd[1])) <= 1e-10 or 0.0 in xi or 1.0 in xi


def test_bspline_extraction_and_multipatch_nearest():
    """B-spline rigid bodies: Bezier extraction by knot insertion reproduces the B-spline (scipy evaluates the original),
    and the multi-patch closest point is the minimum over the patches (checked against dense sampling). Parity unpinned
    against splinepy like the single-patch body."""
    from scipy.interpolate import BSpline
    rng = np.random.default_rng(11)
    p = 3
    U = np.array([0, 0, 0, 0, 0.2, 0.5, 0.5, 0.8, 1, 1, 1, 1.0])
    P = rng.standard_normal((len(U) - p - 1, 2))
    patches = O.bspline_to_bezier(p, U, P)
    assert patches.shape == (4, p + 1, 2)
    spl = BSpline(U, P, p)
    breaks = [0, 0.2, 0.5, 0.8, 1.0]
    from math import comb
    for s in range(4):
        for t in np.linspace(0, 1, 7):
            B = np.array([comb(p, i) * t ** i * (1 - t) ** (p - i) for i in range(p + 1)])
            assert np.allclose(B @ patches[s], spl(breaks[s] + t * (breaks[s + 1] - breaks[s])), atol=1e-13)
    body = O.rigid_bezier_patches(2, patches, n_seed=32)
    tt = np.linspace(0, 1, 20001)
    dense = spl(tt)
    for _ in range(25):
        q = rng.standard_normal(2) * 1.5
        xi, pmq, dist, fd = O.rigid_query(body, q)
        brute = np.sqrt(((dense - q) ** 2).sum(1)).min()
        assert dist <= brute + 1e-12 and abs(dist - brute) <= 1e-6
    # surface: 2 x 2 patches reproduce the tensor-product B-spline at the patch corners' interiors
    ku = np.array([0, 0, 0, 0.4, 1, 1, 1.0])
    kv = np.array([0, 0, 0, 0.6, 1, 1, 1.0])
    S = rng.standard_normal((4, 4, 3))
    pt = O.bspline_surface_to_bezier(2, 2, ku, kv, S)
    assert pt.shape == (4, 3, 3, 3)

    def surf(u, v):
        Bu = BSpline(ku, np.eye(4), 2)(u)
        Bv = BSpline(kv, np.eye(4), 2)(v)
        return np.einsum("j,i,jid->d", Bv, Bu, S)

    ub, vb = [0, 0.4, 1.0], [0, 0.6, 1.0]
    for b in range(2):
        for a in range(2):
            for (s, t) in [(0.0, 0.0), (0.3, 0.7), (1.0, 1.0), (0.5, 0.1)]:
                Bs = np.array([comb(2, i) * s ** i * (1 - s) ** (2 - i) for i in range(3)])
                Bt = np.array([comb(2, i) * t ** i * (1 - t) ** (2 - i) for i in range(3)])
                val = np.einsum("j,i,jid->d", Bt, Bs, pt[b * 2 + a])
                assert np.allclose(val, surf(ub[a] + s * (ub[a + 1] - ub[a]), vb[b] + t * (vb[b + 1] - vb[b])), atol=1e-13)


def _arc(c, R, a0, a1):
    """Rational quadratic Bezier arc of the circle (c, R) from angle a0 to a1 (< 180 degrees apart): control points and
    weights (1, cos(half angle), 1)."""
    c = np.asarray(c, np.float64)
    h = 0.5 * (a1 - a0)
    m = 0.5 * (a0 + a1)
    P = np.array([c + R * np.array([np.cos(a0), np.sin(a0)]),
                  c + R / np.cos(h) * np.array([np.cos(m), np.sin(m)]),
                  c + R * np.array([np.cos(a1), np.sin(a1)])])
    return P, np.array([1.0, np.cos(h), 1.0])


def test_rational_arc_reproduces_the_closed_form_circle():
    """Rational patches (NURBS rigid bodies): numerator / denominator evaluation + quotient rule. A rational quadratic arc
    IS a circle, so the search must return the closed-form circle's distance, closest point and tangent direction
    (RigidSphere of the reference harness, pinned) wherever the foot point lies on the arc."""
    rng = np.random.default_rng(3)
    c, R = np.array([1.0, 1.9995]), 1.0
    P, w = _arc(c, R, np.deg2rad(210.0), np.deg2rad(330.0))
    arc = O.rigid_rational_patches(2, O.homogeneous(P, w)[None])
    sph = O.rigid_from_spec(2, ("sphere", [c[0], c[1], R]))
    n = 0
    for _ in range(200):
        q = c + rng.uniform(0.3, 1.6) * np.array([np.cos(t := np.deg2rad(rng.uniform(215.0, 325.0))), np.sin(t)])
        xi, pmq, d, fd = O.rigid_query(arc, q)
        _, pmq_s, d_s, fd_s = O.rigid_query(sph, q)
        assert abs(d - d_s) <= 1e-13 and np.abs(pmq - pmq_s).max() <= 1e-13
        tan = fd[0] / np.linalg.norm(fd[0])
        assert abs(tan @ (q + pmq - c)) <= 1e-13                     # tangent is perpendicular to the radius
        assert abs(np.array([tan[1], -tan[0]]) @ (q + pmq - c) / R - 1.0) <= 1e-12      # normal (d1, -d0) points outward
        n += 1
    assert n == 200


def test_unit_weights_equal_the_polynomial_patch_and_nurbs_extraction_keeps_the_circle():
    rng = np.random.default_rng(4)
    cps = np.array([[-2.5, 1.3], [0.3, 0.7], [0.7, 0.7], [1.5, 1.3]]) + [0.05, 1.0]
    poly = O.rigid_bezier(2, cps)
    rat = O.rigid_rational_patches(2, O.homogeneous(cps, np.ones(4))[None], n_seed=64)
    for _ in range(50):
        q = rng.uniform([-2, 0.5], [1.5, 2.5])
        a, b = O.rigid_query(poly, q), O.rigid_query(rat, q)
        assert abs(a[2] - b[2]) <= 1e-13 and np.abs(a[1] - b[1]).max() <= 1e-13 and np.abs(a[3] - b[3]).max() <= 1e-11
    # knot insertion on homogeneous points (what MB2_RIGID_NURBS does on the host): the arc with an extra knot at 0.3,
    # split back into two rational Bezier patches, is still the circle
    c, R = np.array([0.2, -0.1]), 0.7
    P, w = _arc(c, R, np.deg2rad(20.0), np.deg2rad(140.0))
    H = O.homogeneous(P, w)
    a = 0.3
    Hn = np.array([H[0], a * H[1] + (1 - a) * H[0], a * H[2] + (1 - a) * H[1], H[2]])     # Boehm, one insertion
    patches = O.bspline_to_bezier(2, [0, 0, 0, 0.3, 1, 1, 1], Hn)
    assert patches.shape == (2, 3, 3)
    body = O.rigid_rational_patches(2, patches, n_seed=32)
    for _ in range(100):
        q = c + rng.uniform(0.2, 1.5) * np.array([np.cos(t := np.deg2rad(rng.uniform(25.0, 135.0))), np.sin(t)])
        _, pmq, d, fd = O.rigid_query(body, q)
        assert abs(d - abs(np.linalg.norm(q - c) - R)) <= 1e-13
        assert abs(np.linalg.norm(q + pmq - c) - R) <= 1e-13


def test_rational_cylinder_surface():
    """3D: rational quadratic arc (u) extruded along y (v, degree 1) = a cylinder; distance = | rho - R |."""
    rng = np.random.default_rng(5)
    c, R = np.array([0.5, 2.0]), 1.0                        # axis along y through (x, z) = c
    P2, w = _arc(c, R, np.deg2rad(210.0), np.deg2rad(330.0))
    cp = np.zeros((2, 3, 3))
    for j, y in enumerate([-0.5, 1.5]):
        cp[j, :, 0], cp[j, :, 1], cp[j, :, 2] = P2[:, 0], y, P2[:, 1]
    body = O.rigid_rational_patches(3, O.homogeneous(cp, np.stack([w, w]))[None])
    for _ in range(100):
        t = np.deg2rad(rng.uniform(215.0, 325.0))
        rho = rng.uniform(0.3, 1.5)
        q = np.array([c[0] + rho * np.cos(t), rng.uniform(-0.4, 1.4), c[1] + rho * np.sin(t)])
        xi, pmq, d, fd = O.rigid_query(body, q)
        assert abs(d - abs(rho - R)) <= 1e-13 and abs(pmq[1]) <= 1e-13
        assert max(abs(pmq @ fd[0]), abs(pmq @ fd[1])) <= 1e-12


def _rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(np.asarray(b)).max(), 1e-300)


@pytest.mark.parametrize("name", SPLINE_CONTACT_CASES)
def test_spline_bodies_pinned_against_the_reference_on_real_splinepy(name):
    """A12 pinned: the golden vectors come from the reference's own NearestDistanceToSplines holding a real splinepy
    Bezier / RationalBezier / BSpline (kd-tree seed + Newton -> SLSQP -> LM, proximity.cpp:643-733) inside the unmodified
    MortarContact. The restated search (sampled seed + bound-clipped Newton) is a different iteration that converges to
    the same point: nodal gap / pressure / residual within 1e-12 relative (measured 3e-15 .. 3e-13; the gap is a 1e-4
    difference of O(1) coordinates). Everything downstream of the projection is bit-exact: fed the reference's own nodal
    pressures, residual AND FD tangent equal the reference's bit for bit (the FD loop freezes the pressure and never
    calls the projection, mortar_contact.cpp:263-295). With its own pressures the tangent inherits their 1e-13 through
    1/h: stated 1e-4 of the largest entry (measured <= 3.5e-5)."""
    D = load_contact_golden(name)
    dim = D["dim"]
    assert D["body"][0] == "spline"
    body = O.rigid_from_spec(dim, D["body"])
    local, nm, marked = O.contact_prepare(dim, D["c_v_dofs"])
    assert nm == D["n_marked_nodes"] and np.array_equal(marked, D["c_marked_v_dofs"])
    gap, area, pr, tot = O.contact_pressure(dim, D["c_v_dofs"], local, nm, D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"],
                                            D["u"], body, D["penalty"])
    assert (D["c_pressure"] != 0).sum() >= 4                       # the fixture is in contact
    assert np.array_equal(area, D["c_area"])                        # areas do not depend on the projection
    assert np.array_equal(pr != 0, D["c_pressure"] != 0)            # the same nodes are active
    assert _rel(gap, D["c_gap"]) <= 1e-12 and _rel(pr, D["c_pressure"]) <= 1e-12
    r = D["c_r0"].copy()
    O.contact_assemble(dim, D["c_v_dofs"], local, None, D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"], D["u"], pr, False,
                       0.0, r, None)
    assert _rel(r - D["c_r0"], D["c_r"] - D["c_r0"]) <= 1e-12
    r2, A = np.zeros_like(r), D["mass"].copy()
    O.contact_assemble(dim, D["c_v_dofs"], local, D["c_A_ids"], D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"], D["u"], pr,
                       True, D["gf"], r2, A)
    assert _rel(r2, D["c_r2"]) <= 1e-12
    assert _rel(A - D["mass"], D["c_A"] - D["mass"]) <= 1e-4
    assert abs(O.contact_gap_norm(dim, D["c_v_dofs"], D["c_N"], D["c_w"], D["c_x_ref"], D["u"], body) - D["c_scalars"][5]) \
        <= 1e-12 * D["c_scalars"][5]
    # downstream of the projection: bit-exact
    r3, A3 = np.zeros_like(r), D["mass"].copy()
    O.contact_assemble(dim, D["c_v_dofs"], local, D["c_A_ids"], D["c_N"], D["c_dN_dxi"], D["c_w"], D["c_x_ref"], D["u"],
                       D["c_pressure"], True, D["gf"], r3, A3)
    assert np.array_equal(r3, D["c_r2"]) and np.array_equal(A3, D["c_A"])


@pytest.mark.parametrize("name", SPLINE_CONTACT_CASES)
def test_closest_point_query_against_live_splinepy(name):
    """The query itself, point by point, against NearestDistanceToSplines::NearestDistance on real splinepy (harness mode
    "proximity", needs oracle/_ref): parametric coordinate, closest point, first derivatives, unit normal and normal gap
    of random points around the contact boundary agree to 1e-13 absolute (measured <= 4e-15)."""
    if not O.have_ref():
        pytest.skip("oracle/_ref/mimi_ref not present on this box")
    D = load_contact_golden(name)
    dim = D["dim"]
    desc = D["body"][1]
    rng = np.random.default_rng(7)
    lo = [0.0] * (dim - 1) + [0.97]
    hi = ([2.0] if dim == 2 else [1.0, 1.0]) + [1.03]
    pts = rng.uniform(lo, hi, size=(96, dim))
    R = O.run_ref_proximity(pts, desc)
    body = O.rigid_from_desc(dim, desc)
    has_knots = O.parse_spline_desc(dim, desc)["knots"] is not None
    for i, q in enumerate(pts):
        xi, pmq, dist, fd = O.rigid_query(body, q)
        assert np.abs(pmq - R["pmq"][i]).max() <= 1e-13
        assert abs(dist - R["distance"][i]) <= 1e-13
        if not has_knots:   # Bezier patches extracted from a B-spline are re-parameterised on [0, 1]: only directions compare
            assert np.abs(fd - R["first_derivatives"][i]).max() <= 1e-12
        # normal gap as MortarContact forms it (nearest_distance.hpp:139-193)
        if dim == 2:
            n = np.array([fd[0, 1], -fd[0, 0]]) / np.hypot(fd[0, 0], fd[0, 1])
        else:
            n = np.cross(fd[0], fd[1])
            n /= np.linalg.norm(n)
        assert np.abs(n - R["normal"][i]).max() <= 1e-12
        assert abs(-(n @ pmq) - R["normal_gap"][i]) <= 1e-13
