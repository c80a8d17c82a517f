"""CPU: the plain-C contact oracle (oracle/mimi_oracle_contact.c) against golden vectors dumped from the unmodified
reference's MortarContact (src/mimi/integrators/mortar_contact.cpp) driven by analytic rigid bodies
(oracle/ref_harness.cpp, mode "contact"). Bit-exact."""
import numpy as np
import pytest

from oracle import oracle_py as O
from tests.cases import CONTACT_CASES, load_contact_golden


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
