"""CPU: the oracle's restatement of the setup forms (oracle/mimi_oracle_forms.c: mfem::VectorMassIntegrator,
VectorDiffusionIntegrator, VectorDomainLFIntegrator on one NURBS patch) is BIT-identical to what the compiled reference
dumped: the `mass` array of the fixtures and tests/golden/forms.npz (oracle/make_golden.py bodyforce)."""
import os

import numpy as np
import pytest

from mimi_b200 import meshgen
from oracle import oracle_py as O
from tests.cases import GOLDEN, load_golden

FIX = {"cube_p2_n2": 2, "cube_p1_n2": 1, "cube_p3_n2": 3, "cube_p2_n3_patch": 2, "quad_p2_n4": 2, "beam2d_p3_n2": 3,
       "balken_p3_n2": 3}


def patch_of(D, p):
    """knots / ncp / el_ijk of a box fixture read off its own arrays (see tests/test_gpu_mass.py)."""
    dim, nd = D["dim"], D["nd"]
    xr = D["x_ref"].reshape(-1, dim)
    first = D["v_dofs"].reshape(D["NE"], dim, nd)[:, 0, 0] // dim
    ncp, knots, ijk = [], [], []
    for d in range(dim):
        distinct = np.unique(xr[:, d])
        ncp.append(len(distinct))
        knots.append(meshgen.open_uniform_knots(p, len(distinct) - p))
        ijk.append(np.searchsorted(distinct, xr[first, d]))
    return np.array(ncp, np.int32), knots, np.stack(ijk, 1).astype(np.int32)


def forms():
    z = np.load(os.path.join(GOLDEN, "forms.npz"))
    return {k: z[k] for k in z.files}


@pytest.mark.parametrize("name", sorted(FIX))
def test_oracle_forms_bit_exact_vs_reference(name):
    D, G, p = load_golden(name), forms(), FIX[name]
    dim, nd = D["dim"], D["nd"]
    _, knots, ijk = patch_of(D, p)
    Me = O.form_element_matrices(0, dim, p, D["v_dofs"], ijk, knots, D["x_ref"])
    assert np.array_equal(Me, np.transpose(Me, (0, 2, 1)))
    M = O.assemble_bilinear_form(dim, nd, D["v_dofs"], D["I"], D["J"], Me)
    assert np.array_equal(M, D["mass"]), np.abs(M - D["mass"]).max()
    Se = O.form_element_matrices(2, dim, p, D["v_dofs"], ijk, knots, D["x_ref"], coef=float(G["viscosity"]))
    S = O.assemble_bilinear_form(dim, nd, D["v_dofs"], D["I"], D["J"], Se)
    assert np.array_equal(S, G["visc_" + name]), np.abs(S - G["visc_" + name]).max()
    be = O.form_element_vectors(dim, p, D["v_dofs"], ijk, knots, D["x_ref"], G["f"][:dim])
    b = O.assemble_linear_form(D["NV"], D["v_dofs"], be)
    assert np.array_equal(b, G[name]), np.abs(b - G[name]).max()
    # body force, then the traction of the loaded boundary on top: the reference's rhs linear form
    rhs = O.add_traction_form(b.copy(), dim, G["tvd_" + name], G["tN_" + name], G["tw_" + name], G["tdet_" + name],
                              G["traction"][:dim])
    assert not np.array_equal(rhs, b)
    assert np.array_equal(rhs, G["rhs_" + name]), np.abs(rhs - G["rhs_" + name]).max()


def test_oracle_forms_properties():
    D, p = load_golden("cube_p2_n3_patch"), 2
    _, knots, ijk = patch_of(D, p)
    Me = O.form_element_matrices(0, 3, p, D["v_dofs"], ijk, knots, D["x_ref"], coef=2.0)
    assert abs(Me.sum() - 2.0 * 1.5) < 1e-13                      # rho x volume
    Se = O.form_element_matrices(2, 3, p, D["v_dofs"], ijk, knots, D["x_ref"])
    assert np.abs(Se.sum(axis=2)).max() < 1e-13                   # gradients of a partition of unity sum to zero
    # NURBS weights: a uniform scaling of all weights leaves the rational basis unchanged up to rounding
    w = np.full(D["NV"] // 3, 0.5)
    Mw = O.form_element_matrices(0, 3, p, D["v_dofs"], ijk, knots, D["x_ref"], coef=2.0, weights=w)
    assert np.allclose(Mw, Me, rtol=1e-13, atol=0)


from tests.cases import CURVED_CASES, patch_description  # noqa: E402


@pytest.mark.parametrize("name", CURVED_CASES)
def test_oracle_forms_on_rational_and_curved_geometry_bit_exact_vs_reference(name):
    """Rows A2 / f3 on NURBS geometry proper: exact quarter annulus / thick quarter cylinder (weights 1, sqrt(1/2), 1:
    the rational branch of fe_nurbs.cpp:243-311, non-constant Jacobian) and the reference's own non-affine
    square-nurbs.mesh. Mass, viscosity, body force and body force + traction equal MFEM's integrators inside the compiled
    reference bit for bit, with the patch description MFEM itself holds after refinement."""
    D = load_golden(name)
    dim, nd = D["dim"], D["nd"]
    p, ncp, knots, ijk, w = patch_description(D)
    Me = O.form_element_matrices(0, dim, p, D["v_dofs"], ijk, knots, D["x_ref"], weights=w)
    M = O.assemble_bilinear_form(dim, nd, D["v_dofs"], D["I"], D["J"], Me)
    assert np.array_equal(M, D["mass"]), np.abs(M - D["mass"]).max()
    Se = O.form_element_matrices(2, dim, p, D["v_dofs"], ijk, knots, D["x_ref"], coef=0.37, weights=w)
    S = O.assemble_bilinear_form(dim, nd, D["v_dofs"], D["I"], D["J"], Se)
    assert np.array_equal(S, D["viscform"]), np.abs(S - D["viscform"]).max()
    f = np.array([0.3, -1.7, 0.9])[:dim]
    be = O.form_element_vectors(dim, p, D["v_dofs"], ijk, knots, D["x_ref"], f, weights=w)
    b = O.assemble_linear_form(D["NV"], D["v_dofs"], be)
    assert np.array_equal(b, D["bodyforce"]), np.abs(b - D["bodyforce"]).max()
    NF, ndf, nqf = (int(x) for x in D["t_meta"])
    rhs = O.add_traction_form(b.copy(), dim, D["t_v_dofs"].reshape(NF, -1), D["t_N"].reshape(NF, nqf, ndf),
                              D["t_w"].reshape(NF, nqf), D["t_detW"].reshape(NF, nqf), np.array([0.5, -0.25, 1.5])[:dim])
    assert np.array_equal(rhs, D["rhs_bf_traction"])
    if name != "square_nurbs_p2_n4":
        assert w.min() < 0.9        # genuinely rational
        r = np.hypot(D["x_ref"].reshape(-1, dim)[:, 0], D["x_ref"].reshape(-1, dim)[:, 1])
        assert r.max() > 2.0        # control points off the arc: the geometry is curved, not interpolated
