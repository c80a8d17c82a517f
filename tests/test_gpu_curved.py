"""GPU: rational / curved / non-affine NURBS geometry pinned against the compiled
reference: exact quarter annulus and thick quarter cylinder (weights 1, sqrt(1/2), 1) and the reference's own non-affine
tests/data/square-nurbs.mesh (tests/golden/{annulus_p2_n2,cylinder_p2_n2,square_nurbs_p2_n4}.npz, oracle/make_golden.py
curved_main). The device regenerates the quadrature tables from the patch MFEM itself holds after refinement
(mb2_generate_patch_tables: rational basis with the quotient rule, non-constant Jacobian inverse) and assembles the four
setup forms; everything must equal PrecomputeElementQuadData / MFEM's integrators BIT for bit."""
import numpy as np
import pytest

from tests.cases import CURVED_CASES, MATS, load_golden, patch_description

pytestmark = pytest.mark.gpu


def generated(ctx, D):
    p, ncp, knots, ijk, w = patch_description(D)
    ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
    ctx.set_nodes(D["x_ref"], w)
    ctx.generate_patch_tables(p, ncp, knots, ijk, want_N=True)
    return p, ncp, knots, ijk


@pytest.mark.parametrize("name", CURVED_CASES)
def test_generated_tables_bit_exact_on_rational_and_curved_patches(ctx_factory, name):
    D = load_golden(name)
    ctx = ctx_factory()
    generated(ctx, D)
    dN, wd = ctx.download_quad_tables()
    assert np.array_equal(wd, D["w"] * D["detJ"])
    assert np.array_equal(dN.reshape(D["dN_dX"].shape), D["dN_dX"])
    assert np.array_equal(ctx.download_shape_values().reshape(D["N"].shape), D["N"])
    # the geometry really is non-affine: det J varies inside an element
    dj = D["detJ"].reshape(D["NE"], -1)
    assert (dj.max(axis=1) / dj.min(axis=1)).max() > 1.05


@pytest.mark.parametrize("name", CURVED_CASES)
def test_assembly_from_generated_tables_on_curved_patches(ctx_factory, name):
    D = load_golden(name)
    ctx = ctx_factory()
    generated(ctx, D)
    assert ctx.build_pattern() == D["nnz"]
    I, J = ctx.get_pattern()
    assert np.array_equal(I, D["I"]) and np.array_equal(J, D["J"]) and np.array_equal(ctx.get_A_ids(), D["A_ids"])
    for material in ("neohookean", "stvk"):
        ctx.set_material(MATS[material], list(D["lambda_mu"]))
        r = np.zeros(D["NV"])
        A = D["mass"].copy()
        ctx.add_domain_residual_and_grad(D["u"], D["gf"], r, A)
        assert np.array_equal(r, D["r2_" + material]) and np.array_equal(A, D["A_" + material])
        r1 = np.zeros(D["NV"])
        ctx.add_domain_residual(D["u"], r1)
        assert np.array_equal(r1, D["r_" + material])


@pytest.mark.parametrize("name", CURVED_CASES)
def test_setup_forms_bit_exact_on_curved_patches(ctx_factory, name):
    """mass (VectorMassIntegrator), viscosity (VectorDiffusionIntegrator, 0.37), body force (VectorDomainLFIntegrator) and
    body force + traction (the reference's whole rhs LinearForm) on the device."""
    D = load_golden(name)
    dim = D["dim"]
    ctx = ctx_factory()
    p, ncp, knots, ijk = generated(ctx, D)
    ctx.build_pattern()
    ctx.set_A(None)
    ctx.assemble_mass_device(p, ncp, knots, ijk)
    assert np.array_equal(ctx.get_A(), D["mass"]), np.abs(ctx.get_A() - D["mass"]).max()
    ctx.set_A(None)
    ctx.assemble_viscosity_device(p, ncp, knots, ijk, 0.37)
    assert np.array_equal(ctx.get_A(), D["viscform"]), np.abs(ctx.get_A() - D["viscform"]).max()
    f = np.array([0.3, -1.7, 0.9])[:dim]
    ctx.set_r(None)
    ctx.assemble_body_force_device(p, ncp, knots, ijk, f)
    assert np.array_equal(ctx.get_r(), D["bodyforce"])
    NF, ndf, nqf = (int(x) for x in D["t_meta"])
    ctx.assemble_traction_device(D["t_v_dofs"].reshape(NF, -1), D["t_N"].reshape(NF, nqf, ndf), D["t_w"].reshape(NF, nqf),
                                 D["t_detW"].reshape(NF, nqf), np.array([0.5, -0.25, 1.5])[:dim])
    assert np.array_equal(ctx.get_r(), D["rhs_bf_traction"])
    vol = float((D["w"] * D["detJ"]).sum())
    print("\n[%s] tables, mass, viscosity, body force, rhs bit-equal to MFEM on %d elements; area / volume %.12f" % (
        name, D["NE"], vol))
    if name == "annulus_p2_n2":
        assert abs(vol - 0.75 * np.pi) < 1e-6        # quarter annulus r in [1, 2]: exact geometry, Gauss rule on a rational map
