"""GPU: mass matrix, viscosity matrix and body-force vector assembled on the device. The viscosity
matrix (mb2_assemble_viscosity_device) is bit-identical to the reference's mfem::VectorDiffusionIntegrator. The body-force vector
(mb2_assemble_body_force_device) is bit-identical to the reference's mfem::VectorDomainLFIntegrator (tests/golden/forms.npz,
oracle/make_golden.py bodyforce). The mass matrix (mb2_assemble_mass_device) is BIT-identical to
what the compiled reference dumps after  mass->AddDomainIntegrator(new mfem::VectorMassIntegrator(rho));
mass->Assemble(0)  (src/mimi/py/py_nonlinear_solid.cpp:154-172; oracle/ref_harness.cpp dumps `mass` into every golden
fixture): own quadrature rule (2p + OrderW), shape values and geometry Jacobian regenerated from the knot vectors, points
then elements added in MFEM's order. Also: FormSystemMatrix's essential-dof elimination when the result becomes the mass
of the Newton sweeps, density scaling, adds-into semantics, and that the domain tables are untouched.
"""
import numpy as np
import pytest

from mimi_b200 import meshgen
from tests.cases import load_contact_golden, load_golden

pytestmark = pytest.mark.gpu

# fixture -> degree of the refined patch (oracle/make_golden.py: single tensor-product box patches, open uniform knots)
FIX = {"cube_p2_n2": 2, "cube_p1_n2": 1, "cube_p3_n2": 3, "cube_p2_n3_patch": 2, "quad_p2_n4": 2, "beam2d_p3_n2": 3,
       "balken_p3_n2": 3, "contact_quad_p3_n4_sphere": 3, "contact_cube_p2_n4_sphere": 2}


def patch_of(D, p):
    """knots / ncp / el_ijk of a box fixture, read off its own arrays: control points sit at the Greville abscissae, so
    the rank of a coordinate among the distinct ones is the control-point index, and local node 0 of an element is its
    lowest corner = its knot-span index."""
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


def setup(ctx, D):
    ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
    ctx.set_nodes(D["x_ref"], None)
    ctx.build_pattern()
    I, J = ctx.get_pattern()
    assert np.array_equal(I, D["I"]) and np.array_equal(J, D["J"])


@pytest.mark.parametrize("name", sorted(FIX))
def test_mass_bit_exact_vs_reference(ctx_factory, name):
    if name.startswith("contact_"):      # the contact fixtures carry the domain connectivity and the mass dump too
        D = load_contact_golden(name)
        D["NE"] = D["v_dofs"].size // (D["dim"] * D["nd"])
    else:
        D = load_golden(name)
    p = FIX[name]
    assert (p + 1) ** D["dim"] == D["nd"]
    ctx = ctx_factory()
    setup(ctx, D)
    ncp, knots, ijk = patch_of(D, p)
    ctx.set_A(None)
    ctx.set_r(None)
    ctx.assemble_mass_device(p, ncp, knots, ijk)
    M = ctx.get_A()
    assert np.abs(D["mass"]).max() > 0
    assert np.array_equal(M, D["mass"]), "max |dM| = %.3e" % np.abs(M - D["mass"]).max()
    assert not np.any(ctx.get_r())                      # the residual is untouched (r += 0)
    print("\n[mass %s] %d CSR values bit-equal to the reference's VectorMassIntegrator" % (name, M.size))


def test_mass_density_addsinto_and_total(ctx_factory):
    D = load_golden("cube_p2_n3_patch")
    ctx = ctx_factory()
    setup(ctx, D)
    ncp, knots, ijk = patch_of(D, 2)
    ctx.set_A(None)
    ctx.assemble_mass_device(2, ncp, knots, ijk, density=2.0)
    assert np.array_equal(ctx.get_A(), 2.0 * D["mass"])          # (w detJ) * rho, rho a power of two: exact
    ctx.assemble_mass_device(2, ncp, knots, ijk, density=2.0)     # adds into the resident values
    assert np.array_equal(ctx.get_A(), 4.0 * D["mass"])
    # partition of unity: the entries of one component block sum to the volume (1.5 x 1 x 1), dim blocks in total
    assert abs(D["mass"].sum() - 3 * 1.5) < 1e-12
    # a rule that integrates the integrand exactly gives the same matrix up to rounding
    ctx.set_A(None)
    ctx.assemble_mass_device(2, ncp, knots, ijk, quad_order=13)
    M13 = ctx.get_A()
    assert not np.array_equal(M13, D["mass"])
    assert np.allclose(M13, D["mass"], rtol=1e-12, atol=1e-16)


def test_mass_then_domain_assembly_unchanged(ctx_factory):
    """Mass assembly borrows the Jacobian pipeline: a domain assembly afterwards is the golden one (J = M + gf K)."""
    from oracle import oracle_py as O
    D = load_golden("cube_p2_n2")
    ctx = ctx_factory()
    setup(ctx, D)
    ctx.upload_quad_tables(D["dN_dX"], D["w"], D["detJ"])
    ctx.set_material(0, list(O.lame()))
    ncp, knots, ijk = patch_of(D, 2)
    ctx.set_A(None)
    ctx.set_r(None)
    ctx.assemble_mass_device(2, ncp, knots, ijk)
    ctx.set_u(D["u"])
    ctx.assemble_residual_and_grad_device(D["gf"])
    assert np.array_equal(ctx.get_r(), D["r2_neohookean"]) and np.array_equal(ctx.get_A(), D["A_neohookean"])


def test_mass_becomes_newton_mass_with_elimination(ctx_factory):
    """FormSystemMatrix(zero_dofs): the device result equals the host elimination tests/beam_driver.py applies to the
    reference's dump, and a Newton sweep that uses it reproduces the reference's (sweeps fixture)."""
    from tests.beam_driver import Beam
    D = load_golden("beam2d_p3_n2")
    ctx = ctx_factory()
    setup(ctx, D)
    ncp, knots, ijk = patch_of(D, 3)
    ctx.set_A(None)
    ctx.assemble_mass_device(3, ncp, knots, ijk)

    class _NoBackend:
        pass
    beam = Beam(D, _NoBackend(), {1: -1.0}, 0.1)                # only for its ess / eliminated mass
    ctx.set_essential_dofs(beam.ess)
    ctx.newton_mass_from_A(True)
    assert np.array_equal(ctx.newton_get_mass(), beam.mass)
    ctx.newton_mass_from_A(False)
    assert np.array_equal(ctx.newton_get_mass(), D["mass"])


def test_mass_rejects_bad_input(ctx_factory):
    from mimi_b200 import capi
    D = load_golden("cube_p2_n2")
    ctx = ctx_factory()
    ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
    ncp, knots, ijk = patch_of(D, 2)
    with pytest.raises(capi.Mb2Error, match="set_nodes"):
        ctx.assemble_mass_device(2, ncp, knots, ijk)
    ctx.set_nodes(D["x_ref"], None)
    with pytest.raises(capi.Mb2Error, match="build_pattern"):
        ctx.assemble_mass_device(2, ncp, knots, ijk)
    ctx.build_pattern()
    with pytest.raises(capi.Mb2Error, match="quadrature order"):
        ctx.assemble_mass_device(2, ncp, knots, ijk, quad_order=17)


# ---- body-force linear form -------------------------------------------------------------------------------------------
BF = {k: v for k, v in FIX.items() if not k.startswith("contact_")}


def bodyforce():
    import os
    from tests.cases import GOLDEN
    z = np.load(os.path.join(GOLDEN, "forms.npz"))
    return {k: z[k] for k in z.files}


@pytest.mark.parametrize("name", sorted(BF))
def test_body_force_bit_exact_vs_reference(ctx_factory, name):
    D = load_golden(name)
    G = bodyforce()
    p = BF[name]
    ctx = ctx_factory()
    setup(ctx, D)
    ncp, knots, ijk = patch_of(D, p)
    ctx.set_r(None)
    ctx.assemble_body_force_device(p, ncp, knots, ijk, G["f"][:D["dim"]])
    b = ctx.get_r()
    assert np.abs(G[name]).max() > 0
    assert np.array_equal(b, G[name]), "max |db| = %.3e" % np.abs(b - G[name]).max()
    # adds into r; partition of unity: component c sums to f_c x volume
    ctx.assemble_body_force_device(p, ncp, knots, ijk, G["f"][:D["dim"]])
    assert np.array_equal(ctx.get_r(), 2.0 * G[name])
    vol = float((D["w"] * D["detJ"]).sum())
    for c in range(D["dim"]):
        assert abs(b[c::D["dim"]].sum() - G["f"][c] * vol) < 1e-12 * max(1.0, vol)
    print("\n[body force %s] %d entries bit-equal to the reference's VectorDomainLFIntegrator" % (name, b.size))


@pytest.mark.parametrize("name", sorted(BF))
def test_rhs_body_force_plus_traction_bit_exact_vs_reference(ctx_factory, name):
    """The reference's whole rhs linear form (VectorDomainLFIntegrator + VectorBoundaryLFIntegrator in one
    LinearForm::Assemble, py_nonlinear_solid.cpp:221-282) on the device: body force through the residual gather, then
    mb2_assemble_traction_device on the loaded faces' tables (dumped from the reference's boundary precompute)."""
    from oracle import oracle_py as O
    D = load_golden(name)
    G = bodyforce()
    p, dim = BF[name], D["dim"]
    ctx = ctx_factory()
    setup(ctx, D)
    ncp, knots, ijk = patch_of(D, p)
    ctx.set_r(None)
    ctx.assemble_body_force_device(p, ncp, knots, ijk, G["f"][:dim])
    ctx.assemble_traction_device(G["tvd_" + name], G["tN_" + name], G["tw_" + name], G["tdet_" + name], G["traction"][:dim])
    rhs = ctx.get_r()
    assert np.array_equal(rhs, G["rhs_" + name]), "max |d rhs| = %.3e" % np.abs(rhs - G["rhs_" + name]).max()
    assert not np.array_equal(rhs, G[name])
    # traction alone on a zero vector == the oracle's
    ctx.set_r(None)
    ctx.assemble_traction_device(G["tvd_" + name], G["tN_" + name], G["tw_" + name], G["tdet_" + name], G["traction"][:dim])
    want = O.add_traction_form(np.zeros(D["NV"]), dim, G["tvd_" + name], G["tN_" + name], G["tw_" + name], G["tdet_" + name],
                               G["traction"][:dim])
    assert np.array_equal(ctx.get_r(), want)
    print("\n[rhs %s] body force + traction: %d entries bit-equal to the reference's LinearForm" % (name, rhs.size))


@pytest.mark.parametrize("name", sorted(BF))
def test_viscosity_bit_exact_vs_reference(ctx_factory, name):
    """mb2_assemble_viscosity_device against the reference's mfem::VectorDiffusionIntegrator matrix (forms.npz, visc_*)."""
    D = load_golden(name)
    G = bodyforce()
    p = BF[name]
    ctx = ctx_factory()
    setup(ctx, D)
    ncp, knots, ijk = patch_of(D, p)
    ctx.set_A(None)
    ctx.assemble_viscosity_device(p, ncp, knots, ijk, float(G["viscosity"]))
    S = ctx.get_A()
    ref = G["visc_" + name]
    assert np.abs(ref).max() > 0
    assert np.array_equal(S, ref), "max |dS| = %.3e" % np.abs(S - ref).max()
    # a diffusion matrix annihilates constants: every row sums to ~0 over each component block
    rows = np.repeat(np.arange(D["NV"]), np.diff(D["I"]))
    rs = np.zeros(D["NV"])
    np.add.at(rs, rows, S)
    assert np.abs(rs).max() < 1e-11 * np.abs(S).max()
    # FormSystemMatrix elimination when it becomes the viscosity of the Newton sweeps
    xr = D["x_ref"].reshape(-1, D["dim"])
    nodes = np.nonzero(xr[:, 0] == 0.0)[0]
    ess = np.sort(np.concatenate([D["dim"] * nodes + c for c in range(D["dim"])])).astype(np.int32)
    ctx.set_essential_dofs(ess)
    ctx.newton_viscosity_from_A(True)
    is_ess = np.zeros(D["NV"], bool)
    is_ess[ess] = True
    want = ref.copy()
    want[is_ess[rows] | is_ess[D["J"]]] = 0.0
    want[(rows == D["J"]) & is_ess[rows]] = 1.0
    assert np.array_equal(ctx.newton_get_viscosity(), want)
    print("\n[viscosity %s] %d CSR values bit-equal to the reference's VectorDiffusionIntegrator" % (name, S.size))


def test_body_force_becomes_newton_rhs_and_drives_the_beam(ctx_factory):
    """Mass and rhs of the reference's own beam test assembled on the device (nothing from MFEM but the mesh): the
    device Newton loop reproduces the reference's committed NeoHookean history."""
    from oracle import oracle_py as O
    from tests.beam_driver import DeviceNewtonBeam
    from tests.test_reference_histories import DeviceBackend, SETUP, histories
    D = load_golden("balken_p3_n2")
    force, dt = SETUP["neohook"]
    ctx = ctx_factory()
    beam = DeviceNewtonBeam(D, DeviceBackend(D, "neohook", ctx), ctx, {1: force}, dt)
    ncp, knots, ijk = patch_of(D, 3)
    ctx.set_nodes(D["x_ref"], None)
    ctx.set_A(None)
    ctx.assemble_mass_device(3, ncp, knots, ijk)
    ctx.newton_mass_from_A(True)              # essential dofs were set by DeviceNewtonBeam
    assert np.array_equal(ctx.newton_get_mass(), beam.mass)
    ctx.set_r(None)
    ctx.assemble_body_force_device(3, ncp, knots, ijk, [0.0, force])
    ctx.newton_rhs_from_r(True)
    b = ctx.get_r()
    b[beam.ess] = 0.0
    assert np.allclose(b, beam.rhs, rtol=1e-13, atol=1e-15)     # beam_driver's numpy rhs uses the domain rule
    beam.rhs = b                                                # the host-side a0 solve uses the same vector
    ref = histories()["neohook"]
    for i in range(10):
        assert np.allclose(beam.step(), ref[i]), i               # the reference's own assertion
    _ = O


def test_forms_with_nurbs_weights_vs_oracle(ctx_factory):
    """Rational basis (NURBS weights != 1; no reference fixture has them): device == CPU oracle bit for bit, the oracle
    itself being pinned against the reference on the unit-weight fixtures (tests/test_oracle_forms.py)."""
    from oracle import oracle_py as O
    D = load_golden("cube_p2_n2")
    G = bodyforce()
    dim, nd, p = D["dim"], D["nd"], 2
    w = np.random.default_rng(11).uniform(0.7, 1.3, D["NV"] // dim)
    ctx = ctx_factory()
    ctx.set_connectivity(dim, nd, D["nq"], D["NV"], D["v_dofs"])
    ctx.set_nodes(D["x_ref"], w)
    ctx.build_pattern()
    ncp, knots, ijk = patch_of(D, p)
    for kind, coef, run in ((0, 1.7, lambda: ctx.assemble_mass_device(p, ncp, knots, ijk, density=1.7)),
                            (2, 0.37, lambda: ctx.assemble_viscosity_device(p, ncp, knots, ijk, 0.37))):
        ctx.set_A(None)
        run()
        want = O.assemble_bilinear_form(dim, nd, D["v_dofs"], D["I"], D["J"],
                                        O.form_element_matrices(kind, dim, p, D["v_dofs"], ijk, knots, D["x_ref"], coef=coef,
                                                                weights=w))
        assert np.array_equal(ctx.get_A(), want), kind
        assert not np.array_equal(want, D["mass"])
    ctx.set_r(None)
    ctx.assemble_body_force_device(p, ncp, knots, ijk, G["f"])
    want = O.assemble_linear_form(D["NV"], D["v_dofs"], O.form_element_vectors(dim, p, D["v_dofs"], ijk, knots, D["x_ref"],
                                                                               G["f"], weights=w))
    assert np.array_equal(ctx.get_r(), want)
