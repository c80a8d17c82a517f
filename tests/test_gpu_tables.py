"""GPU: device-regenerated quadrature tables (mb2_generate_patch_tables) are BIT-identical to the tables the
unmodified reference precomputes (PrecomputeElementQuadData, src/mimi/utils/precomputed.cpp:264-330), on the
golden fixtures and, where oracle/_ref travelled, on fresh reference dumps; and the assembly that consumes them
reproduces the reference residual / Jacobian bit for bit."""
import numpy as np
import pytest

from mimi_b200 import meshgen
from oracle import oracle_py as O
from tests.cases import MATS, load_golden, ref_case

pytestmark = pytest.mark.gpu

# fixture name -> (p, spans, lengths): how oracle/make_golden.py built the mesh
FIX = {"cube_p2_n2": (2, (2, 2, 2), (1.0, 1.0, 1.0)), "cube_p1_n2": (1, (2, 2, 2), (1.0, 1.0, 1.0)),
       "cube_p2_n3_patch": (2, (3, 2, 2), (1.5, 1.0, 1.0))}


def _gen(ctx, D, p, spans, lengths):
    M = meshgen.patch_connectivity(p, spans, lengths)
    assert np.array_equal(M["v_dofs"], D["v_dofs"])          # MFEM numbering reproduced
    assert np.array_equal(M["x_ref"], D["x_ref"])
    ctx.set_connectivity(3, D["nd"], D["nq"], D["NV"], M["v_dofs"])
    ctx.set_nodes(M["x_ref"], M["weights"])
    ctx.generate_patch_tables(p, M["ncp"], M["knots"], M["el_ijk"], want_N=True)
    return M


@pytest.mark.parametrize("name", sorted(FIX))
def test_tables_bit_exact_vs_golden(ctx_factory, name):
    D = load_golden(name)
    ctx = ctx_factory()
    _gen(ctx, D, *FIX[name])
    dN, wd = ctx.download_quad_tables()
    assert np.array_equal(wd, D["w"] * D["detJ"])
    assert np.array_equal(dN, D["dN_dX"])
    assert np.array_equal(ctx.download_shape_values(), D["N"])


@pytest.mark.parametrize("name", sorted(FIX))
def test_assembly_from_generated_tables(ctx_factory, name):
    D = load_golden(name)
    ctx = ctx_factory()
    _gen(ctx, D, *FIX[name])
    assert ctx.build_pattern() == D["nnz"]
    for material in ("neohookean", "stvk"):
        ctx.set_material(MATS[material], list(D["lambda_mu"]))
        r = np.zeros(D["NV"])
        A = D["mass"].copy()
        ctx.add_domain_residual_and_grad(D["u"], D["gf"], r, A)
        assert np.array_equal(r, D["r2_" + material]) and np.array_equal(A, D["A_" + material])


@pytest.mark.parametrize("p,spans", [(2, (3, 4, 5)), (3, (2, 3, 2))])
def test_tables_vs_live_reference(ctx_factory, p, spans):
    if not O.have_ref():
        pytest.skip("oracle/_ref/mimi_ref not present on this box")
    import os, tempfile
    tmp = tempfile.mkdtemp(prefix="mb2tab_")
    mesh = os.path.join(tmp, "m.mesh")
    L = (1.0, 2.0, 0.5)
    meshgen.write_box_patch(mesh, p, spans, L)
    D = O.run_ref_dump(mesh, 0, 0, "stvk", os.path.join(tmp, "d"), gf=1.0, lite=True)
    ctx = ctx_factory()
    _gen(ctx, D, p, spans, L)
    dN, wd = ctx.download_quad_tables()
    assert np.array_equal(wd, D["w"] * D["detJ"]) and np.array_equal(dN, D["dN_dX"])


def test_generate_rejects_bad_input(ctx_factory):
    from mimi_b200 import capi
    D = load_golden("cube_p2_n2")
    ctx = ctx_factory()
    M = meshgen.patch_connectivity(2, (2, 2, 2))
    ctx.set_connectivity(3, 27, 64, D["NV"], M["v_dofs"])
    with pytest.raises(capi.Mb2Error, match="set_nodes"):
        ctx.generate_patch_tables(2, M["ncp"], M["knots"], M["el_ijk"])
    ctx.set_nodes(M["x_ref"])
    with pytest.raises((capi.Mb2Error, AssertionError)):
        ctx.generate_patch_tables(3, M["ncp"], M["knots"], M["el_ijk"])   # degree does not match nd
    bad = M["x_ref"].copy().reshape(-1, 3)
    bad[:, 0] *= -1.0                                                   # inverted geometry -> det J < 0
    ctx.set_nodes(bad.ravel())
    with pytest.raises(capi.Mb2Error, match="Jacobian"):
        ctx.generate_patch_tables(2, M["ncp"], M["knots"], M["el_ijk"])
