"""GPU: multi-patch NURBS meshes (BASELINE config 4: patches glued along x). The fixtures are dumps of the unmodified
reference on 4 x (2 x 1 x 1) p = 3 and 3 x (2 x 2 x 2) p = 2 beams (tests/golden/beam4_p3.npz, beam3_p2.npz): MFEM numbers
the interface nodes once, PrepareSparsity / A_ids run across the interfaces (precomputed.cpp:39-218). Pattern, index map,
residual and Jacobian on uploaded tables are covered by tests/test_gpu_parity.py::test_golden (both names are in
GOLDEN_CASES) and the partitioned runs by tests/test_gpu_partition.py; here: the tables REGENERATED on the device patch by
patch (mb2_generate_patch_tables once per patch into one ascending element range) and the assembly that consumes them."""
import numpy as np
import pytest

from tests.cases import MATS, MULTIPATCH_CASES, load_golden, patches_of

pytestmark = pytest.mark.gpu


def generated(ctx, D):
    ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
    ctx.set_nodes(D["x_ref"], D["node_w"])
    for (e0, e1), ncp, knots in patches_of(D):
        ctx.generate_patch_tables(int(D["p"]), ncp, knots, D["el_ijk"][e0:e1], e0=e0, want_N=True)
    return ctx


@pytest.mark.parametrize("name", MULTIPATCH_CASES)
def test_tables_generated_patch_by_patch_bit_exact(ctx_factory, name):
    D = load_golden(name)
    assert int(D["n_patches"]) >= 3
    ctx = generated(ctx_factory(), D)
    dN, wd = ctx.download_quad_tables()
    assert np.array_equal(wd, D["w"] * D["detJ"])
    assert np.array_equal(dN.reshape(D["dN_dX"].shape), D["dN_dX"])
    assert np.array_equal(ctx.download_shape_values().reshape(D["N"].shape), D["N"])


@pytest.mark.parametrize("material", ["neohookean", "stvk"])
@pytest.mark.parametrize("name", MULTIPATCH_CASES)
def test_multipatch_assembly_from_generated_tables(ctx_factory, name, material):
    D = load_golden(name)
    ctx = generated(ctx_factory(), D)
    assert ctx.build_pattern() == D["nnz"]
    I, J = ctx.get_pattern()
    assert np.array_equal(I, D["I"]) and np.array_equal(J, D["J"])          # rows of interface nodes see both patches
    assert np.array_equal(ctx.get_A_ids(), D["A_ids"])
    ctx.set_material(MATS[material], list(D["lambda_mu"]))
    r = np.zeros(D["NV"])
    A = D["mass"].copy()
    ctx.add_domain_residual_and_grad(D["u"], D["gf"], r, A)
    assert np.array_equal(r, D["r2_" + material]) and np.array_equal(A, D["A_" + material])
    # an interface node really couples two patches: its row holds more neighbours than an interior node of one patch
    # than one patch has nodes (beam3_p2: 80 > 4 * 4 * 4)
    nn = np.diff(I)[::3] // 3
    assert nn.max() > max(int(np.prod(ncp)) for _, ncp, _ in patches_of(D))
