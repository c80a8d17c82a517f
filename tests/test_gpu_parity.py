"""GPU: the CUDA path through the C ABI against (a) the golden vectors dumped from the unmodified
reference, (b) the plain-C oracle, (c) live runs of the compiled reference where oracle/_ref travelled.
Bar: bit-exact pattern / indices AND bit-exact FP64 residual and Jacobian (the FD tangent is only
reproducible to 1e-12 by reproducing every bit; north_star's tolerance is 1e-12 relative)."""
import numpy as np
import pytest

from oracle import oracle_py as O
from tests.cases import GOLDEN_CASES, MATS, load_golden, ref_case, setup_ctx

pytestmark = pytest.mark.gpu


def _check_case(ctx, D, material, r_ref, r2_ref, A_ref):
    setup_ctx(ctx, D, material)
    # pattern + DOF indexing: bit-exact
    I, J = ctx.get_pattern()
    assert ctx.nnz == D["nnz"]
    assert np.array_equal(I, D["I"]) and np.array_equal(J, D["J"])
    if D.get("A_ids") is not None:
        assert np.array_equal(ctx.get_A_ids(), D["A_ids"])
    # residual only (AddDomainResidual)
    r = np.zeros(D["NV"])
    ctx.add_domain_residual(D["u"], r)
    assert np.array_equal(r, r_ref), np.abs(r - r_ref).max()
    # residual + tangent (AddDomainResidualAndGrad), A pre-loaded with the mass
    r2 = np.zeros(D["NV"])
    A = D["mass"].copy()
    ctx.add_domain_residual_and_grad(D["u"], D["gf"], r2, A)
    assert np.array_equal(r2, r2_ref)
    rel = np.abs(A - A_ref).max() / np.abs(A_ref).max()
    assert rel <= 1e-12, rel          # north_star tolerance
    assert np.array_equal(A, A_ref)   # and in fact bit-exact
    # "adds into": a second call on top of non-zero r
    r3 = r2.copy()
    ctx.add_domain_residual(D["u"], r3)
    assert np.array_equal(r3, r2 + r_ref)


@pytest.mark.parametrize("material", ["neohookean", "stvk"])
@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_golden(ctx_factory, name, material):
    D = load_golden(name)
    _check_case(ctx_factory(), D, material, D["r_" + material], D["r2_" + material], D["A_" + material])


@pytest.mark.parametrize("kind,p,n,material", [("refine", 2, 4, "neohookean"), ("patch", 2, 5, "stvk"),
                                               ("refine", 2, 8, "neohookean")])
def test_live_reference_or_oracle(ctx_factory, kind, p, n, material):
    """Larger meshes: the compiled reference when it travelled to this box, else the C oracle on
    tables from the golden-validated path is not possible (tables come from the reference) -> skip."""
    if not O.have_ref():
        pytest.skip("oracle/_ref/mimi_ref not present on this box")
    D = ref_case(kind, p, n, material, gf=0.3)
    D["A_ids"] = None
    _check_case(ctx_factory(), D, material, D["r"], D["r2"], D["A"])


def test_dirichlet_zero_dofs_take_the_1e10_branch(ctx_factory):
    """u == 0 exactly on some dofs (h = 1e-10 branch, nonlinear_solid.cpp:65-66) and u == 0 everywhere
    (first Newton iteration). Checked against the C oracle."""
    D = load_golden("cube_p2_n2")
    m = O.mat(0)
    ctx = setup_ctx(ctx_factory(), D, "neohookean")
    for u in (np.where(D["x_ref"].reshape(-1, 3)[:, :1] == 0.0, 0.0, D["u"].reshape(-1, 3)).ravel(),
              np.zeros(D["NV"])):
        u = np.ascontiguousarray(u)
        r_o, A_o = O.assemble_residual_and_grad(3, 27, 64, m, D["v_dofs"], D["A_ids"], D["dN_dX"], D["w"], D["detJ"],
                                                u, 1.0, D["mass"])
        r = np.zeros(D["NV"])
        A = D["mass"].copy()
        ctx.add_domain_residual_and_grad(u, 1.0, r, A)
        assert np.array_equal(r, r_o) and np.array_equal(A, A_o)


def test_rerun_is_deterministic(ctx_factory):
    D = load_golden("cube_p2_n3_patch")
    ctx = setup_ctx(ctx_factory(), D, "neohookean")
    outs = []
    for _ in range(3):
        r = np.zeros(D["NV"])
        A = np.zeros(D["nnz"])
        ctx.add_domain_residual_and_grad(D["u"], 1.0, r, A)
        outs.append((r, A))
    assert all(np.array_equal(outs[0][0], o[0]) and np.array_equal(outs[0][1], o[1]) for o in outs[1:])


def test_errors_are_loud(ctx_factory):
    from mimi_b200 import capi
    ctx = ctx_factory()
    with pytest.raises(capi.Mb2Error):
        ctx.build_pattern()  # nothing set
    D = load_golden("cube_p1_n2")
    bad = D["v_dofs"].copy()
    bad[0, 0], bad[0, 1] = bad[0, 1], bad[0, 0]  # break component-major / byVDIM consistency
    ctx.set_connectivity(3, 8, 27, D["NV"], bad)
    with pytest.raises(capi.Mb2Error, match="byVDIM"):
        ctx.build_pattern()


@pytest.mark.parametrize("name,chunk", [("cube_p2_n3_patch", 5), ("cube_p2_n3_patch", 1), ("cube_p2_n2", 3), ("beam2d_p3_n2", 1)])
def test_chunked_pipeline_is_bit_exact(ctx_factory, name, chunk):
    """Elements processed in chunks (nodes straddling chunks carry their partial sums): identical bits, through
    the host-buffer entry (rows streamed per chunk) and the device-resident entry."""
    D = load_golden(name)
    ctx = ctx_factory()
    ctx.set_option("chunk_elems", chunk)
    setup_ctx(ctx, D, "neohookean")
    ec, nch = ctx.plan_info()
    assert ec == chunk and nch == -(-D["NE"] // chunk)
    r = np.zeros(D["NV"])
    A = D["mass"].copy()
    ctx.add_domain_residual_and_grad(D["u"], D["gf"], r, A)
    assert np.array_equal(r, D["r2_neohookean"]) and np.array_equal(A, D["A_neohookean"])
    ctx.set_u(D["u"]); ctx.set_r(None); ctx.set_A(D["mass"])
    ctx.assemble_residual_and_grad_device(D["gf"])
    assert np.array_equal(ctx.get_r(), D["r2_neohookean"]) and np.array_equal(ctx.get_A(), D["A_neohookean"])
    assert ctx.last_launch_count() == 2 * nch + 1
    t = ctx.last_timing_ms()
    assert all(x >= 0 for x in t)


@pytest.mark.parametrize("variant", [1])
@pytest.mark.parametrize("material", ["neohookean", "stvk"])
@pytest.mark.parametrize("name", ["cube_p2_n2", "cube_p2_n3_patch"])
def test_every_element_kernel_variant_is_bit_exact(ctx_factory, name, material, variant):
    """mb2_set_option("elem_variant"): 1 = the generic one-thread-per-perturbed-dof kernel instead of the two-phase
    kernel. Same bits from both, also with zero dofs (h = 1e-10 branch) in the state."""
    D = load_golden(name)
    ctx = ctx_factory()
    ctx.set_option("elem_variant", variant)
    setup_ctx(ctx, D, material)
    r = np.zeros(D["NV"])
    A = D["mass"].copy()
    ctx.add_domain_residual_and_grad(D["u"], D["gf"], r, A)
    assert np.array_equal(r, D["r2_" + material]) and np.array_equal(A, D["A_" + material])
    # clamped face: exact zeros in u; compared with the default kernel of another context
    u0 = np.ascontiguousarray(np.where(D["x_ref"].reshape(-1, 3)[:, :1] == 0.0, 0.0, D["u"].reshape(-1, 3)).ravel())
    ref = setup_ctx(ctx_factory(), D, material)
    ra, Aa, rb, Ab = np.zeros(D["NV"]), np.zeros(D["nnz"]), np.zeros(D["NV"]), np.zeros(D["nnz"])
    ctx.add_domain_residual_and_grad(u0, 1.0, ra, Aa)
    ref.add_domain_residual_and_grad(u0, 1.0, rb, Ab)
    assert np.array_equal(ra, rb) and np.array_equal(Aa, Ab)


def test_mass_resident_entry_equals_the_adds_into_entry(ctx_factory):
    """mb2_domain_residual_and_grad_on_mass (mass uploaded once, CSR only downloaded) == mb2_add_domain_residual_and_grad
    on an A preloaded with the mass == the reference's dump."""
    D = load_golden("cube_p2_n3_patch")
    ctx = setup_ctx(ctx_factory(), D, "neohookean")
    ctx.set_mass(D["mass"])
    for _ in range(2):
        r = np.zeros(D["NV"])
        A = np.full(D["nnz"], np.nan)   # every entry must be overwritten
        ctx.domain_residual_and_grad_on_mass(D["u"], D["gf"], r, A)
        assert np.array_equal(r, D["r2_neohookean"]) and np.array_equal(A, D["A_neohookean"])


@pytest.mark.parametrize("material", ["neohookean", "stvk", "j2linear"])
@pytest.mark.parametrize("name", ["cube_p2_n3_patch", "beam3_p2", "cylinder_p2_n2", "cube_p3_n2", "beam4_p3"])
def test_both_residual_only_kernels_are_bit_exact(ctx_factory, name, material):
    """mb2_set_option("res_variant"): 0 = element_res2_kernel (p = 2: one thread per quadrature point for F + stress, one warp
    with three accumulators per lane for R) / element_res3_kernel (p = 3: the table streamed in chunks of 16 points by
    row-wise bulk copies), 1 = the first form (p = 2) / the generic kernel (p = 3). Same bits from both = the reference's
    residual."""
    D = load_golden(name)
    for variant in (0, 1):
        ctx = ctx_factory()
        ctx.set_option("res_variant", variant)
        setup_ctx(ctx, D, material)
        r = np.zeros(D["NV"])
        ctx.add_domain_residual(D["u"], r)
        assert np.array_equal(r, D["r_" + material])
