"""GPU: the partitioned assembly. Several ranks are emulated on ONE GPU (one context per rank, run one after the
other, the packed element matrices moved by the test through host memory = the library's manual transport), so the
round-end single-GPU run covers everything but the NCCL calls; tests/mg_nccl_check.py covers those on >= 2 GPUs.
The union of the rows the ranks own must equal the reference's single-thread assembly bit for bit."""
import numpy as np
import pytest

from mimi_b200 import partition as P
from oracle import oracle_py as O
from tests.cases import MATS, load_golden

pytestmark = pytest.mark.gpu


def _rank_ctx(ctx, D, R, material, chunk):
    if chunk:
        ctx.set_option("chunk_elems", chunk)
    loc = R["local_elems"]
    ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], R["v_dofs"])
    ctx.upload_quad_tables(D["dN_dX"][loc], D["w"][loc], D["detJ"][loc])
    ctx.mg_setup(R, manual=True)
    ctx.build_pattern()
    ctx.set_material(MATS[material], list(D["lambda_mu"]))
    return ctx


@pytest.mark.parametrize("name,world,chunk", [("cube_p2_n3_patch", 2, 0), ("cube_p2_n3_patch", 3, 2), ("cube_p2_n2", 2, 1),
                                              ("beam2d_p3_n2", 2, 0),
                                              # multi-patch beams: the cuts fall ON patch interfaces (4 ranks / 4 patches), INSIDE
                                              # patches (2 and 3 ranks over 3 patches of 8 elements) -- ghosts cross both
                                              ("beam4_p3", 4, 0), ("beam4_p3", 2, 0), ("beam3_p2", 2, 0), ("beam3_p2", 4, 5),
                                              ("beam3_p2", 3, 0)])
@pytest.mark.parametrize("grad", [True, False])
def test_emulated_ranks_reproduce_reference(ctx_factory, name, world, chunk, grad, material="neohookean"):
    D = load_golden(name)
    dim = D["dim"]
    ranges = P.element_ranges(D["NE"], world)
    ranks = [P.rank_problem(D["v_dofs"], dim, D["NV"], ranges, r) for r in range(world)]
    ctxs = [_rank_ctx(ctx_factory(), D, R, material, chunk) for R in ranks]
    gf = D["gf"]
    # global CSR offsets of every rank's (compact) row block
    for ctx, R in zip(ctxs, ranks):
        ctx.set_u(D["u"]); ctx.set_r(None)
        I_loc, J_loc = ctx.get_pattern()
        rows = np.repeat(R["owned"], dim)
        lens = np.diff(I_loc.astype(np.int64))
        assert (lens[~rows] == 0).all() and np.array_equal(lens[rows], np.diff(D["I"].astype(np.int64))[rows])
        R["_lens"] = lens
        # scatter the global mass values of the owned rows into the local value array
        gI = D["I"].astype(np.int64)
        idx = np.concatenate([np.arange(gI[v], gI[v + 1]) for v in np.nonzero(rows)[0]]) if rows.any() else np.zeros(0, int)
        R["_gidx"] = idx
        assert np.array_equal(J_loc, D["J"][idx])
        ctx.set_A(np.ascontiguousarray(D["mass"][idx]))
        ctx.mg_begin(gf, grad)
    for r, (ctx, R) in enumerate(zip(ctxs, ranks)):       # "send"
        for peer in R["send"]:
            K, Re = ctx.mg_export(peer, grad)
            ctxs[peer].mg_import(r, K, Re)
    A_glob = np.full(D["nnz"], np.nan); r_glob = np.full(D["NV"], np.nan)
    for ctx, R in zip(ctxs, ranks):
        ctx.mg_finish(gf, grad)
        rows = np.repeat(R["owned"], dim)
        r_glob[rows] = ctx.get_r()[rows]
        if grad:
            A_glob[R["_gidx"]] = ctx.get_A()
    if grad:
        assert np.array_equal(r_glob, D["r2_" + material]) and np.array_equal(A_glob, D["A_" + material])
    else:
        assert np.array_equal(r_glob, D["r_" + material])


@pytest.mark.parametrize("world", [2, 4, 8])
def test_config4_element_type_partitioned_across_patches(ctx_factory, world):
    """BASELINE config 4 in small: 4-patch p = 3 StVenantKirchhoff beam, element-partitioned over 2 / 4 / 8 emulated ranks
    (8 ranks = one element each: every patch interface and every in-patch cut carries ghosts), bit-identical to the
    reference's single-thread assembly."""
    test_emulated_ranks_reproduce_reference(ctx_factory, "beam4_p3", world, 0, True, material="stvk")


def test_partition_call_order_is_checked(ctx_factory):
    from mimi_b200 import capi
    D = load_golden("cube_p2_n2")
    R = P.rank_problem(D["v_dofs"], 3, D["NV"], P.element_ranges(D["NE"], 2), 0)
    ctx = ctx_factory()
    ctx.set_connectivity(3, 27, 64, D["NV"], R["v_dofs"])
    ctx.build_pattern()
    with pytest.raises(capi.Mb2Error, match="precede"):
        ctx.set_partition(R["n_compute"], R["owned"])


@pytest.mark.parametrize("name,world,material", [("cube_p2_n3_patch", 3, "neohookean"), ("beam4_p3", 4, "stvk"), ("beam3_p2", 2, "neohookean")])
def test_one_process_group_of_contexts(ctx_factory, name, world, material):
    """The C++ route to several GPUs: mb2_plan_partition + mb2_plan_apply per context, mb2_mg_group_assemble moving the
    packed element matrices with cudaMemcpyPeerAsync (here: all contexts on device 0, the copy engine path is the same).
    Two assemblies back to back without a host sync; owned rows bit-identical to the reference."""
    from mimi_b200 import capi
    D = load_golden(name)
    dim = D["dim"]
    plans = [capi.Plan(D["v_dofs"], dim, D["NV"], world, r) for r in range(world)]
    ctxs = []
    for pl in plans:
        ctx = ctx_factory()
        pl.apply(ctx, D["nq"])
        loc = pl.local_elems
        ctx.upload_quad_tables(D["dN_dX"][loc], D["w"][loc], D["detJ"][loc])
        ctx.build_pattern()
        ctx.set_material(MATS[material], list(D["lambda_mu"]))
        ctxs.append(ctx)
    arr = capi.group_init(ctxs)
    gI = D["I"].astype(np.int64)
    idxs = []
    for ctx, pl in zip(ctxs, plans):
        rows = np.repeat(pl.owned.astype(bool), dim)
        idx = np.concatenate([np.arange(gI[v], gI[v + 1]) for v in np.nonzero(rows)[0]])
        idxs.append((rows, idx))
        ctx.set_u(D["u"]); ctx.set_r(None); ctx.set_A(np.ascontiguousarray(D["mass"][idx]))
    capi.group_assemble(ctxs, arr, D["gf"], True)
    A_glob = np.full(D["nnz"], np.nan); r_glob = np.full(D["NV"], np.nan)
    for ctx, (rows, idx) in zip(ctxs, idxs):
        r_glob[rows] = ctx.get_r()[rows]
        A_glob[idx] = ctx.get_A()
    assert np.array_equal(r_glob, D["r2_" + material]) and np.array_equal(A_glob, D["A_" + material])
    # second and third assembly queued without a host sync in between: A = mass + 3 gf K up to the order of additions
    capi.group_assemble(ctxs, arr, D["gf"], True)
    capi.group_assemble(ctxs, arr, D["gf"], True)
    K = D["A_" + material] - D["mass"]
    for ctx, (rows, idx) in zip(ctxs, idxs):
        assert np.allclose(ctx.get_A(), D["mass"][idx] + 3.0 * K[idx], rtol=1e-12, atol=1e-12 * np.abs(K).max())
    # residual only
    for ctx in ctxs:
        ctx.set_r(None)
    capi.group_assemble(ctxs, arr, 0.0, False)
    for ctx, (rows, idx) in zip(ctxs, idxs):
        r_glob[rows] = ctx.get_r()[rows]
    assert np.array_equal(r_glob, D["r_" + material])
    for pl in plans:
        pl.close()
