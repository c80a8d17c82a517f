"""CPU, world_size 2 over gloo: the multi-GPU host logic (mimi_b200/partition.py). Each rank computes the element
matrices of its own range with the C oracle, sends the raw matrices of its send-set to the lower rank, and sums
own + ghost contributions in ascending element id into the rows of the nodes it owns. The union of the owned rows
must equal the reference's single-thread assembly bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mimi_b200 import partition as P
from oracle import oracle_py as O
from tests.cases import load_golden


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    D = load_golden(name)
    dim, nd, nq, tdof = D["dim"], D["nd"], D["nq"], D["tdof"]
    ranges = P.element_ranges(D["NE"], world)
    R = P.rank_problem(D["v_dofs"], dim, D["NV"], ranges, rank)
    m = O.mat(0)
    own = np.arange(R["e0"], R["e1"])
    Re, Ke = O.element_matrices(dim, nd, nq, m, D["v_dofs"], D["dN_dX"], D["w"], D["detJ"], D["u"], own)
    n_local = len(R["local_elems"])
    Re_l = np.zeros((n_local, tdof)); Ke_l = np.zeros((n_local, tdof, tdof))
    Re_l[:len(own)] = Re; Ke_l[:len(own)] = Ke
    ops, keep = [], []
    for peer, idx in R["send"].items():
        for buf in (torch.from_numpy(np.ascontiguousarray(Ke[idx])), torch.from_numpy(np.ascontiguousarray(Re[idx]))):
            keep.append(buf); ops.append(dist.P2POp(dist.isend, buf, peer))
    recv_bufs = []
    for peer, idx in R["recv"].items():
        kb, rb = torch.zeros(len(idx), tdof, tdof, dtype=torch.float64), torch.zeros(len(idx), tdof, dtype=torch.float64)
        recv_bufs.append((idx, kb, rb)); ops += [dist.P2POp(dist.irecv, kb, peer), dist.P2POp(dist.irecv, rb, peer)]
    if ops:
        for w_ in dist.batch_isend_irecv(ops):
            w_.wait()
    for idx, kb, rb in recv_bufs:
        Ke_l[idx] = kb.numpy(); Re_l[idx] = rb.numpy()
    # ordered gather into owned rows (reference: thread-local sum from 0.0, then A += gf * A_tl)
    owned_v = np.repeat(R["owned"], dim)
    Atl = np.zeros(D["nnz"]); rtl = np.zeros(D["NV"])
    for x, e in enumerate(R["local_elems"]):
        vd = D["v_dofs"][e]
        rows_owned = owned_v[vd]
        np.add.at(rtl, vd[rows_owned], Re_l[x][rows_owned])
        aid = D["A_ids"][e].reshape(tdof, tdof)            # [col i][row j]
        sel = np.broadcast_to(rows_owned[None, :], (tdof, tdof))
        np.add.at(Atl, aid[sel], Ke_l[x][sel])
    A = D["mass"] + D["gf"] * Atl
    I = D["I"].astype(np.int64)
    row_of = np.repeat(np.arange(D["NV"]), np.diff(I))
    ok_A = np.array_equal(A[owned_v[row_of]], D["A_neohookean"][owned_v[row_of]])
    ok_r = np.array_equal(rtl[owned_v], D["r2_neohookean"][owned_v])
    cover = torch.from_numpy(owned_v.astype(np.int64))
    dist.all_reduce(cover)
    out[rank] = (ok_A, ok_r, bool((cover.numpy() == 1).all()), len(R["send"]), len(R["recv"]))
    dist.destroy_process_group()


@pytest.mark.parametrize("name", ["cube_p2_n3_patch", "cube_p2_n2"])
def test_two_rank_partition_reproduces_reference(name):
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), name, out), nprocs=world, join=True)
    for r in range(world):
        ok_A, ok_r, cover, ns, nr = out[r]
        assert ok_A and ok_r and cover, (r, out[r])
    assert out[0][4] == 1 and out[1][3] == 1       # rank 0 receives ghosts from rank 1, rank 1 sends to rank 0


def test_ranges_and_ownership():
    from mimi_b200 import meshgen
    M = meshgen.patch_connectivity(2, (3, 3, 8))
    rg = P.element_ranges(M["NE"], 4)
    assert rg[0] == 0 and rg[-1] == M["NE"] and (np.diff(rg) > 0).all()
    owned = np.zeros(M["Nn"], int)
    for r in range(4):
        R = P.rank_problem(M["v_dofs"], 3, M["NV"], rg, r)
        owned += R["owned"]
        # every ghost has a larger id than every own element and ghosts are ascending
        assert (np.diff(R["local_elems"]) > 0).all()
        for peer, idx in R["send"].items():
            assert peer < r
    assert (owned == 1).all()


@pytest.mark.parametrize("p,spans,world", [(2, (3, 2, 8), 2), (2, (2, 3, 12), 4), (1, (2, 2, 6), 3), (3, (2, 2, 6), 2),
                                           # uneven slabs (strong scaling of a 100-layer block over 8 ranks: 13,13,13,13,12,12,12,12)
                                           (2, (2, 2, 11), 4), (2, (2, 3, 7), 3), (1, (3, 2, 13), 8)])
def test_slab_problem_equals_generic(p, spans, world):
    from mimi_b200 import meshgen
    M = meshgen.patch_connectivity(p, spans)
    rg = P.slab_bounds(spans[2], world) * (spans[0] * spans[1])       # whole layers per rank
    if spans[2] % world == 0:
        assert np.array_equal(rg, P.element_ranges(M["NE"], world))
    for r in range(world):
        G = P.rank_problem(M["v_dofs"], 3, M["NV"], rg, r)
        S = P.slab_rank_problem(p, spans, world, r)
        assert S["n_compute"] == G["n_compute"] and np.array_equal(S["local_elems"], G["local_elems"])
        assert np.array_equal(S["v_dofs"], G["v_dofs"]) and np.array_equal(S["owned"], G["owned"])
        assert set(S["send"]) == set(G["send"]) and set(S["recv"]) == set(G["recv"])
        for k in S["send"]:
            assert np.array_equal(S["send"][k], G["send"][k])
        for k in S["recv"]:
            assert np.array_equal(S["recv"][k], G["recv"][k])
        assert np.array_equal(S["x_ref"], M["x_ref"]) and np.array_equal(S["el_ijk"], M["el_ijk"][G["local_elems"]])


@pytest.mark.parametrize("p,spans,M,world", [(2, (2, 2, 4), 2, 4), (3, (2, 1, 6), 2, 4), (2, (3, 2, 2), 4, 2), (2, (2, 2, 2), 4, 4),
                                             (1, (2, 3, 4), 2, 8), (2, (2, 2, 4), 1, 2)])
def test_beam_rank_problem_equals_the_generic_planner(p, spans, M, world):
    """Config 4's analytic planner (glued multi-patch beam, no global connectivity) against rank_problem on the full
    connectivity: same own range, ghosts, owned nodes, send / receive lists, local connectivity."""
    from mimi_b200 import partition as P
    v_dofs, NV = P.beam_connectivity(p, spans, M)
    ranges = P.element_ranges(len(v_dofs), world)
    for rank in range(world):
        A = P.rank_problem(v_dofs, 3, NV, ranges, rank)
        B = P.beam_rank_problem(p, spans, M, world, rank)
        assert (A["e0"], A["e1"], A["n_compute"]) == (B["e0"], B["e1"], B["n_compute"])
        assert np.array_equal(A["local_elems"], B["local_elems"]) and np.array_equal(A["v_dofs"], B["v_dofs"])
        touched = A["owner"] >= 0
        assert np.array_equal(A["owned"][touched], B["owned"][touched]) and B["NV"] == NV
        assert sorted(A["send"]) == sorted(B["send"]) and sorted(A["recv"]) == sorted(B["recv"])
        for k in A["send"]:
            assert np.array_equal(A["send"][k], B["send"][k])
        for k in A["recv"]:
            assert np.array_equal(A["recv"][k], B["recv"][k])
        segs = B["patch_segments"]
        assert sum(n for _, n, _ in segs) == B["n_compute"] and segs[0][0] == 0


@pytest.mark.parametrize("name,world", [("cube_p2_n3_patch", 2), ("cube_p2_n3_patch", 3), ("cube_p2_n2", 8), ("beam2d_p3_n2", 2),
                                        ("beam4_p3", 4), ("beam3_p2", 5)])
def test_library_planner_equals_the_python_oracle(name, world):
    """mb2_plan_partition (C++, in libmimi_b200.so; pure host code, so it runs here without a GPU) against
    mimi_b200/partition.py::rank_problem on the reference's own connectivities, multi-patch ones included."""
    from mimi_b200 import capi, partition as P
    from tests.cases import load_golden
    D = load_golden(name)
    ranges = P.element_ranges(D["NE"], world)
    for rank in range(world):
        A = P.rank_problem(D["v_dofs"], D["dim"], D["NV"], ranges, rank)
        B = capi.Plan(D["v_dofs"], D["dim"], D["NV"], world, rank)
        assert (B.e0, B.n_compute) == (A["e0"], A["n_compute"])
        assert np.array_equal(B.local_elems, A["local_elems"]) and np.array_equal(B.owned.astype(bool), A["owned"])
        assert sorted(B.send) == sorted(A["send"]) and sorted(B.recv) == sorted(A["recv"])
        for k in A["send"]:
            assert np.array_equal(B.send[k], A["send"][k])
        for k in A["recv"]:
            assert np.array_equal(B.recv[k], A["recv"][k])
        B.close()


def test_planner_flags_partitions_whose_interface_elements_come_late():
    """mb2_plan_size(plan, 5): z-slabs of one patch send their first p layers (early: the NCCL exchange may start behind the
    pack); a beam whose patches are split over ranks in x sends every i < p column of a patch, all through the element
    range (late: mb2_plan_apply sets "mg_exchange_late")."""
    from mimi_b200 import capi, meshgen
    M = meshgen.patch_connectivity(2, (2, 2, 24))       # 6 layers per rank, the first 2 of them go to the rank below
    assert not any(capi.Plan(M["v_dofs"], 3, M["NV"], 4, r).late_exchange for r in range(4))
    v_dofs, NV = P.beam_connectivity(2, (3, 2, 2), 4)
    assert all(capi.Plan(v_dofs, 3, NV, 4, r).late_exchange for r in range(4))
