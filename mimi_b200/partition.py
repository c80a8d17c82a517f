"""Element partition of one assembly across the GPUs of a box.

Rules (they keep the multi-GPU result bit-identical to the reference's single-thread assembly, which adds
element contributions in ascending element id, src/mimi/integrators/nonlinear_solid.cpp:126-143):

* rank g computes the contiguous element range [E_g, E_{g+1}) of the global ascending numbering;
* a node (= dim CSR rows and dim residual entries) is owned by the rank that computes its LOWEST element;
* elements of higher ranks that touch a node owned by rank g are g's ghost elements: their raw element
  matrices / residuals (not partial sums) are sent to g and appended after g's own elements, so g's ordered
  gather still sums in ascending global element id. Ranges are contiguous and ascending, hence all ghosts of
  g have larger ids than its own elements and the ghosts of one source rank form one contiguous block.
"""
from __future__ import annotations

import numpy as np


def element_ranges(NE: int, world: int) -> np.ndarray:
    """E_0..E_world: balanced contiguous ranges."""
    return np.array([(NE * r) // world for r in range(world + 1)], np.int64)


def node_first_element(v_dofs: np.ndarray, dim: int, Nn: int) -> np.ndarray:
    nd = v_dofs.shape[1] // dim
    nodes = v_dofs[:, :nd] // dim
    first = np.full(Nn, np.iinfo(np.int64).max, np.int64)
    e = np.repeat(np.arange(v_dofs.shape[0], dtype=np.int64), nd)
    np.minimum.at(first, nodes.ravel(), e)
    return first


def rank_problem(v_dofs: np.ndarray, dim: int, NV: int, ranges: np.ndarray, rank: int) -> dict:
    """Everything rank `rank` needs: local connectivity (own + ghost elements, global node ids), owned-node
    mask, and per-peer send / receive element lists (local indices, ascending)."""
    Nn = NV // dim
    nd = v_dofs.shape[1] // dim
    world = len(ranges) - 1
    first = node_first_element(v_dofs, dim, Nn)
    touched = first < np.iinfo(np.int64).max
    owner = np.full(Nn, -1, np.int64)
    owner[touched] = np.searchsorted(ranges, first[touched], side="right") - 1
    e0, e1 = int(ranges[rank]), int(ranges[rank + 1])
    nodes = v_dofs[:, :nd] // dim
    el_owner_min = owner[nodes].min(axis=1)          # lowest owner among an element's nodes
    el_touches_mine = (owner[nodes] == rank).any(axis=1)
    # ghosts: elements of higher ranks touching one of my nodes
    ghost = np.nonzero(el_touches_mine & (np.arange(len(v_dofs)) >= e1))[0]
    ghost_src = np.searchsorted(ranges, ghost, side="right") - 1
    recv = {int(s): np.nonzero(ghost_src == s)[0] + (e1 - e0) for s in np.unique(ghost_src)}  # local indices
    # sends: my elements touching a node owned by a lower rank r  (one list per r)
    send = {}
    own_nodes_owner = owner[nodes[e0:e1]]
    for r in range(rank):
        idx = np.nonzero((own_nodes_owner == r).any(axis=1))[0]
        if len(idx):
            send[r] = idx.astype(np.int64)
    local = np.concatenate([np.arange(e0, e1), ghost])
    return dict(rank=rank, world=world, e0=e0, e1=e1, n_compute=e1 - e0, local_elems=local,
                v_dofs=np.ascontiguousarray(v_dofs[local]), owned=(owner == rank), owner=owner,
                send=send, recv=recv, el_owner_min=el_owner_min)


def slab_rank_problem(p: int, spans, world: int, rank: int, lengths=None) -> dict:
    """rank_problem for a single-patch box cut into z-slabs of whole element layers, built WITHOUT the global
    connectivity (the 8-GPU weak-scaling mesh has 8e6 elements). Same rules as rank_problem; equality with it is
    checked in tests/test_partition.py. Returns the rank_problem dict plus the mesh pieces bench.py needs."""
    from . import meshgen
    n0, n1, n2 = (int(s) for s in spans)
    assert n2 % world == 0 and n2 // world >= p, "z spans must split into slabs of at least p layers"
    lz = n2 // world
    k0, k1 = rank * lz, (rank + 1) * lz
    kg = min(k1 + p, n2)                                   # ghost layers [k1, kg)
    kvs = [meshgen.open_uniform_knots(p, n) for n in spans]
    ncp = [len(kv) - p - 1 for kv in kvs]
    g = meshgen.patch_node_grid(ncp)
    ii, jj = np.meshgrid(np.arange(n0), np.arange(n1), indexing="ij")
    layer_ij = np.stack([ii.T.ravel(), jj.T.ravel()], 1)   # i fastest within a layer
    ks = np.arange(k0, kg)
    el_ijk = np.concatenate([np.column_stack([layer_ij, np.full(len(layer_ij), k)]) for k in ks]).astype(np.int32)
    nd = (p + 1) ** 3
    nodes = np.empty((len(el_ijk), nd), np.int32)
    a = 0
    for c in range(p + 1):
        for b in range(p + 1):
            for a_ in range(p + 1):
                nodes[:, a] = g[el_ijk[:, 0] + a_, el_ijk[:, 1] + b, el_ijk[:, 2] + c]
                a += 1
    v_dofs = np.concatenate([3 * nodes + c for c in range(3)], 1).astype(np.int32)
    Nn = g.size
    # owner of a node = rank of the layer of its lowest element = max(0, k_node - p)
    knode = np.arange(ncp[2])
    owner_k = np.maximum(knode - p, 0) // lz
    owned = np.zeros(Nn, bool)
    owned[g[:, :, owner_k == rank].ravel()] = True
    n_layer = n0 * n1
    n_compute = lz * n_layer
    send = {rank - 1: np.arange(p * n_layer, dtype=np.int64)} if rank > 0 else {}
    recv = {rank + 1: np.arange(n_compute, n_compute + (kg - k1) * n_layer, dtype=np.int64)} if kg > k1 else {}
    lengths = lengths or (1.0, 1.0, 1.0)
    grev = [meshgen.greville(p, kv) * L for kv, L in zip(kvs, lengths)]
    x = np.empty((Nn, 3))
    x[g.ravel(), 0] = np.repeat(grev[0], ncp[1] * ncp[2])
    x[g.ravel(), 1] = np.tile(np.repeat(grev[1], ncp[2]), ncp[0])
    x[g.ravel(), 2] = np.tile(grev[2], ncp[0] * ncp[1])
    e0 = k0 * n_layer
    return dict(rank=rank, world=world, e0=e0, e1=e0 + n_compute, n_compute=n_compute,
                local_elems=np.arange(e0, e0 + len(el_ijk)), v_dofs=v_dofs, owned=owned, send=send, recv=recv,
                el_ijk=el_ijk, knots=kvs, ncp=np.array(ncp, np.int32), x_ref=x.ravel(), NV=3 * Nn, Nn=Nn, nd=nd, p=p,
                NE=len(el_ijk), NE_global=n_layer * n2)
