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


def slab_bounds(n_layers: int, world: int) -> np.ndarray:
    """Balanced cut of n_layers element layers into `world` slabs: the first n_layers % world ranks get one layer more."""
    base, rem = divmod(int(n_layers), int(world))
    return np.concatenate([[0], np.cumsum([base + (1 if r < rem else 0) for r in range(world)])]).astype(np.int64)


def slab_rank_problem(p: int, spans, world: int, rank: int, lengths=None) -> dict:
    """rank_problem for a single-patch box cut into z-slabs of whole element layers, built WITHOUT the global
    connectivity (the 8-GPU weak-scaling mesh has 8e6 elements). Same rules as rank_problem; equality with it is
    checked in tests/test_partition.py. Returns the rank_problem dict plus the mesh pieces bench.py needs."""
    from . import meshgen
    n0, n1, n2 = (int(s) for s in spans)
    kb = slab_bounds(n2, world)                            # layers [kb[r], kb[r+1]) belong to rank r (balanced, may be uneven)
    assert min(np.diff(kb)) >= p, "z spans must split into slabs of at least p layers"
    k0, k1 = int(kb[rank]), int(kb[rank + 1])
    lz = k1 - k0
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
    owner_k = np.searchsorted(kb[1:], np.maximum(knode - p, 0), side="right")
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
                NE=len(el_ijk), NE_global=n_layer * n2, node_grid_top=g[:, :, -1].copy())


# ---------------------------------------------------------------------------------------------------
# BASELINE config 4: multi-patch beam, M patches glued along x, element-partitioned across the ranks.
# Numbering used by the throughput bench (NOT MFEM's multi-patch numbering, which the parity fixtures carry):
# nodes lexicographic over the glued control grid, ix fastest; elements patch by patch, i fastest, then j, then k.
# ---------------------------------------------------------------------------------------------------
def _beam_dims(p, spans, M):
    nx, ny, nz = (int(s) for s in spans)
    cx = nx + p
    return nx, ny, nz, cx, M * (cx - 1) + 1, ny + p, nz + p


def beam_connectivity(p: int, spans, n_patches: int):
    """Global connectivity of the glued beam (small sizes: tests): v_dofs [NE, 3 (p+1)^3], NV."""
    M = int(n_patches)
    nx, ny, nz, cx, NX, NY, NZ = _beam_dims(p, spans, M)
    s, k, j, i = np.meshgrid(np.arange(M), np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    return _beam_vdofs(p, spans, M, s.ravel(), i.ravel(), j.ravel(), k.ravel()), 3 * NX * NY * NZ


def _beam_vdofs(p, spans, M, s, i, j, k):
    nx, ny, nz, cx, NX, NY, NZ = _beam_dims(p, spans, M)
    nd = (p + 1) ** 3
    nodes = np.empty((len(s), nd), np.int64)
    a = 0
    for c in range(p + 1):
        for b in range(p + 1):
            for a_ in range(p + 1):
                nodes[:, a] = (s * (cx - 1) + i + a_) + NX * ((j + b) + NY * (k + c))
                a += 1
    return np.concatenate([3 * nodes + c for c in range(3)], 1).astype(np.int32)


def beam_rank_problem(p: int, spans, n_patches: int, world: int, rank: int, lengths=(1.0, 1.0, 1.0)) -> dict:
    """rank_problem for the glued beam WITHOUT the global connectivity (config 4 has 4e6 p = 3 elements). world >= M: every
    patch is cut into world / M z-slabs of whole element layers; world < M: every rank takes M / world whole patches.
    Same rules as rank_problem (node owner = rank of its lowest element; raw element matrices of higher ranks that touch
    an owned node arrive as ghosts); equality with rank_problem on beam_connectivity is checked in tests/test_partition.py."""
    from . import meshgen
    M = int(n_patches)
    nx, ny, nz, cx, NX, NY, NZ = _beam_dims(p, spans, M)
    if world >= M:
        assert world % M == 0 and nz % (world // M) == 0 and nz // (world // M) >= p, "slabs of >= p whole layers per rank"
        R, lz = world // M, nz // (world // M)
        rank_of = lambda s_, k_: s_ * R + k_ // lz                      # noqa: E731
    else:
        assert M % world == 0
        G = M // world
        rank_of = lambda s_, k_: s_ // G + 0 * k_                       # noqa: E731
    npe, nlay = nx * ny * nz, nx * ny
    ks = np.arange(nz)
    own_blocks, ghost_parts, send_parts = [], [], {}
    lay = np.arange(nlay)
    lay_i0 = np.arange(ny) * nx                                          # elements with i == 0 inside a layer
    for s_ in range(M):
        for k_ in range(nz):
            re = int(rank_of(s_, k_))
            kk = np.arange(max(0, k_ - p), k_ + 1)
            set_s = set(int(x) for x in rank_of(s_, kk))
            set_sm = set(int(x) for x in rank_of(s_ - 1, kk)) if s_ > 0 else set()
            base = s_ * npe + k_ * nlay
            if re == rank:
                own_blocks.append(base + lay)
                for r2 in sorted((set_s | set_sm) - {rank}):
                    assert r2 < rank
                    send_parts.setdefault(r2, []).append(base + (lay if r2 in set_s else lay_i0))
            elif re > rank:
                if rank in set_s:
                    ghost_parts.append((re, base + lay))
                elif rank in set_sm:
                    ghost_parts.append((re, base + lay_i0))
    own = np.concatenate(own_blocks)
    e0, n_compute = int(own[0]), len(own)
    assert np.array_equal(own, np.arange(e0, e0 + n_compute))
    ghosts = np.concatenate([g for _, g in ghost_parts]) if ghost_parts else np.zeros(0, np.int64)
    gsrc = np.concatenate([np.full(len(g), r2) for r2, g in ghost_parts]) if ghost_parts else np.zeros(0, np.int64)
    order = np.argsort(ghosts, kind="stable")
    ghosts, gsrc = ghosts[order], gsrc[order]
    local = np.concatenate([own, ghosts])
    recv = {int(r2): np.nonzero(gsrc == r2)[0] + n_compute for r2 in np.unique(gsrc)}
    send = {int(r2): np.sort(np.concatenate(v)) - e0 for r2, v in send_parts.items()}
    s_ = local // npe
    rem = local - s_ * npe
    k_, rem = rem // nlay, rem % nlay
    j_, i_ = rem // nx, rem % nx
    v_dofs = _beam_vdofs(p, spans, M, s_, i_, j_, k_)
    # node owners
    ix, iz = np.arange(NX), np.arange(NZ)
    smin = np.where(ix == 0, 0, np.minimum((ix - 1) // (cx - 1), M - 1))
    kmin = np.maximum(iz - p, 0)
    owner_xz = rank_of(smin[:, None], kmin[None, :])                     # [NX, NZ]
    owned = np.zeros((NZ, NY, NX), bool)
    owned[:] = (owner_xz.T == rank)[:, None, :]
    kvs = [meshgen.open_uniform_knots(p, n) for n in (nx, ny, nz)]
    grev = [meshgen.greville(p, kv) * L for kv, L in zip(kvs, lengths)]
    sx = np.minimum(ix // (cx - 1), M - 1)
    X = sx * lengths[0] + grev[0][ix - sx * (cx - 1)]
    x = np.empty((NZ, NY, NX, 3))
    x[..., 0] = X[None, None, :]
    x[..., 1] = grev[1][None, :, None]
    x[..., 2] = grev[2][:, None, None]
    # own elements patch by patch (mb2_generate_patch_tables is called once per segment)
    segs = []
    so = s_[:n_compute]
    for q in np.unique(so):
        idx = np.nonzero(so == q)[0]
        segs.append((int(idx[0]), int(len(idx)), int(q)))
    el_ijk = np.stack([i_, j_, k_], 1).astype(np.int32)
    Nn = NX * NY * NZ
    return dict(rank=rank, world=world, e0=e0, e1=e0 + n_compute, n_compute=n_compute, local_elems=local, v_dofs=v_dofs,
                owned=owned.ravel(), send=send, recv=recv, el_ijk=el_ijk, knots=kvs,
                ncp=np.array([cx, NY, NZ], np.int32), x_ref=x.reshape(-1), NV=3 * Nn, Nn=Nn, nd=(p + 1) ** 3, p=p,
                NE=len(local), NE_global=M * npe, patch_segments=segs)
