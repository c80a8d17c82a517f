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
