"""Writers for single-patch NURBS meshes in MFEM's "MFEM NURBS mesh v1.0" text format.

The reference loads its geometry with ``mfem::Mesh(fname, 1, 1)`` (src/mimi/py/py_solid.cpp:76) and
then calls ``DegreeElevate`` / ``UniformRefinement`` (py_solid.cpp:155,177). These writers produce
the inputs for that path (a degree-1 box) and, for span counts that are not a power of two, the
``patches`` variant of the format (third_party/mfem/mesh/nurbs.cpp:2037-2044) holding an open
uniform knot vector of any degree with control points at the Greville abscissae (affine geometry).
"""
from __future__ import annotations

import numpy as np

_HEX_TOPOLOGY = """dimension
3

elements
1
1 5 0 1 2 3 4 5 6 7

boundary
6
1 3 3 2 1 0
2 3 4 5 6 7
3 3 0 1 5 4
4 3 1 2 6 5
5 3 2 3 7 6
6 3 3 0 4 7

edges
12
0 0 1
0 3 2
0 4 5
0 7 6
1 0 3
1 1 2
1 4 7
1 5 6
2 0 4
2 1 5
2 2 6
2 3 7

vertices
8
"""

_QUAD_TOPOLOGY = """dimension
2

elements
1
1 3 0 1 2 3

boundary
4
1 1 0 1
2 1 2 3
3 1 3 0
4 1 1 2

edges
4
0 0 1
0 3 2
1 0 3
1 1 2

vertices
4
"""


def write_box_p1(path: str, lengths=(1.0, 1.0, 1.0)) -> None:
    """Degree-1, one-element box (2D or 3D) with vertex coordinates in MFEM's vertex order."""
    dim = len(lengths)
    if dim == 3:
        lx, ly, lz = lengths
        verts = [(0, 0, 0), (lx, 0, 0), (lx, ly, 0), (0, ly, 0), (0, 0, lz), (lx, 0, lz), (lx, ly, lz), (0, ly, lz)]
        topo = _HEX_TOPOLOGY
    elif dim == 2:
        lx, ly = lengths
        verts = [(0, 0), (lx, 0), (lx, ly), (0, ly)]
        topo = _QUAD_TOPOLOGY
    else:
        raise ValueError("dim must be 2 or 3")
    with open(path, "w") as f:
        f.write("MFEM NURBS mesh v1.0\n\n" + topo + "\nknotvectors\n%d\n" % dim)
        for _ in range(dim):
            f.write("1 2 0 0 1 1\n")
        f.write("\nweights\n" + "1\n" * len(verts))
        f.write("\nFiniteElementSpace\nFiniteElementCollection: NURBS1\nVDim: %d\nOrdering: 1\n\n" % dim)
        for v in verts:
            f.write(" ".join(repr(float(c)) for c in v) + "\n")


def open_uniform_knots(p: int, n: int) -> np.ndarray:
    """Open uniform knot vector of degree p with n spans on [0, 1] (n + 2p + 1 knots)."""
    inner = np.arange(0, n + 1, dtype=np.float64) / n
    return np.concatenate([np.zeros(p), inner, np.ones(p)])


def greville(p: int, knots: np.ndarray) -> np.ndarray:
    ncp = len(knots) - p - 1
    return np.array([knots[i + 1:i + p + 1].sum() / p for i in range(ncp)])


def write_box_patch(path: str, p: int, spans, lengths=None) -> None:
    """Single-patch box of degree p with spans[d] knot spans per direction, 'patches' format."""
    dim = len(spans)
    lengths = lengths or (1.0,) * dim
    topo = _HEX_TOPOLOGY if dim == 3 else _QUAD_TOPOLOGY
    kvs = [open_uniform_knots(p, n) for n in spans]
    grev = [greville(p, kv) * L for kv, L in zip(kvs, lengths)]
    with open(path, "w") as f:
        f.write("MFEM NURBS mesh v1.0\n\n" + topo + "\npatches\n\nknotvectors\n%d\n" % dim)
        for kv in kvs:
            f.write("%d %d " % (p, len(kv) - p - 1) + " ".join(repr(float(k)) for k in kv) + "\n")
        f.write("\ndimension\n%d\n\ncontrolpoints_cartesian\n" % dim)
        if dim == 3:
            for z in grev[2]:
                for y in grev[1]:
                    for x in grev[0]:
                        f.write("%r %r %r 1\n" % (float(x), float(y), float(z)))
        else:
            for y in grev[1]:
                for x in grev[0]:
                    f.write("%r %r 1\n" % (float(x), float(y)))
