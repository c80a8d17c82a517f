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


def write_cell_p1(path: str, verts, weights=None) -> None:
    """Degree-1, one-element NURBS mesh with the given vertices (4 in 2D / 8 in 3D, MFEM's vertex order): a general
    (non-affine) quadrilateral / hexahedron, e.g. the reference's tests/data/square-nurbs.mesh = [(0,0),(2,0),(1,1),(-1,2)];
    optional NURBS weights per vertex make the patch rational."""
    verts = [tuple(float(c) for c in v) for v in verts]
    dim = len(verts[0])
    if (dim, len(verts)) not in ((2, 4), (3, 8)):
        raise ValueError("4 vertices in 2D, 8 in 3D")
    topo = _QUAD_TOPOLOGY if dim == 2 else _HEX_TOPOLOGY
    w = [1.0] * len(verts) if weights is None else [float(x) for x in weights]
    with open(path, "w") as f:
        f.write("MFEM NURBS mesh v1.0\n\n" + topo + "\nknotvectors\n%d\n" % dim)
        for _ in range(dim):
            f.write("1 2 0 0 1 1\n")
        f.write("\nweights\n" + "".join(repr(x) + "\n" for x in w))
        f.write("\nFiniteElementSpace\nFiniteElementCollection: NURBS1\nVDim: %d\nOrdering: 1\n\n" % dim)
        for v in verts:
            f.write(" ".join(repr(c) for c in v) + "\n")


SQUARE_NURBS_VERTS = [(0.0, 0.0), (2.0, 0.0), (1.0, 1.0), (-1.0, 2.0)]   # the reference's tests/data/square-nurbs.mesh


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


def write_patch(path: str, p: int, knots, cps, weights) -> None:
    """Single NURBS patch of degree p in MFEM's 'patches' format: knots[d] per direction, control points
    cps[k][j][i][dim] (3D) / cps[j][i][dim] (2D) with direction 0 fastest, weights of the same shape (Cartesian
    coordinates, not pre-multiplied)."""
    cps = np.asarray(cps, np.float64)
    w = np.asarray(weights, np.float64)
    dim = cps.shape[-1]
    topo = _HEX_TOPOLOGY if dim == 3 else _QUAD_TOPOLOGY
    with open(path, "w") as f:
        f.write("MFEM NURBS mesh v1.0\n\n" + topo + "\npatches\n\nknotvectors\n%d\n" % dim)
        for kv in knots:
            f.write("%d %d " % (p, len(kv) - p - 1) + " ".join(repr(float(k)) for k in kv) + "\n")
        f.write("\ndimension\n%d\n\ncontrolpoints_cartesian\n" % dim)
        for x, ww in zip(cps.reshape(-1, dim), w.reshape(-1)):
            f.write(" ".join(repr(float(c)) for c in x) + " %r\n" % float(ww))


def write_beam_patches(path: str, p: int, spans, n_patches: int, lengths=(1.0, 1.0, 1.0)) -> None:
    """Multi-patch beam (BASELINE config 4): n_patches hexahedral patches of degree p glued along x, each with spans =
    (nx, ny, nz) knot spans and size lengths (per patch), open uniform knots, control points at the Greville abscissae.
    Vertices: cross-section s holds 4 s + {0: (y0, z0), 1: (y1, z0), 2: (y1, z1), 3: (y0, z1)}; knot vector s belongs to
    the x edges of patch s, n_patches / n_patches + 1 to all y / z edges."""
    M = int(n_patches)
    kvs = [open_uniform_knots(p, n) for n in spans]
    grev = [greville(p, kv) * L for kv, L in zip(kvs, lengths)]
    with open(path, "w") as f:
        f.write("MFEM NURBS mesh v1.0\n\ndimension\n3\n\nelements\n%d\n" % M)
        for s_ in range(M):
            a, b = 4 * s_, 4 * (s_ + 1)
            f.write("1 5 %d %d %d %d %d %d %d %d\n" % (a, b, b + 1, a + 1, a + 3, b + 3, b + 2, a + 2))
        f.write("\nboundary\n%d\n" % (4 * M + 2))
        for s_ in range(M):
            a, b = 4 * s_, 4 * (s_ + 1)
            f.write("1 3 %d %d %d %d\n" % (a + 1, b + 1, b, a))          # z = 0
            f.write("2 3 %d %d %d %d\n" % (a + 3, b + 3, b + 2, a + 2))  # z = 1
            f.write("3 3 %d %d %d %d\n" % (a, b, b + 3, a + 3))          # y = 0
            f.write("5 3 %d %d %d %d\n" % (b + 1, a + 1, a + 2, b + 2))  # y = 1
        f.write("6 3 %d %d %d %d\n" % (1, 0, 3, 2))                       # x = 0
        e = 4 * M
        f.write("4 3 %d %d %d %d\n" % (e, e + 1, e + 2, e + 3))           # x = L
        f.write("\nedges\n%d\n" % (4 * M + 4 * (M + 1)))
        for s_ in range(M):
            for i in range(4):
                f.write("%d %d %d\n" % (s_, 4 * s_ + i, 4 * (s_ + 1) + i))
        for s_ in range(M + 1):
            a = 4 * s_
            f.write("%d %d %d\n%d %d %d\n" % (M, a, a + 1, M, a + 3, a + 2))
            f.write("%d %d %d\n%d %d %d\n" % (M + 1, a, a + 3, M + 1, a + 1, a + 2))
        f.write("\nvertices\n%d\n\npatches\n\n" % (4 * (M + 1)))
        for s_ in range(M):
            f.write("knotvectors\n3\n")
            for kv in kvs:
                f.write("%d %d " % (p, len(kv) - p - 1) + " ".join(repr(float(k)) for k in kv) + "\n")
            f.write("\ndimension\n3\n\ncontrolpoints_cartesian\n")
            for z in grev[2]:
                for y in grev[1]:
                    for x in grev[0]:
                        f.write("%r %r %r 1\n" % (float(x + s_ * lengths[0]), float(y), float(z)))
            f.write("\n")


def write_quarter_annulus(path: str, r0=1.0, r1=2.0, height=None) -> None:
    """Rational, curved geometry: the exact quarter annulus r0 <= r <= r1 (2D), or the thick quarter cylinder of the given
    height (3D), p = 2 in every direction, one knot span each (refine with Mesh::UniformRefinement). Direction 0 = radial,
    direction 1 = angular (rational quadratic arc, weights 1, sqrt(1/2), 1), direction 2 = axial."""
    s = np.sqrt(0.5)
    kv = [0.0, 0.0, 0.0, 1.0, 1.0, 1.0]
    rs = [r0, 0.5 * (r0 + r1), r1]
    ang = [((1.0, 0.0), 1.0), ((1.0, 1.0), s), ((0.0, 1.0), 1.0)]
    if height is None:
        cps = [[[r * a[0], r * a[1]] for r in rs] for a, _ in ang]
        w = [[ww for _ in rs] for _, ww in ang]
        write_patch(path, 2, [kv, kv], cps, w)
    else:
        zs = [0.0, 0.5 * height, height]
        cps = [[[[r * a[0], r * a[1], z] for r in rs] for a, _ in ang] for z in zs]
        w = [[[ww for _ in rs] for _, ww in ang] for _ in zs]
        write_patch(path, 2, [kv, kv, kv], cps, w)


# ---------------------------------------------------------------------------------------------------
# Synthetic connectivity of a single 3D patch in MFEM's own numbering, so that large meshes (which the
# reference cannot precompute on a host) get the same DOF indexing the reference would produce.
#
# MFEM numbers the control points of a patch as: 8 patch vertices, then the interiors of the 12 patch
# edges, then of the 6 patch faces, then the patch interior (NURBSExtension::GenerateOffsets,
# third_party/mfem/mesh/nurbs.cpp:3255-3330), each entity traversed with the orientation the patch
# topology gives it (NURBSPatchMap::operator()(i,j,k), third_party/mfem/mesh/nurbs.hpp:1201-1240). The
# entity order/orientation below is the one of the hexahedron topology written by write_box_p1 /
# write_box_patch; tests/test_meshgen.py checks it against dumps of the compiled reference.
# Elements run i fastest, then j, then k, and list their (p+1)^3 nodes ii fastest
# (Generate3DElementDofTable, nurbs.cpp:3759-3823).
# ---------------------------------------------------------------------------------------------------
def patch_node_grid(ncp) -> np.ndarray:
    """Node id of control point (i, j, k) for a single hexahedral patch: int32 array [ncp0, ncp1, ncp2]."""
    n0, n1, n2 = (int(x) for x in ncp)
    I, J, K = n0 - 2, n1 - 2, n2 - 2
    g = np.full((n0, n1, n2), -1, np.int64)
    # vertices
    for v, (i, j, k) in enumerate([(0, 0, 0), (-1, 0, 0), (-1, -1, 0), (0, -1, 0),
                                   (0, 0, -1), (-1, 0, -1), (-1, -1, -1), (0, -1, -1)]):
        g[i, j, k] = v
    c = 8
    ax, ay, az = np.arange(I), np.arange(J), np.arange(K)
    # edges: 4 along x (the 2nd and 4th reversed), 4 along y, 4 along z
    for (j, k, rev) in [(0, 0, 0), (-1, 0, 1), (0, -1, 0), (-1, -1, 1)]:
        g[1:-1, j, k] = c + (ax[::-1] if rev else ax)
        c += I
    for (i, k) in [(0, 0), (-1, 0), (0, -1), (-1, -1)]:
        g[i, 1:-1, k] = c + ay
        c += J
    for (i, j) in [(0, 0), (-1, 0), (-1, -1), (0, -1)]:
        g[i, j, 1:-1] = c + az
        c += K
    # faces: z=0, y=0, x=1, y=1, x=0, z=1
    g[1:-1, 1:-1, 0] = c + ax[:, None] + I * (J - 1 - ay)[None, :]
    c += I * J
    g[1:-1, 0, 1:-1] = c + ax[:, None] + I * az[None, :]
    c += I * K
    g[-1, 1:-1, 1:-1] = c + ay[:, None] + J * az[None, :]
    c += J * K
    g[1:-1, -1, 1:-1] = c + (I - 1 - ax)[:, None] + I * az[None, :]
    c += I * K
    g[0, 1:-1, 1:-1] = c + (J - 1 - ay)[:, None] + J * az[None, :]
    c += J * K
    g[1:-1, 1:-1, -1] = c + ax[:, None] + I * ay[None, :]
    c += I * J
    # interior
    g[1:-1, 1:-1, 1:-1] = c + ax[:, None, None] + I * (ay[None, :, None] + J * az[None, None, :])
    assert (g >= 0).all() and c + I * J * K == n0 * n1 * n2
    return g.astype(np.int32)


def patch_connectivity(p: int, spans, lengths=None):
    """Synthetic single-patch box: returns dict(v_dofs [NE, 3*(p+1)^3] int32 component-major byVDIM,
    el_ijk [NE,3] int32 knot-span indices, knots (3 arrays), ncp, x_ref [NV] byVDIM (Greville control
    points = affine geometry), weights [Nn] (all 1), NV)."""
    lengths = lengths or (1.0, 1.0, 1.0)
    kvs = [open_uniform_knots(p, n) for n in spans]
    ncp = [len(kv) - p - 1 for kv in kvs]
    g = patch_node_grid(ncp)
    n0, n1, n2 = spans
    ei, ej, ek = np.meshgrid(np.arange(n0), np.arange(n1), np.arange(n2), indexing="ij")
    # element order: i fastest
    order = np.argsort((ek * n1 + ej).ravel() * n0 + ei.ravel(), kind="stable")
    el_ijk = np.stack([ei.ravel()[order], ej.ravel()[order], ek.ravel()[order]], 1).astype(np.int32)
    nd = (p + 1) ** 3
    nodes = np.empty((len(el_ijk), nd), np.int32)
    a = 0
    for kk in range(p + 1):
        for jj in range(p + 1):
            for ii in range(p + 1):
                nodes[:, a] = g[el_ijk[:, 0] + ii, el_ijk[:, 1] + jj, el_ijk[:, 2] + kk]
                a += 1
    v_dofs = np.concatenate([3 * nodes + c for c in range(3)], 1).astype(np.int32)
    Nn = g.size
    x = np.empty((Nn, 3))
    grev = [greville(p, kv) * L for kv, L in zip(kvs, lengths)]
    x[g.ravel(), 0] = np.repeat(grev[0], ncp[1] * ncp[2])
    x[g.ravel(), 1] = np.tile(np.repeat(grev[1], ncp[2]), ncp[0])
    x[g.ravel(), 2] = np.tile(grev[2], ncp[0] * ncp[1])
    return dict(v_dofs=v_dofs, el_ijk=el_ijk, knots=kvs, ncp=np.array(ncp, np.int32), x_ref=x.ravel(),
                weights=np.ones(Nn), NV=3 * Nn, Nn=Nn, NE=len(el_ijk), nd=nd, p=p)


def synth_u(NV: int, seed: int = 12345) -> np.ndarray:
    """u_i = 1e-3 (U_i - 0.5), U_i = splitmix64(seed + i) >> 11 scaled to [0,1): SURVEY.md section 8(d);
    the same rule oracle/ref_harness.cpp applies inside the compiled reference."""
    x = (np.arange(NV, dtype=np.uint64) + np.uint64(seed)) + np.uint64(0x9E3779B97F4A7C15)
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    x = x ^ (x >> np.uint64(31))
    U = (x >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    return 1.0e-3 * (U - 0.5)


def bspline_basis_1d(p: int, knots: np.ndarray, span: int, xi: np.ndarray):
    """Values and reference-coordinate derivatives of the p+1 B-spline basis functions that are non-zero on knot span
    `span` (between knots[p+span] and knots[p+span+1]) at local coordinates xi in [0,1] (Cox-de Boor).
    Returns N [len(xi), p+1], dN/dxi [len(xi), p+1]."""
    i = p + span
    t0, t1 = knots[i], knots[i + 1]
    t = t0 + (t1 - t0) * xi

    def basis(deg):
        N = np.zeros((len(xi), deg + 1))
        N[:, 0] = 1.0
        left, right = np.zeros((deg + 1, len(xi))), np.zeros((deg + 1, len(xi)))
        for j in range(1, deg + 1):
            left[j] = t - knots[i + 1 - j]
            right[j] = knots[i + j] - t
            saved = np.zeros(len(xi))
            for r in range(j):
                tmp = N[:, r] / (right[r + 1] + left[j - r])
                N[:, r] = saved + right[r + 1] * tmp
                saved = left[j - r] * tmp
            N[:, j] = saved
        return N

    N = basis(p)
    Nm = basis(p - 1) if p > 0 else np.zeros((len(xi), 1))
    dN = np.zeros_like(N)
    for a in range(p + 1):          # function index i - p + a
        idx = i - p + a
        if a > 0:
            dN[:, a] += p * Nm[:, a - 1] / (knots[idx + p] - knots[idx])
        if a < p:
            dN[:, a] -= p * Nm[:, a] / (knots[idx + p + 1] - knots[idx + 1])
    return N, dN * (t1 - t0)


def top_face_tables(M, spans, quad_points=None):
    """Boundary-face data of the z = max face of a patch_connectivity() box, in the layout mb2_contact_set_faces takes:
    v_dofs [NF, 3*(p+1)^2] (component-major byVDIM; face nodes ii fastest, then jj), N [NF, nq, (p+1)^2],
    dN_dxi [NF, nq, 2, (p+1)^2], w [NF, nq]. Face parameters (xi0, xi1) = (x, y): the face normal x_xi0 x x_xi1 = +z
    points out of the body, as MFEM orients this boundary element; node order, values and derivatives agree with the
    reference's tables for this face to rounding (tests/test_meshgen.py). Synthetic workloads (bench.py); parity tests
    use the reference's own face tables."""
    p = M["p"]
    n0, n1, n2 = spans
    nq1 = quad_points or (p + 2)                      # order 2p+3 Gauss-Legendre
    gx, gw = np.polynomial.legendre.leggauss(nq1)
    gx, gw = 0.5 * (gx + 1.0), 0.5 * gw
    el = M["el_ijk"]
    top = np.nonzero(el[:, 2] == n2 - 1)[0]
    nd1 = p + 1
    vd = M["v_dofs"][top]
    nd = nd1 ** 3
    loc = np.array([ii + nd1 * (jj + nd1 * p) for jj in range(nd1) for ii in range(nd1)])
    nodes = vd[:, loc] // 3
    fv = np.concatenate([3 * nodes + c for c in range(3)], 1).astype(np.int32)
    NF, ndf, nq = len(top), nd1 * nd1, nq1 * nq1
    N = np.empty((NF, nq, ndf))
    dN = np.empty((NF, nq, 2, ndf))
    bx = [bspline_basis_1d(p, M["knots"][0], s, gx) for s in range(n0)]
    by = [bspline_basis_1d(p, M["knots"][1], s, gx) for s in range(n1)]
    for f, e in enumerate(top):
        Nx, dNx = bx[el[e, 0]]
        Ny, dNy = by[el[e, 1]]
        # q = qx + nq1*qy ; a = ii + nd1*jj
        N[f] = np.einsum("xi,yj->yxji", Nx, Ny).reshape(nq, ndf)
        dN[f, :, 0] = np.einsum("xi,yj->yxji", dNx, Ny).reshape(nq, ndf)
        dN[f, :, 1] = np.einsum("xi,yj->yxji", Nx, dNy).reshape(nq, ndf)
    w = np.tile(np.outer(gw, gw).reshape(1, nq), (NF, 1))
    return dict(v_dofs=fv, N=N, dN_dxi=dN, w=w, NF=NF, ndf=ndf, nqf=nq)


def top_face_tables_from_grid(p, spans_xy, knots, gtop, quad_points=None):
    """top_face_tables for a box whose connectivity is not held as a whole (partitioned runs: every rank evaluates all
    contact faces): gtop[i, j] = node id of control point (i, j) of the top control layer (patch_node_grid(ncp)[:, :, -1]).
    Same layout, face order (i fastest) and values as top_face_tables."""
    n0, n1 = spans_xy
    nq1 = quad_points or (p + 2)
    gx, gw = np.polynomial.legendre.leggauss(nq1)
    gx, gw = 0.5 * (gx + 1.0), 0.5 * gw
    nd1 = p + 1
    ej, ei = np.meshgrid(np.arange(n1), np.arange(n0), indexing="ij")
    ei, ej = ei.ravel(), ej.ravel()                     # i fastest
    NF, ndf, nq = len(ei), nd1 * nd1, nq1 * nq1
    nodes = np.empty((NF, ndf), np.int64)
    a = 0
    for jj in range(nd1):
        for ii in range(nd1):
            nodes[:, a] = gtop[ei + ii, ej + jj]
            a += 1
    fv = np.concatenate([3 * nodes + c for c in range(3)], 1).astype(np.int32)
    bx = [bspline_basis_1d(p, knots[0], s_, gx) for s_ in range(n0)]
    by = [bspline_basis_1d(p, knots[1], s_, gx) for s_ in range(n1)]
    N = np.empty((NF, nq, ndf))
    dN = np.empty((NF, nq, 2, ndf))
    for f in range(NF):
        Nx, dNx = bx[ei[f]]
        Ny, dNy = by[ej[f]]
        N[f] = np.einsum("xi,yj->yxji", Nx, Ny).reshape(nq, ndf)
        dN[f, :, 0] = np.einsum("xi,yj->yxji", dNx, Ny).reshape(nq, ndf)
        dN[f, :, 1] = np.einsum("xi,yj->yxji", Nx, dNy).reshape(nq, ndf)
    w = np.tile(np.outer(gw, gw).reshape(1, nq), (NF, 1))
    return dict(v_dofs=fv, N=N, dN_dxi=dN, w=w, NF=NF, ndf=ndf, nqf=nq)
