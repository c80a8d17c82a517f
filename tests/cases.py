"""Shared helpers: golden fixtures (tests/golden/*.npz, written by oracle/make_golden.py from the
unmodified reference) and live runs of oracle/_ref/mimi_ref on generated meshes."""
import os
import tempfile

import numpy as np

from mimi_b200 import meshgen
from oracle import oracle_py as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# rational / curved / non-affine NURBS geometry (quarter annulus and thick quarter cylinder with weights 1, sqrt(1/2), 1; the
# reference's own non-affine tests/data/square-nurbs.mesh): oracle/make_golden.py curved_main. They also carry the patch as
# MFEM holds it (p, ncp, knots0.., el_ijk, node_w) and MFEM's setup forms (bodyforce, viscform, rhs_bf_traction, t_*).
CURVED_CASES = ["annulus_p2_n2", "cylinder_p2_n2", "square_nurbs_p2_n4"]
GOLDEN_CASES = ["cube_p2_n2", "cube_p1_n2", "beam2d_p3_n2", "balken_p3_n2", "quad_p2_n4", "cube_p2_n3_patch", "cube_p3_n2"] + CURVED_CASES + ["beam4_p3", "beam3_p2"]


# multi-patch beams dumped from the reference (oracle/make_golden.py multipatch_main): MFEM's numbering across patch interfaces
MULTIPATCH_CASES = ["beam4_p3", "beam3_p2"]


def patches_of(D):
    """[(element range e0, e1), ncp, knots] per patch of a multi-patch fixture (elements are ordered patch by patch)."""
    out = []
    for q in range(int(D["n_patches"])):
        idx = np.nonzero(D["el_patch"] == q)[0]
        assert idx.size and np.array_equal(idx, np.arange(idx[0], idx[0] + idx.size))
        out.append(((int(idx[0]), int(idx[-1]) + 1), D["ncp_%d" % q], [D["knots_%d_%d" % (q, d)] for d in range(D["dim"])]))
    return out


def patch_description(D):
    """(p, ncp, knots, el_ijk, node weights) of a curved fixture."""
    return int(D["p"]), D["ncp"], [D["knots%d" % d] for d in range(D["dim"])], D["el_ijk"], D["node_w"]
MATS = {"neohookean": 0, "stvk": 1, "j2linear": 2}


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    D = {k: z[k] for k in z.files}
    dim, nd, nq, NE, NV, nnz = (int(x) for x in D["meta"])
    D.update(dim=dim, nd=nd, nq=nq, NE=NE, NV=NV, nnz=nnz, tdof=dim * nd, gf=float(D["gf"]))
    return D


CONTACT_CASES = ["contact_cube_p2_n2_plane", "contact_cube_p2_n4_sphere", "contact_quad_p3_n4_sphere",
                 "contact_quad_p2_n4_plane_clamped"]


# rigid bodies = the reference's NearestDistanceToSplines on REAL splinepy splines (closest point by splinepy's own search)
SPLINE_CONTACT_CASES = ["contact_quad_p2_n4_bezier", "contact_cube_p2_n4_bezier", "contact_quad_p3_n4_rational_arc",
                        "contact_quad_p2_n4_bspline"]


def load_contact_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    D = {k: z[k] for k in z.files}
    dim, NF, ndf, nqf, NV, nnz, nm, nd, nq = (int(x) for x in D["meta"])
    D.update(dim=dim, NF=NF, ndf=ndf, nqf=nqf, NV=NV, nnz=nnz, n_marked_nodes=nm, nd=nd, nq=nq, gf=float(D["gf"]),
             penalty=float(D["penalty"]), body=(str(D["body_kind"]), [float(x) for x in D["body_params"]]))
    return D


J2_CASES = ["j2_cube_p2_n2", "j2_quad_p2_n4"]
J2_LAWS = ["powerlaw", "voce", "jc", "jcrate", "jcthermo", "jcconst"]
SIMO_LAWS = ["voce", "jcthermo"]
LOG_LAWS = ["voce", "jcthermo"]


def load_j2_golden(name):
    """Results of J2 + every hardening law on the tables of the base fixture."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    D = load_golden(str(z["base"]))
    D.update({k: z[k] for k in z.files if k != "base"})
    D["gf"] = float(z["gf"])
    return D


def j2_params(law, hp=None):
    """Parameter list of mb2_set_material(MB2_MAT_J2, ...) for the harness' material 'j2_<law>' (hp: override of the
    hardening parameters, oracle layout)."""
    lam, mu = O.lame()
    K = 2100.0 / (3.0 * (1.0 - (2.0 * 0.3)))
    T = O.J2_THERMO
    hp = (O.hardening_params(law) if hp is None else np.asarray(hp, dtype=np.float64)).copy()
    hp[6] = hp[8] = 0.0   # the library derives T_melt / temperature_contribution_ itself (J2::Setup)
    return [lam, mu, K, mu, O.HARD[law]] + list(hp) + [T["dt"], T["heat_fraction"], T["density"], T["specific_heat"],
                                                      T["T0"], T["melting"]]


SWEEP_CASES = ["sweeps_cube_p2_n2", "sweeps_quad_p2_n4"]


def load_sweeps_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    W = {k: z[k] for k in z.files if k != "base"}
    W["fac0"], W["fac1"] = float(W["fac0"]), float(W["fac1"])
    return load_golden(str(z["base"])), W


def ref_case(kind, p, n, material, gf=1.0):
    """Live reference run (nthreads=1). kind: 'refine' (p1 cube + DegreeElevate(p-1) + log2(n) refinements)
    or 'patch' (direct p, n spans)."""
    tmp = tempfile.mkdtemp(prefix="mb2case_")
    mesh = os.path.join(tmp, "m.mesh")
    if kind == "refine":
        meshgen.write_box_p1(mesh, (1.0, 1.0, 1.0))
        ref = int(round(np.log2(n)))
        assert 2 ** ref == n
        D = O.run_ref_dump(mesh, p - 1, ref, material, os.path.join(tmp, "d"), gf=gf, lite=True)
    else:
        meshgen.write_box_patch(mesh, p, (n, n, n))
        D = O.run_ref_dump(mesh, 0, 0, material, os.path.join(tmp, "d"), gf=gf, lite=True)
    return D


def setup_ctx(ctx, D, material):
    ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
    ctx.upload_quad_tables(D["dN_dX"], D["w"], D["detJ"])
    ctx.build_pattern()
    lam, mu = O.lame()
    if material == "j2linear":
        K = 2100.0 / (3.0 * (1.0 - (2.0 * 0.3)))  # MaterialBase::SetYoungPoisson, materials.cpp:7-14
        P = O.J2LINEAR_PARAMS
        ctx.set_material(MATS[material], [lam, mu, K, mu, P["iso"], P["kin"], P["sigma_y"]])
    else:
        ctx.set_material(MATS[material], [lam, mu])
    return ctx
