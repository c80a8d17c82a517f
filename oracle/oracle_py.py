"""TEST INFRASTRUCTURE: ctypes loader for the plain-C oracle (oracle/mimi_oracle.c) and a runner for the
compiled reference (oracle/_ref/mimi_ref). Imported only by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs -- never by mimi_b200/."""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libmimi_oracle.so")
REF_BIN = os.path.join(HERE, "_ref", "mimi_ref")

NEOHOOKEAN, STVK = 0, 1


class Material(C.Structure):
    _fields_ = [("kind", C.c_int), ("lam", C.c_double), ("mu", C.c_double), ("K", C.c_double), ("G", C.c_double),
                ("iso", C.c_double), ("kin", C.c_double), ("sigma_y", C.c_double), ("sqrt_6", C.c_double),
                ("sqrt_3_2", C.c_double), ("sqrt_2_3", C.c_double), ("state", C.c_void_p), ("ns", C.c_int),
                ("hard_kind", C.c_int), ("hp", C.c_double * 10), ("dt", C.c_double), ("heat_fraction", C.c_double),
                ("density", C.c_double), ("specific_heat", C.c_double), ("error", C.POINTER(C.c_int))]


J2LINEAR = 2
J2LINEAR_NS = 19
# J2Linear test parameters used by oracle/ref_harness.cpp (material "j2linear")
J2LINEAR_PARAMS = dict(iso=80.0, kin=40.0, sigma_y=2.0)


def mat_j2linear(state, young=2100.0, poisson=0.3, iso=80.0, kin=40.0, sigma_y=2.0):
    """state: float64 array [NE, nq, 19] (modified in place by post_time_advance)."""
    lam, mu = lame(young, poisson)
    K = young / (3.0 * (1.0 - (2.0 * poisson)))     # MaterialBase::SetYoungPoisson, materials.cpp:7-14
    m = Material(J2LINEAR, lam, mu, K, mu, iso, kin, sigma_y, np.sqrt(6.0), np.sqrt(3.0 / 2.0), np.sqrt(2.0 / 3.0),
                 state.ctypes.data, J2LINEAR_NS, 0, (C.c_double * 10)(), 0, 0, 0, 0, None)
    m._keep = state
    return m


J2 = 3
J2_NS = 11
HARD = {"powerlaw": 0, "voce": 1, "jc": 2, "jcrate": 3, "jcthermo": 4, "jcconst": 5}
# the parameter sets oracle/ref_harness.cpp uses for material "j2_<law>"; hp layout of oracle/mimi_oracle.c
J2_THERMO = dict(melting=1500.0, T0=20.0, specific_heat=450.0, heat_fraction=0.9, dt=0.5, density=1.0)


def hardening_params(law):
    hp = np.zeros(10)
    if law == "powerlaw":
        hp[:3] = [2.0, 3.0, 0.01]
    elif law == "voce":
        hp[:3] = [2.0, 3.0, 0.05]
    else:
        hp[:3] = [2.0, 4.0, 0.2835]
        if law != "jc":
            hp[3:5] = [0.015, 1.0e-5]
        if law in ("jcthermo", "jcconst"):
            hp[5:8] = [20.0 if law == "jcthermo" else 10.0, J2_THERMO["melting"], 1.3558]
        if law == "jcconst":
            # JohnsonCookConstantTemperatureHardening::SetTemperature (material_hardening.hpp:310-321)
            hp[8] = 1.0 - np.power((J2_THERMO["T0"] - hp[5]) / (hp[6] - hp[5]), hp[7])
    return hp


def j2_initial_state(NE, nq):
    """J2::CreateState (materials.cpp:151-166): zero plastic strain, zero eqps, initial temperature."""
    S = np.zeros((NE, nq, J2_NS))
    S[:, :, 10] = J2_THERMO["T0"]
    return S


J2SIMO = 4
J2SIMO_NS = 20


def j2simo_initial_state(NE, nq, dim):
    """J2Simo::CreateState (materials.cpp:184-203): be_old = F_old = I, zero eqps, initial temperature."""
    S = np.zeros((NE, nq, J2SIMO_NS))
    for k in range(dim):
        S[:, :, k * (dim + 1)] = 1.0
        S[:, :, 9 + k * (dim + 1)] = 1.0
    S[:, :, 19] = J2_THERMO["T0"]
    return S


J2LOG = 5
J2LOG_NS = 11


def j2log_initial_state(NE, nq, dim):
    """J2Log::CreateState (materials.cpp:221-240): Fp_inv = I, zero eqps, initial temperature."""
    S = np.zeros((NE, nq, J2LOG_NS))
    for k in range(dim):
        S[:, :, k * (dim + 1)] = 1.0
    S[:, :, 10] = J2_THERMO["T0"]
    return S


def eig_sym(A):
    """mfem::DenseMatrix::CalcEigenvalues on a column-major dim x dim symmetric matrix -> (lambda, vectors[k] = k-th)."""
    A = np.ascontiguousarray(A, dtype=np.float64)
    dim = A.shape[0]
    lam = np.zeros(dim)
    vec = np.zeros(dim * dim)
    lib().orc_eig_sym(C.c_int(dim), _p(np.ascontiguousarray(A.T).ravel(), C.c_double), _p(lam, C.c_double), _p(vec, C.c_double))
    return lam, vec.reshape(dim, dim)


def mat_j2(state, law, young=2100.0, poisson=0.3, simo=False, log=False, hp=None):
    """hp: hardening parameters (oracle hp[] layout) overriding the harness' set for `law`."""
    lam, mu = lame(young, poisson)
    K = young / (3.0 * (1.0 - (2.0 * poisson)))
    err = C.c_int(0)
    m = Material(J2LOG if log else (J2SIMO if simo else J2), lam, mu, K, mu, 0, 0, 0, 0, np.sqrt(3.0 / 2.0), 0,
                 state.ctypes.data, J2LOG_NS if log else (J2SIMO_NS if simo else J2_NS), HARD[law],
                 (C.c_double * 10)(*(hardening_params(law) if hp is None else hp)), J2_THERMO["dt"], J2_THERMO["heat_fraction"],
                 J2_THERMO["density"], J2_THERMO["specific_heat"], C.pointer(err))
    m._keep = (state, err)
    return m


def post_time_advance(dim, nd, nq, m, v_dofs, dN_dX, u):
    rc = lib().orc_post_time_advance(C.c_int(dim), C.c_int(nd), C.c_int(nq), C.c_int64(v_dofs.shape[0]), C.byref(m),
                                     _p(v_dofs, C.c_int32), _p(dN_dX, C.c_double), _p(u, C.c_double))
    assert rc == 0


def lame(young=2100.0, poisson=0.3):
    """MaterialBase::SetYoungPoisson, src/mimi/materials/materials.cpp:7-14 (same expression order)."""
    lam = young * poisson / ((1 + poisson) * (1 - 2 * poisson))
    mu = young / (2.0 * (1.0 + poisson))
    return lam, mu


def build_oracle(force=False):
    srcs = [os.path.join(HERE, x) for x in ("mimi_oracle.c", "mimi_oracle_contact.c", "mimi_oracle_eig.c", "mimi_oracle_forms.c")]
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < max(os.path.getmtime(x) for x in srcs):
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", LIB] + srcs + ["-lm"])
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build_oracle()
        _lib = C.CDLL(LIB)
        _lib.orc_build_pattern.restype = C.c_int64
        _lib.orc_contact_gap_norm.restype = C.c_double
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def mat(kind, young=2100.0, poisson=0.3):
    lam, mu = lame(young, poisson)
    return Material(kind, lam, mu, 0, 0, 0, 0, 0, 0, 0, 0, None, 0, 0, (C.c_double * 10)(), 0, 0, 0, 0, None)


def assemble_residual(dim, nd, nq, m, v_dofs, dN_dX, w, detJ, u):
    NE = v_dofs.shape[0]
    r = np.zeros_like(u)
    rc = lib().orc_assemble_residual(C.c_int(dim), C.c_int(nd), C.c_int(nq), C.c_int64(NE), C.c_int64(u.size),
                                     C.byref(m), _p(v_dofs, C.c_int32), _p(dN_dX, C.c_double), _p(w, C.c_double),
                                     _p(detJ, C.c_double), _p(u, C.c_double), _p(r, C.c_double))
    assert rc == 0
    return r


def assemble_residual_and_grad(dim, nd, nq, m, v_dofs, A_ids, dN_dX, w, detJ, u, gf, A0):
    NE = v_dofs.shape[0]
    r = np.zeros_like(u)
    A = A0.copy()
    rc = lib().orc_assemble_residual_and_grad(C.c_int(dim), C.c_int(nd), C.c_int(nq), C.c_int64(NE), C.c_int64(u.size),
                                              C.c_int64(A.size), C.byref(m), _p(v_dofs, C.c_int32),
                                              _p(A_ids, C.c_int32), _p(dN_dX, C.c_double), _p(w, C.c_double),
                                              _p(detJ, C.c_double), _p(u, C.c_double), C.c_double(gf),
                                              _p(r, C.c_double), _p(A, C.c_double))
    assert rc == 0
    return r, A


def build_pattern(tdof, v_dofs, NV):
    NE = v_dofs.shape[0]
    I = np.zeros(NV + 1, np.int32)
    nnz = lib().orc_build_pattern(C.c_int(tdof), C.c_int64(NE), C.c_int64(NV), _p(v_dofs, C.c_int32),
                                  _p(I, C.c_int32), None)
    J = np.zeros(nnz, np.int32)
    lib().orc_build_pattern(C.c_int(tdof), C.c_int64(NE), C.c_int64(NV), _p(v_dofs, C.c_int32), _p(I, C.c_int32),
                            _p(J, C.c_int32))
    return I, J


def build_A_ids(tdof, v_dofs, I, J):
    NE = v_dofs.shape[0]
    A_ids = np.zeros((NE, tdof * tdof), np.int32)
    lib().orc_build_A_ids(C.c_int(tdof), C.c_int64(NE), _p(v_dofs, C.c_int32), _p(I, C.c_int32), _p(J, C.c_int32),
                          _p(A_ids, C.c_int32))
    return A_ids


# ---------------------------------------------------------------- compiled reference (oracle/_ref)
def have_ref():
    return os.path.exists(REF_BIN) and os.access(REF_BIN, os.X_OK)


def load_dump(d):
    meta = json.load(open(os.path.join(d, "meta.json")))
    NE, nd, nq, dim, tdof = meta["NE"], meta["nd"], meta["nq"], meta["dim"], meta["tdof"]
    out = dict(meta)

    def rd(name, dt, shape=None):
        p = os.path.join(d, name)
        if not os.path.exists(p):
            return None
        a = np.fromfile(p, dtype=dt)
        return a.reshape(shape) if shape else a

    out["I"] = rd("I.i32", np.int32)
    out["J"] = rd("J.i32", np.int32)
    out["v_dofs"] = rd("v_dofs.i32", np.int32, (NE, tdof))
    out["A_ids"] = rd("A_ids.i32", np.int32, (NE, tdof * tdof))
    out["w"] = rd("w.f64", np.float64, (NE, nq))
    out["detJ"] = rd("detJ.f64", np.float64, (NE, nq))
    out["dN_dX"] = rd("dN_dX.f64", np.float64, (NE, nq, dim, nd))
    out["dN_dxi"] = rd("dN_dxi.f64", np.float64, (NE, nq, dim, nd))
    out["N"] = rd("N.f64", np.float64, (NE, nq, nd))
    out["state"] = rd("state.f64", np.float64)
    if out["state"] is not None:
        out["state"] = out["state"].reshape(NE, nq, -1)
    pm = rd("p_meta.i32", np.int32)     # the patches as MFEM holds them (degree, ncp, knots per patch; el_patch, el_ijk, weights)
    if pm is not None:
        kn = rd("p_knots.f64", np.float64)
        NP = int(pm[2])
        out["p"], out["n_patches"] = int(pm[0]), NP
        out["patch_ncp"], out["patch_knots"] = [], []
        o, k = 3, 0
        for _ in range(NP):
            out["patch_ncp"].append(pm[o:o + dim].astype(np.int32))
            sizes = pm[o + dim:o + 2 * dim]
            out["patch_knots"].append([kn[k + int(sizes[:d].sum()):k + int(sizes[:d + 1].sum())].copy() for d in range(dim)])
            k += int(sizes.sum())
            o += 2 * dim
        out["ncp"], out["knots"] = out["patch_ncp"][0], out["patch_knots"][0]      # single-patch shorthand
        out["el_ijk"] = rd("el_ijk.i32", np.int32, (NE, dim))
        out["el_patch"] = rd("el_patch.i32", np.int32)
        out["node_w"] = rd("node_w.f64", np.float64)
    out["t_meta"] = rd("t_meta.i32", np.int32)
    out["t_v_dofs"] = rd("t_v_dofs.i32", np.int32)
    for k in ("x_ref", "u", "r", "r2", "mass", "bodyforce", "viscform", "mass_b200", "viscform_b200", "bodyforce_b200", "rhs_b200", "rhs_bf_traction", "t_w", "t_detW", "t_N", "A", "r_b", "A_b", "r_mg", "r2_mg", "A_mg"):
        out[k] = rd(k + ".f64", np.float64)
    return out


REF_B200_BIN = os.path.join(HERE, "_ref", "mimi_ref_b200")   # the harness with adapter/b200_nonlinear_solid.hpp


def have_ref_b200():
    return os.path.exists(REF_B200_BIN) and os.access(REF_B200_BIN, os.X_OK)


def run_ref_dump(mesh, elevate, refine, material, outdir, gf=1.0, nthreads=1, u_file="-", lite=False, binary=None, env=None):
    os.makedirs(outdir, exist_ok=True)
    cmd = [binary or REF_BIN, "dump", mesh, str(elevate), str(refine), material, str(nthreads), repr(gf), u_file, outdir]
    if lite:
        cmd.append("lite")
    subprocess.check_call(cmd, stdout=subprocess.DEVNULL, env=None if env is None else {**os.environ, **env})
    return load_dump(outdir)


def run_ref_time(mesh, elevate, refine, material, nthreads, reps, warmup=0):
    cmd = [REF_BIN, "time", mesh, str(elevate), str(refine), material, str(nthreads), str(reps), str(warmup)]
    # torch.distributed.run exports OMP_NUM_THREADS=1 to its children; the reference sizes its own OpenMP team
    # (py_solid.cpp:303-318) and the harness does too, but the environment must not cap it either
    env = {**os.environ, "OMP_NUM_THREADS": str(nthreads), "OMP_DYNAMIC": "false"}
    p = subprocess.run(cmd, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, check=True, text=True, env=env)
    return json.loads(p.stderr.strip().splitlines()[-1])


def element_matrices(dim, nd, nq, m, v_dofs, dN_dX, w, detJ, u, elems):
    """Per-element residual R_e [n, tdof] and FD tangent K_e [n, tdof(col i), tdof(row j)] (K_e(j, i) column-major,
    as ElementResidualAndGrad fills it) for the listed elements."""
    tdof = dim * nd
    n = len(elems)
    R = np.zeros((n, tdof))
    K = np.zeros((n, tdof, tdof))
    Rf = np.zeros(tdof)
    L = lib()
    for x, e in enumerate(elems):
        ue = np.ascontiguousarray(u[v_dofs[e]])
        L.orc_element_residual_and_grad(C.c_int(dim), C.c_int(nd), C.c_int(nq), C.byref(m), _p(ue, C.c_double),
                                        _p(np.ascontiguousarray(dN_dX[e]), C.c_double),
                                        _p(np.ascontiguousarray(w[e]), C.c_double),
                                        _p(np.ascontiguousarray(detJ[e]), C.c_double), _p(R[x], C.c_double),
                                        _p(Rf, C.c_double), _p(K[x], C.c_double))
    return R, K


# ---------------------------------------------------------------- mortar contact (oracle/mimi_oracle_contact.c)
RIGID_PLANE, RIGID_SPHERE = 0, 1


class Rigid(C.Structure):
    _fields_ = [("kind", C.c_int), ("dim", C.c_int), ("p", C.c_double * 16), ("deg_u", C.c_int), ("deg_v", C.c_int),
                ("n_seed", C.c_int), ("cp", C.c_void_p), ("n_patches", C.c_int), ("rational", C.c_int)]


RIGID_BEZIER = 2


def rigid_bezier(dim, cps, n_seed=None):
    """cps: curve [deg_u+1, 2] or surface [deg_v+1, deg_u+1, 3] (u fastest). PARITY UNPINNED (see the C file)."""
    cps = np.ascontiguousarray(cps, np.float64)
    if dim == 2:
        deg_u, deg_v = cps.shape[0] - 1, 0
    else:
        deg_v, deg_u = cps.shape[0] - 1, cps.shape[1] - 1
    n_seed = n_seed or (64 if dim == 2 else 16)
    r = Rigid(RIGID_BEZIER, dim, (C.c_double * 16)(), deg_u, deg_v, n_seed, cps.ctypes.data, 1)
    r._keep = cps
    return r


def bspline_to_bezier(p, knots, P):
    """Bezier extraction of a B-spline curve by Boehm knot insertion (every interior knot raised to multiplicity p):
    P [n, dim] -> patches [n_spans, p+1, dim]. The same formula and order as mb2_contact.cu (bspline_extract):
    Q_i = alpha P_i + (1 - alpha) P_{i-1}, alpha = (u - t_i) / (t_{i+p} - t_i)."""
    U = [float(x) for x in knots]
    P = [np.array(x, np.float64) for x in P]
    for u in sorted(set(U[p + 1:len(U) - p - 1])):
        while U.count(u) < p:
            k = max(i for i in range(len(U) - 1) if U[i] <= u < U[i + 1])
            Q = []
            for i in range(len(P) + 1):
                if i <= k - p:
                    Q.append(P[i])
                elif i > k:
                    Q.append(P[i - 1])
                else:
                    a = (u - U[i]) / (U[i + p] - U[i])
                    Q.append(a * P[i] + (1.0 - a) * P[i - 1])
            P = Q
            U.insert(k + 1, u)
    n_spans = (len(P) - 1) // p
    return np.array([[P[s * p + i] for i in range(p + 1)] for s in range(n_spans)])


def bspline_surface_to_bezier(pu, pv, ku, kv, P):
    """Tensor-product extraction: P [ncp_v, ncp_u, dim] (u fastest) -> patches [n_spans_v * n_spans_u, pv+1, pu+1, dim],
    rows first (u direction), then columns; patch order: v span slow, u span fast."""
    P = np.asarray(P, np.float64)
    rows = [bspline_to_bezier(pu, ku, P[j]) for j in range(P.shape[0])]           # [ncp_v][su][pu+1][dim]
    su = rows[0].shape[0]
    out = []
    cols = {}
    for a in range(su):
        for i in range(pu + 1):
            cols[(a, i)] = bspline_to_bezier(pv, kv, [rows[j][a][i] for j in range(P.shape[0])])   # [sv][pv+1][dim]
    sv = cols[(0, 0)].shape[0]
    for b in range(sv):
        for a in range(su):
            out.append([[cols[(a, i)][b][jj] for i in range(pu + 1)] for jj in range(pv + 1)])
    return np.array(out)


def rigid_bezier_patches(dim, patches, n_seed=None):
    """patches: curves [np, deg_u+1, 2] / surfaces [np, deg_v+1, deg_u+1, 3]."""
    patches = np.ascontiguousarray(patches, np.float64)
    r = rigid_bezier(dim, patches[0], n_seed)
    r.cp = patches.ctypes.data
    r.n_patches = patches.shape[0]
    r._keep = patches
    return r


def bezier_patches_params(dim, patches, n_seed=None):
    """mb2_contact_set_rigid_body(MB2_RIGID_BEZIER) with several patches: {deg_u, deg_v, n_seed, patch 0, patch 1, ...}."""
    b = rigid_bezier_patches(dim, patches, n_seed)
    return np.concatenate([[b.deg_u, b.deg_v, b.n_seed], b._keep.ravel()])


def bspline_params(dim, pu, pv, ku, kv, P, n_seed=None):
    """mb2_contact_set_rigid_body(MB2_RIGID_BSPLINE): {deg_u, deg_v, n_seed, ncp_u, ncp_v, knots_u, knots_v, control
    points [ncp_v][ncp_u][dim]} (curves: deg_v = 0, ncp_v = 1, no knots_v)."""
    P = np.ascontiguousarray(P, np.float64)
    n_seed = n_seed or (64 if dim == 2 else 16)
    ncp_u = P.shape[-2]
    ncp_v = P.shape[0] if dim == 3 else 1
    return np.concatenate([[pu, pv, n_seed, ncp_u, ncp_v], np.asarray(ku, np.float64),
                           np.asarray(kv, np.float64) if dim == 3 else [], P.ravel()])


def homogeneous(cps, weights):
    """Control points [..., dim] + weights [...] -> {w x, w} [..., dim + 1], the layout rational patches are queried in."""
    cps = np.asarray(cps, np.float64)
    w = np.asarray(weights, np.float64)
    return np.ascontiguousarray(np.concatenate([cps * w[..., None], w[..., None]], axis=-1))


def rigid_rational_patches(dim, patches_h, n_seed=None):
    """patches_h: homogeneous patches, curves [np, deg_u+1, 3] / surfaces [np, deg_v+1, deg_u+1, 4]."""
    patches_h = np.ascontiguousarray(patches_h, np.float64)
    r = rigid_bezier_patches(dim, patches_h, n_seed)
    r.rational = 1
    return r


def rational_bezier_params(dim, patches, weights, n_seed=None):
    """mb2_contact_set_rigid_body(MB2_RIGID_RATIONAL_BEZIER): {deg_u, deg_v, n_seed, patches of {x, w} per point}."""
    patches = np.asarray(patches, np.float64)
    w = np.asarray(weights, np.float64)
    b = rigid_bezier_patches(dim, patches, n_seed)
    return np.concatenate([[b.deg_u, b.deg_v, b.n_seed], np.concatenate([patches, w[..., None]], axis=-1).ravel()])


def nurbs_params(dim, pu, pv, ku, kv, P, weights, n_seed=None):
    """mb2_contact_set_rigid_body(MB2_RIGID_NURBS): the MB2_RIGID_BSPLINE list with {x, w} per control point."""
    P = np.asarray(P, np.float64)
    w = np.asarray(weights, np.float64)
    return bspline_params(dim, pu, pv, ku, kv, np.concatenate([P, w[..., None]], axis=-1), n_seed)


def bezier_params(dim, cps, n_seed=None):
    """Flat parameter list of mb2_contact_set_rigid_body(MB2_RIGID_BEZIER): {deg_u, deg_v, n_seed, control points}."""
    b = rigid_bezier(dim, cps, n_seed)
    return np.concatenate([[b.deg_u, b.deg_v, b.n_seed], b._keep.ravel()])


# ---- splinepy scenes of the reference harness (body "spline", oracle/ref_harness.cpp struct SplineDesc) -----------------
def spline_desc(resolution, cps, weights=None, knots=None, degrees=None):
    """Flat description the harness parses: resolution (kd-tree sampling per direction, nl_contact.py plant_kd_tree),
    rational, has_knots, deg_u, deg_v, ncp_u, ncp_v, [knots], control points [ncp_v][ncp_u][dim (+ weight)].
    cps: curve [ncp_u, 2] or surface [ncp_v, ncp_u, 3]."""
    cps = np.asarray(cps, np.float64)
    dim = cps.shape[-1]
    ncp = (cps.shape[0], 1) if dim == 2 else (cps.shape[1], cps.shape[0])
    if degrees is None:
        degrees = (ncp[0] - 1, ncp[1] - 1 if dim == 3 else 0)
    out = [resolution, int(weights is not None), int(knots is not None), degrees[0], degrees[1], ncp[0], ncp[1]]
    if knots is not None:
        for k in knots:
            out += [float(x) for x in k]
    pts = cps.reshape(-1, dim)
    w = None if weights is None else np.asarray(weights, np.float64).ravel()
    for i in range(pts.shape[0]):
        out += [float(x) for x in pts[i]]
        if w is not None:
            out.append(float(w[i]))
    return np.array(out, np.float64)


def parse_spline_desc(dim, desc):
    d = [float(x) for x in desc]
    res, rational, has_knots, du, dv, nu, nv = (int(x) for x in d[:7])
    k = 7
    knots = []
    if has_knots:
        for deg, n in ((du, nu), (dv, nv))[:dim - 1]:
            knots.append(np.array(d[k:k + n + deg + 1]))
            k += n + deg + 1
    n = nu * (nv if dim == 3 else 1)
    stride = dim + (1 if rational else 0)
    pts = np.array(d[k:k + n * stride]).reshape(n, stride)
    cps = pts[:, :dim].reshape((nu, dim) if dim == 2 else (nv, nu, dim))
    w = pts[:, dim].reshape(cps.shape[:-1]) if rational else None
    return dict(resolution=res, rational=bool(rational), knots=knots if has_knots else None, deg=(du, dv), ncp=(nu, nv), cps=cps,
                weights=w)


SEED_CAP = {2: 256, 3: 32}   # adapter/b200_mortar_contact.hpp B200NearestDistanceToSplines: max_seed_curve_ / _surface_


def _n_seed(dim, S):
    return max(1, min(S["resolution"] - 1, SEED_CAP[dim]))


def rigid_from_desc(dim, desc):
    """Oracle Rigid for a harness spline scene (Bezier extraction for splines with knots), seed sampling as the adapter
    derives it from the kd-tree resolution."""
    S = parse_spline_desc(dim, desc)
    ns = _n_seed(dim, S)
    cps = homogeneous(S["cps"], S["weights"]) if S["rational"] else S["cps"]
    if S["knots"] is None:
        patches = cps[None]
    elif dim == 2:
        patches = bspline_to_bezier(S["deg"][0], S["knots"][0], cps)
    else:
        patches = bspline_surface_to_bezier(S["deg"][0], S["deg"][1], S["knots"][0], S["knots"][1], cps)
    return rigid_rational_patches(dim, patches, ns) if S["rational"] else rigid_bezier_patches(dim, patches, ns)


def mb2_body_from_desc(dim, desc):
    """(kind, params) of mb2_contact_set_rigid_body for a harness spline scene: what B200NearestDistanceToSplines builds."""
    S = parse_spline_desc(dim, desc)
    ns = _n_seed(dim, S)
    pu, pv = S["deg"]
    if S["knots"] is None:
        if S["rational"]:
            return 4, rational_bezier_params(dim, S["cps"][None], S["weights"][None], ns)
        return 2, bezier_patches_params(dim, S["cps"][None], ns)
    ku = S["knots"][0]
    kv = S["knots"][1] if dim == 3 else []
    if S["rational"]:
        return 5, nurbs_params(dim, pu, pv, ku, kv, S["cps"], S["weights"], ns)
    return 3, bspline_params(dim, pu, pv, ku, kv, S["cps"], ns)


def run_ref_proximity(points, desc, nthreads=1):
    """NearestDistanceToSplines::NearestDistance (the reference's class on real splinepy, harness mode "proximity") on
    points [n, dim] -> dict of parametric, physical, pmq, distance, convergence, first_derivatives, normal, normal_gap."""
    import tempfile
    pts = np.ascontiguousarray(points, np.float64)
    n, dim = pts.shape
    pd = dim - 1
    with tempfile.TemporaryDirectory() as tmp:
        pts.tofile(os.path.join(tmp, "p.f64"))
        cmd = [REF_BIN, "proximity", os.path.join(tmp, "p.f64"), str(dim), str(n), "-", str(nthreads), tmp] + [repr(float(x)) for x in desc]
        subprocess.check_call(cmd, stdout=subprocess.DEVNULL)
        rec = pd + 2 * dim + 2 + pd * dim + dim + 1
        R = np.fromfile(os.path.join(tmp, "prox.f64")).reshape(n, rec)
    o = 0
    out = {}
    for name, w in (("parametric", pd), ("physical", dim), ("pmq", dim), ("distance", 1), ("convergence", 1),
                    ("first_derivatives", pd * dim), ("normal", dim), ("normal_gap", 1)):
        out[name] = R[:, o:o + w] if w > 1 else R[:, o]
        o += w
    out["first_derivatives"] = R[:, pd + 2 * dim + 2:pd + 2 * dim + 2 + pd * dim].reshape(n, pd, dim)
    return out


def rigid_query(body, q):
    q = np.ascontiguousarray(q, np.float64)
    xi, pmq, fd = np.zeros(2), np.zeros(3), np.zeros(6)
    dist = C.c_double()
    lib().orc_rigid_query(C.byref(body), _p(q, C.c_double), _p(xi, C.c_double), _p(pmq, C.c_double), C.byref(dist),
                          _p(fd, C.c_double))
    return xi, pmq[:body.dim], dist.value, fd[:(body.dim - 1) * body.dim].reshape(body.dim - 1, body.dim)


def rigid_plane(dim, o, t0, t1=None, lo=-1e300, hi=1e300):
    p = np.zeros(16)
    p[0:dim] = o
    p[3:3 + dim] = t0
    if t1 is not None:
        p[6:6 + dim] = t1
    p[9], p[10] = lo, hi
    return Rigid(RIGID_PLANE, dim, (C.c_double * 16)(*p), 0, 0, 0, None)


def rigid_sphere(dim, c, R):
    p = np.zeros(16)
    p[0:dim] = c
    p[3] = R
    return Rigid(RIGID_SPHERE, dim, (C.c_double * 16)(*p), 0, 0, 0, None)


def rigid_from_spec(dim, spec):
    """spec = ("plane", [o..., t0..., (t1...), lo, hi]) or ("sphere", [c..., R]) -- the harness' command line."""
    kind, v = spec
    v = list(v)
    if kind == "spline":
        return rigid_from_desc(dim, v)
    if kind == "plane":
        o, t0 = v[0:dim], v[dim:2 * dim]
        t1 = v[2 * dim:3 * dim] if dim == 3 else None
        lo, hi = v[dim * dim], v[dim * dim + 1]
        return rigid_plane(dim, o, t0, t1, lo, hi)
    return rigid_sphere(dim, v[0:dim], v[dim])


def rigid_params(dim, spec):
    """Flat parameter list of the C ABI (mb2_contact_set_rigid_body): the Rigid.p layout."""
    return np.array(list(rigid_from_spec(dim, spec).p))


def contact_prepare(dim, v_dofs):
    NF, tdf = v_dofs.shape
    ndf = tdf // dim
    local = np.zeros((NF, ndf), np.int32)
    marked = np.zeros(NF * tdf, np.int32)
    nm = C.c_int64()
    n = lib().orc_contact_prepare(C.c_int(dim), C.c_int64(NF), C.c_int(ndf), _p(v_dofs, C.c_int32), _p(local, C.c_int32),
                                  _p(marked, C.c_int32), C.byref(nm))
    return local, n, marked[:nm.value].copy()


def contact_pressure(dim, v_dofs, local, n_marked, N, dN_dxi, w, x_ref, u, body, penalty):
    NF, tdf = v_dofs.shape
    ndf, nqf = tdf // dim, w.shape[1]
    gap, area, pr = np.zeros(n_marked), np.zeros(n_marked), np.zeros(n_marked)
    tot = C.c_double(0.0)
    rc = lib().orc_contact_pressure(C.c_int(dim), C.c_int64(NF), C.c_int(ndf), C.c_int(nqf), C.c_int(n_marked),
                                    _p(v_dofs, C.c_int32), _p(local, C.c_int32), _p(N, C.c_double), _p(dN_dxi, C.c_double),
                                    _p(w, C.c_double), _p(x_ref, C.c_double), _p(u, C.c_double), C.byref(body),
                                    C.c_double(penalty), _p(gap, C.c_double), _p(area, C.c_double), _p(pr, C.c_double),
                                    C.byref(tot))
    assert rc == 0
    return gap, area, pr, tot.value


def contact_assemble(dim, v_dofs, local, A_ids, N, dN_dxi, w, x_ref, u, pressure, grad, gf, r, A):
    """Adds into r (and A when grad) in place; returns (force[3], pressure_integral)."""
    NF, tdf = v_dofs.shape
    ndf, nqf = tdf // dim, w.shape[1]
    force = np.zeros(3)
    pint = C.c_double(0.0)
    rc = lib().orc_contact_assemble(C.c_int(dim), C.c_int64(NF), C.c_int(ndf), C.c_int(nqf), _p(v_dofs, C.c_int32),
                                    _p(local, C.c_int32), _p(A_ids, C.c_int32) if A_ids is not None else None,
                                    _p(N, C.c_double), _p(dN_dxi, C.c_double), _p(w, C.c_double), _p(x_ref, C.c_double),
                                    _p(u, C.c_double), _p(pressure, C.c_double), C.c_int(int(grad)), C.c_double(gf),
                                    _p(r, C.c_double), _p(A, C.c_double) if A is not None else None,
                                    _p(force, C.c_double), C.byref(pint))
    assert rc == 0
    return force, pint.value


def contact_gap_norm(dim, v_dofs, N, w, x_ref, u, body):
    NF, tdf = v_dofs.shape
    return lib().orc_contact_gap_norm(C.c_int(dim), C.c_int64(NF), C.c_int(tdf // dim), C.c_int(w.shape[1]),
                                      _p(v_dofs, C.c_int32), _p(N, C.c_double), _p(x_ref, C.c_double), _p(u, C.c_double),
                                      C.byref(body))


def load_contact_dump(d):
    meta = json.load(open(os.path.join(d, "meta.json")))
    NF, ndf, nqf, dim = meta["NF"], meta["ndf"], meta["nqf"], meta["dim"]
    tdf = ndf * dim
    out = dict(meta)

    def rd(name, dt, shape=None):
        a = np.fromfile(os.path.join(d, name), dtype=dt)
        return a.reshape(shape) if shape else a

    out["v_dofs"] = rd("v_dofs.i32", np.int32, (meta["NE"], dim * meta["nd"]))
    out["marked"] = rd("c_marked.i32", np.int32)
    out["c_v_dofs"] = rd("c_v_dofs.i32", np.int32, (NF, tdf))
    out["c_A_ids"] = rd("c_A_ids.i32", np.int32, (NF, tdf * tdf))
    out["c_w"] = rd("c_w.f64", np.float64, (NF, nqf))
    out["c_N"] = rd("c_N.f64", np.float64, (NF, nqf, ndf))
    out["c_dN_dxi"] = rd("c_dN_dxi.f64", np.float64, (NF, nqf, dim - 1, ndf))
    out["c_x_ref"] = rd("c_x_ref.f64", np.float64, (NF, dim, ndf))
    out["c_marked_v_dofs"] = rd("c_marked_v_dofs.i32", np.int32)
    out["I"] = rd("I.i32", np.int32)
    out["J"] = rd("J.i32", np.int32)
    for k in ("x_ref", "u", "mass", "c_r0", "c_r", "c_gap", "c_area", "c_pressure", "c_r2", "c_A", "c_scalars"):
        out[k] = rd(k + ".f64", np.float64)
    return out


def run_ref_contact(mesh, elevate, refine, outdir, bdr_attr, penalty, body, gf=1.0, material="neohookean", u_file="-",
                    nthreads=1, binary=None):
    """body = ("plane", [...]) / ("sphere", [...]) / ("spline", spline_desc(...))."""
    os.makedirs(outdir, exist_ok=True)
    cmd = [binary or REF_BIN, "contact", mesh, str(elevate), str(refine), material, str(nthreads), repr(gf), u_file, outdir,
           str(bdr_attr), repr(float(penalty)), body[0]] + [repr(float(x)) for x in body[1]]
    subprocess.check_call(cmd, stdout=subprocess.DEVNULL)
    return load_contact_dump(outdir)


# The contact time loop of examples/nl_contact.py as the harness restates it (mode "steps_contact"): its configuration
NL_CONTACT = dict(verts=[(0.0, 0.0), (2.0, 0.0), (1.0, 1.0), (-1.0, 2.0)],       # tests/data/square-nurbs.mesh
                  elevate=1, refine=3, density=7e4, young=1e10, dt=0.001, penalty=1e11, resolution=10000,
                  cps=[[-2.45, 2.3], [0.35, 1.7], [0.75, 1.7], [1.55, 2.3]],      # sp.Bezier([3], ...) + [0.05, 1]
                  shift=(0.0, 0.005), bdr_contact=2, bdr_clamp=1)


def run_ref_steps_contact(mesh, outdir, nsteps, cfg=None, nthreads=1, binary=None):
    """-> x [steps, NV], Newton iterations [steps], contact history [steps, 6] (area, pressure, force xyz, gap norm)."""
    c = dict(NL_CONTACT)
    c.update(cfg or {})
    os.makedirs(outdir, exist_ok=True)
    desc = spline_desc(c["resolution"], np.array(c["cps"], np.float64))
    env = {**os.environ, "MIMI_REF_DENSITY": repr(float(c["density"])), "MIMI_REF_YOUNG": repr(float(c["young"]))}
    cmd = [binary or REF_BIN, "steps_contact", mesh, str(c["elevate"]), str(c["refine"]), "neohookean", str(nthreads),
           repr(float(c["dt"])), str(nsteps), outdir, str(c["bdr_contact"]), str(c["bdr_clamp"]), repr(float(c["penalty"]))]
    cmd += [repr(float(v)) for v in c["shift"]] + [repr(float(v)) for v in desc]
    subprocess.check_call(cmd, stdout=subprocess.DEVNULL, env=env)
    its = np.fromfile(os.path.join(outdir, "newton_its.i32"), np.int32)
    x = np.fromfile(os.path.join(outdir, "x_hist.f64")).reshape(nsteps, -1)
    ch = np.fromfile(os.path.join(outdir, "contact_hist.f64")).reshape(nsteps, 6)
    return x, its, ch


# ---------------------------------------------------------------- Newton-iteration sweeps
def spmv(I, J, A, x, y=None):
    out = np.zeros(len(I) - 1) if y is None else y
    lib().orc_spmv(C.c_int64(len(I) - 1), _p(I, C.c_int32), _p(J, C.c_int32), _p(A, C.c_double), _p(x, C.c_double),
                   _p(out, C.c_double), C.c_int(0 if y is None else 1))
    return out


def eliminate(ess, I, J, A, y):
    ess = np.ascontiguousarray(ess, np.int32)
    rc = lib().orc_eliminate(C.c_int64(ess.size), _p(ess, C.c_int32), _p(I, C.c_int32), _p(J, C.c_int32), _p(A, C.c_double),
                             _p(y, C.c_double))
    assert rc == 0


def newton_residual_and_grad(dim, nd, nq, m, D, W, contact=None):
    """NonlinearSolid::ResidualAndGrad (operators/nonlinear_solid.cpp:240-283) restated on the oracle's pieces.
    D: domain fixture (tables, A_ids, ...), W: dict(I, J, mass, visc|None, x, v, a, rhs|None, ess, fac0, fac1)."""
    fac0, fac1 = W["fac0"], W["fac1"]
    tx = W["x"] + fac0 * W["a"]
    y = spmv(W["I"], W["J"], W["mass"], W["a"])
    if W.get("visc") is not None:
        tv = W["v"] + fac1 * W["a"]
        spmv(W["I"], W["J"], W["visc"], np.ascontiguousarray(tv), y)
    r, Jm = assemble_residual_and_grad(dim, nd, nq, m, D["v_dofs"], D["A_ids"], D["dN_dX"], D["w"], D["detJ"],
                                       np.ascontiguousarray(tx), fac0, W["mass"])
    y = y + r            # the integrators add their (zero-started, ordered) thread-local sum onto y
    eliminate(W["ess"], W["I"], W["J"], Jm, y)
    if contact is not None:
        contact(np.ascontiguousarray(tx), fac0, y, Jm)
        eliminate(W["ess"], W["I"], W["J"], Jm, y)
    if W.get("visc") is not None:
        Jm += fac1 * W["visc"]
    if W.get("rhs") is not None:
        y += -1.0 * W["rhs"]
    y[W["ess"]] = 0.0
    return y, Jm, tx


def newton_residual(dim, nd, nq, m, D, W, contact=None):
    """NonlinearSolid::Mult(a, y) (operators/nonlinear_solid.cpp:172-205), the line search's residual-only evaluation,
    restated on the oracle's pieces (same W as newton_residual_and_grad)."""
    fac0, fac1 = W["fac0"], W["fac1"]
    tx = np.ascontiguousarray(W["x"] + fac0 * W["a"])
    y = spmv(W["I"], W["J"], W["mass"], W["a"])
    y = y + assemble_residual(dim, nd, nq, m, D["v_dofs"], D["dN_dX"], D["w"], D["detJ"], tx)
    y[W["ess"]] = 0.0
    if W.get("visc") is not None:
        spmv(W["I"], W["J"], W["visc"], np.ascontiguousarray(W["v"] + fac1 * W["a"]), y)
    if contact is not None:
        contact(tx, y)
        y[W["ess"]] = 0.0
    if W.get("rhs") is not None:
        y += -1.0 * W["rhs"]
    y[W["ess"]] = 0.0
    return y


def run_ref_sweeps(mesh, elevate, refine, material, outdir, fac0, fac1, nthreads=1):
    os.makedirs(outdir, exist_ok=True)
    cmd = [REF_BIN, "sweeps", mesh, str(elevate), str(refine), material, str(nthreads), repr(fac0), repr(fac1), outdir]
    subprocess.check_call(cmd, stdout=subprocess.DEVNULL)
    meta = json.load(open(os.path.join(outdir, "meta.json")))
    out = dict(fac0=fac0, fac1=fac1)
    for k in ("I", "J", "ess"):
        out[k] = np.fromfile(os.path.join(outdir, k + ".i32"), np.int32)
    for k in ("mass", "visc", "x", "v", "a", "rhs", "y", "Jac", "tx"):
        out[k] = np.fromfile(os.path.join(outdir, k + ".f64"), np.float64)
    assert out["y"].size == meta["NV"]
    return out


def run_ref_eig(mats, outdir):
    """mfem::DenseMatrix::CalcEigenvalues of the unmodified reference (harness mode 'eig') on mats [n][dim][dim]."""
    os.makedirs(outdir, exist_ok=True)
    n, dim, _ = mats.shape
    f = os.path.join(outdir, "in.f64")
    np.ascontiguousarray(np.transpose(mats, (0, 2, 1))).tofile(f)   # column-major
    subprocess.check_call([REF_BIN, "eig", f, str(dim), str(n), "-", "0", outdir, "-"], stdout=subprocess.DEVNULL)
    lam = np.fromfile(os.path.join(outdir, "lambda.f64")).reshape(n, dim)
    vec = np.fromfile(os.path.join(outdir, "vec.f64")).reshape(n, dim, dim)
    return lam, vec


def run_ref_steps(mesh, elevate, refine, material, force, dt, nsteps, outdir, binary=None, nthreads=1, rho_inf=None, lin=None):
    """Harness mode 'steps': the reference's own time loop (forms::Nonlinear + operators::NonlinearSolid + LineSearchNewton
    + GeneralizedAlpha2) on the clamped-beam set-up of its tests -> (x history [nsteps][NV], Newton iterations).
    lin="gmres": the reference's `use_iterative_solver` option (GMRESSolver + DSmoother on the CPU binary,
    adapter/b200_gmres_solver.hpp on the B200 binary) instead of the dense LU -> additionally the linear iteration count of
    every Newton iteration."""
    os.makedirs(outdir, exist_ok=True)
    env = dict(os.environ) if rho_inf is None else {**os.environ, "MIMI_REF_RHO_INF": repr(float(rho_inf))}
    if lin:
        env["MIMI_REF_LIN"] = lin
    subprocess.check_call([binary or REF_BIN, "steps", mesh, str(elevate), str(refine), material, str(nthreads), repr(force),
                           repr(dt), str(nsteps), outdir], stdout=subprocess.DEVNULL, env=env)
    its = np.fromfile(os.path.join(outdir, "newton_its.i32"), np.int32)
    x = np.fromfile(os.path.join(outdir, "x_hist.f64")).reshape(nsteps, -1)
    if lin == "gmres":
        return x, its, np.fromfile(os.path.join(outdir, "lin_its.i32"), np.int32)
    return x, its


def gmres_jacobi(I, J, A, b, rel_tol=1e-8, abs_tol=1e-12, max_iter=300, m=50):
    """mfem::GMRESSolver + DSmoother as the reference configures them (py_nonlinear_solid.cpp:330-339) -> (x, iterations,
    final norm, converged)."""
    n = len(I) - 1
    x = np.zeros(n)
    it, fn = C.c_int(), C.c_double()
    lib().orc_gmres_jacobi.restype = C.c_int
    conv = lib().orc_gmres_jacobi(C.c_int64(n), _p(I, C.c_int32), _p(J, C.c_int32), _p(A, C.c_double), _p(b, C.c_double),
                                  C.c_double(rel_tol), C.c_double(abs_tol), C.c_int(max_iter), C.c_int(m), _p(x, C.c_double),
                                  C.byref(it), C.byref(fn))
    return x, it.value, fn.value, bool(conv)


def run_ref_gmres(I, J, A, b, workdir):
    """Harness mode 'gmres': the compiled reference's GMRESSolver + DSmoother on the given CSR system."""
    os.makedirs(workdir, exist_ok=True)
    for name, arr, t in (("I.i32", I, np.int32), ("J.i32", J, np.int32), ("A.f64", A, np.float64), ("b.f64", b, np.float64)):
        np.ascontiguousarray(arr, t).tofile(os.path.join(workdir, name))
    subprocess.check_call([REF_BIN, "gmres", workdir, str(len(I) - 1), str(len(J)), "-", "0", workdir, "-"],
                          stdout=subprocess.DEVNULL)
    return np.fromfile(os.path.join(workdir, "x.f64")), np.fromfile(os.path.join(workdir, "info.f64"))


# ---------------------------------------------------------------- setup forms (oracle/mimi_oracle_forms.c)
def _knot_ptrs(knots):
    kv = [np.ascontiguousarray(k, np.float64) for k in knots]
    return kv, (C.c_void_p * len(kv))(*[k.ctypes.data for k in kv])


def form_element_matrices(kind, dim, p, v_dofs, el_ijk, knots, x_ref, coef=1.0, weights=None, quad_order=-1):
    """Scalar element blocks [NE, nd, nd] of mfem::VectorMassIntegrator (kind 0) / VectorDiffusionIntegrator (kind 2)."""
    vd = np.ascontiguousarray(v_dofs, np.int32)
    ijk = np.ascontiguousarray(el_ijk, np.int32)
    NE, nd = vd.shape[0], (p + 1) ** dim
    kv, arr = _knot_ptrs(knots)
    x = np.ascontiguousarray(x_ref, np.float64)
    w = None if weights is None else np.ascontiguousarray(weights, np.float64)
    out = np.empty((NE, nd, nd))
    rc = lib().orc_form_element_matrices(C.c_int(kind), C.c_int(dim), C.c_int(p), C.c_int64(NE), _p(vd, C.c_int32),
                                         _p(ijk, C.c_int32), arr, _p(x, C.c_double),
                                         None if w is None else _p(w, C.c_double), C.c_int(quad_order), C.c_double(coef),
                                         _p(out, C.c_double))
    assert rc == 0
    return out


def form_element_vectors(dim, p, v_dofs, el_ijk, knots, x_ref, force, weights=None, quad_order=-1):
    """Element vectors [NE, dim*nd] of mfem::VectorDomainLFIntegrator for a constant body force."""
    vd = np.ascontiguousarray(v_dofs, np.int32)
    ijk = np.ascontiguousarray(el_ijk, np.int32)
    NE, nd = vd.shape[0], (p + 1) ** dim
    kv, arr = _knot_ptrs(knots)
    x = np.ascontiguousarray(x_ref, np.float64)
    w = None if weights is None else np.ascontiguousarray(weights, np.float64)
    f = np.ascontiguousarray(force, np.float64)
    out = np.empty((NE, dim * nd))
    rc = lib().orc_form_element_vectors(C.c_int(dim), C.c_int(p), C.c_int64(NE), _p(vd, C.c_int32), _p(ijk, C.c_int32), arr,
                                        _p(x, C.c_double), None if w is None else _p(w, C.c_double), C.c_int(quad_order),
                                        _p(f, C.c_double), _p(out, C.c_double))
    assert rc == 0
    return out


def assemble_bilinear_form(dim, nd, v_dofs, I, J, Me):
    """BilinearForm::Assemble(0): element blocks added into the CSR in ascending element order, the same scalar block on
    every diagonal component block (AddSubMatrix also adds the zeros between components: no effect on the values)."""
    A = np.zeros(len(J))
    vd = np.asarray(v_dofs).reshape(len(Me), dim, nd)
    for e in range(len(Me)):
        for c in range(dim):
            rows = vd[e, c]
            for a in range(nd):
                r = rows[a]
                pos = I[r] + np.searchsorted(J[I[r]:I[r + 1]], rows)
                A[pos] += Me[e, a]
    return A


def assemble_linear_form(NV, v_dofs, be):
    """LinearForm::Assemble: element vectors added in ascending element order."""
    b = np.zeros(NV)
    for e in range(len(be)):
        b[np.asarray(v_dofs[e])] += be[e]      # the vdofs of one element are distinct
    return b


def add_traction_form(b, dim, face_vdofs, N, w, detW, traction):
    """mfem::VectorBoundaryLFIntegrator (lininteg.cpp:375-408) on precomputed face tables, added face by face onto b the
    way LinearForm::Assemble does after the domain integrators: vec = t; vec *= Tr.Weight() * ip.weight;
    elvect(ndf k + s) += vec(k) * shape(s)."""
    NF, nqf, ndf = N.shape
    for f in range(NF):
        el = np.zeros(dim * ndf)
        for q in range(nqf):
            c = detW[f, q] * w[f, q]
            for k in range(dim):
                el[k * ndf:(k + 1) * ndf] += (traction[k] * c) * N[f, q]
        b[face_vdofs[f]] += el
    return b
