"""TEST INFRASTRUCTURE. Regenerates tests/golden/*.npz by running the UNMODIFIED reference
(oracle/_ref/mimi_ref, built by `make -f oracle/Makefile.ref`) at nthreads=1 on small meshes.
Run from the repo root where /root/reference exists:  python oracle/make_golden.py
"""
import os
import shutil
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mimi_b200 import meshgen  # noqa: E402
from oracle import oracle_py as O  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

# name: (mesh writer, elevate, refine, gf)
CASES = {
    "cube_p2_n2": (lambda f: meshgen.write_box_p1(f, (1.0, 1.0, 1.0)), 1, 1, 1.0),
    "cube_p1_n2": (lambda f: meshgen.write_box_p1(f, (1.0, 1.0, 1.0)), 0, 1, 0.5),
    "beam2d_p3_n2": (lambda f: meshgen.write_box_p1(f, (5.0, 1.0)), 2, 1, 0.0125),  # tests/test_nonlinear_solid.py mesh
    # the reference's own test mesh (tests/test_nonlinear_solid.py:6-13: balken.mesh, elevate 2, subdivide 1)
    "balken_p3_n2": (lambda f: shutil.copy("/root/reference/tests/data/balken.mesh", f), 2, 1, 0.0625),
    "quad_p2_n4": (lambda f: meshgen.write_box_p1(f, (2.0, 1.0)), 1, 2, 1.0),
    "cube_p3_n2": (lambda f: meshgen.write_box_p1(f, (1.0, 1.0, 1.0)), 2, 1, 0.6),
    "cube_p2_n3_patch": (lambda f: meshgen.write_box_patch(f, 2, (3, 2, 2), (1.5, 1.0, 1.0)), 0, 0, 2.0),
}


# contact fixtures: name: (mesh writer, elevate, refine, gf, boundary attribute, penalty, rigid body spec)
CONTACT_CASES = {
    "contact_cube_p2_n2_plane": (lambda f: meshgen.write_box_p1(f, (1.0, 1.0, 1.0)), 1, 1, 0.7, 2, 1e5,
                                 ("plane", [0, 0, 1.0, 0, 1, 0, 1, 0, 0, -10, 10])),
    "contact_cube_p2_n4_sphere": (lambda f: meshgen.write_box_p1(f, (1.0, 1.0, 1.0)), 1, 2, 0.7, 2, 1e5,
                                  ("sphere", [0.5, 0.5, 1.9996, 1.0])),
    "contact_quad_p3_n4_sphere": (lambda f: meshgen.write_box_p1(f, (2.0, 1.0)), 2, 2, 0.3, 2, 1e5,
                                  ("sphere", [1.0, 1.9995, 1.0])),
    "contact_quad_p2_n4_plane_clamped": (lambda f: meshgen.write_box_p1(f, (2.0, 1.0)), 1, 2, 0.3, 2, 1e5,
                                         ("plane", [0.3, 1.0002, 1, 0, 0.0, 1.0])),
}


# Rigid bodies as the reference's own scenes hold them: NearestDistanceToSplines on REAL splinepy splines (closest point =
# splinepy's kd-tree seed + Newton -> SLSQP -> LM, third_party/splinepy/src/proximity/proximity.cpp:643-733).
def _spline_bodies():
    s2 = np.sqrt(0.5)
    # the cubic Bezier curve of examples/nl_contact.py:20-29 (control polygon shape), lowered onto the top of the 2 x 1 block
    curve = np.array([[-0.5, 1.3], [0.6, 0.9], [1.4, 0.9], [2.5, 1.3]]) + [0.0, -0.0003]
    # biquadratic Bezier surface bulging down onto the face z = 1 of the unit cube; rows reversed: normal = S_u x S_v points down
    g = np.array([-0.5, 0.5, 1.5])
    surf = np.array([[[g[i], g[j], 1.2 - 0.8012 * (i == 1) * (j == 1)] for i in range(3)] for j in range(3)])[::-1].copy()
    # rational quadratic arc = the bottom quarter of the circle of contact_quad_p3_n4_sphere (c = (1, 1.9995), R = 1)
    c, R = np.array([1.0, 1.9995]), 1.0
    arc = np.array([c + R * np.array([-s2, -s2]), c + np.array([0.0, -R / s2]), c + R * np.array([s2, -s2])])
    arc_w = np.array([1.0, s2, 1.0])
    # cubic B-spline curve with two interior knots (what toy_problem.py's extracted boundaries look like)
    bs = np.array([[-0.4, 1.25], [0.3, 1.05], [0.8, 0.9985], [1.2, 0.9985], [1.7, 1.05], [2.4, 1.25]]) - [0.0, 0.002]
    bs_k = [0, 0, 0, 0, 1 / 3, 2 / 3, 1, 1, 1, 1]
    return {
        "contact_quad_p2_n4_bezier": (lambda f: meshgen.write_box_p1(f, (2.0, 1.0)), 1, 2, 0.3, 2, 1e5,
                                      ("spline", O.spline_desc(200, curve))),
        "contact_cube_p2_n4_bezier": (lambda f: meshgen.write_box_p1(f, (1.0, 1.0, 1.0)), 1, 2, 0.7, 2, 1e5,
                                      ("spline", O.spline_desc(64, surf))),
        "contact_quad_p3_n4_rational_arc": (lambda f: meshgen.write_box_p1(f, (2.0, 1.0)), 2, 2, 0.3, 2, 1e5,
                                            ("spline", O.spline_desc(200, arc, weights=arc_w))),
        "contact_quad_p2_n4_bspline": (lambda f: meshgen.write_box_p1(f, (2.0, 1.0)), 1, 2, 0.3, 2, 1e5,
                                       ("spline", O.spline_desc(150, bs, knots=[bs_k], degrees=(3, 0)))),
    }


def contact_main():
    cases = dict(CONTACT_CASES)
    cases.update(_spline_bodies())
    for name, (writer, elevate, refine, gf, attr, penalty, spec) in cases.items():
        with tempfile.TemporaryDirectory() as tmp:
            mesh = os.path.join(tmp, "m.mesh")
            writer(mesh)
            D = O.run_ref_contact(mesh, elevate, refine, os.path.join(tmp, "c"), attr, penalty, spec, gf=gf)
            out = {k: v for k, v in D.items() if isinstance(v, np.ndarray)}
            out["meta"] = np.array([D["dim"], D["NF"], D["ndf"], D["nqf"], D["NV"], D["nnz"], D["n_marked_nodes"], D["nd"], D["nq"]],
                                   np.int64)
            out["gf"], out["penalty"] = np.float64(gf), np.float64(penalty)
            out["body_kind"] = np.array(spec[0])
            out["body_params"] = np.array(spec[1], np.float64)
            np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
            print(name, out["meta"])


def contact_steps_main(nsteps=12):
    """examples/nl_contact.py's time loop (harness mode steps_contact) on the CPU reference with real splinepy, plus the
    reference's OWN spread: the same loop at nthreads = 2, 3, 4 (the thread-ordered reduction changes the last bits,
    nonlinear_base.hpp:111-151) and with kd-tree resolutions 5000 / 20000 (another seed for the same closest point)."""
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        meshgen.write_cell_p1(mesh, O.NL_CONTACT["verts"])
        x, its, ch = O.run_ref_steps_contact(mesh, os.path.join(tmp, "o"), nsteps)
        variants, labels, spread = [its], ["nthreads=1"], [0.0]
        for nt in (2, 3, 4):
            xv, iv, _ = O.run_ref_steps_contact(mesh, os.path.join(tmp, "t%d" % nt), nsteps, nthreads=nt)
            variants.append(iv); labels.append("nthreads=%d" % nt)
            spread.append(float(np.abs(xv - x).max() / np.abs(x).max()))
        for res in (5000, 20000):
            xv, iv, _ = O.run_ref_steps_contact(mesh, os.path.join(tmp, "r%d" % res), nsteps, cfg=dict(resolution=res))
            variants.append(iv); labels.append("resolution=%d" % res)
            spread.append(float(np.abs(xv - x).max() / np.abs(x).max()))
    np.savez_compressed(os.path.join(OUT, "contact_steps_square.npz"), x=x, its=its, contact_hist=ch,
                        its_variants=np.array(variants), variant_labels=np.array(labels), x_rel_spread=np.array(spread))
    print("contact_steps_square: Newton iterations", its.tolist(), "max |u|", float(np.abs(x).max()))
    for lab, iv, sp in zip(labels, variants, spread):
        print("   %-18s %s  rel dx %.2e" % (lab, iv.tolist(), sp))


# J2 + hardening laws: results only (tables / connectivity are those of the named base fixture)
J2_CASES = {"j2_cube_p2_n2": "cube_p2_n2", "j2_quad_p2_n4": "quad_p2_n4"}


SIMO_LAWS = ("voce", "jcthermo")
LOG_LAWS = ("voce", "jcthermo")


def j2_main():
    for name, base in J2_CASES.items():
        writer, elevate, refine, gf = CASES[base]
        with tempfile.TemporaryDirectory() as tmp:
            mesh = os.path.join(tmp, "m.mesh")
            writer(mesh)
            out = {"base": np.array(base), "gf": np.float64(gf)}
            for law in O.HARD:
                D = O.run_ref_dump(mesh, elevate, refine, "j2_" + law, os.path.join(tmp, law), gf=gf)
                for k_out, k_in in (("r", "r"), ("r2", "r2"), ("A", "A"), ("state", "state"), ("rb", "r_b"), ("Ab", "A_b")):
                    out[k_out + "_" + law] = D[k_in]
            for law in SIMO_LAWS:  # J2Simo (finite strain) on two representative laws
                D = O.run_ref_dump(mesh, elevate, refine, "j2simo_" + law, os.path.join(tmp, "simo" + law), gf=gf)
                for k_out, k_in in (("r", "r"), ("r2", "r2"), ("A", "A"), ("state", "state"), ("rb", "r_b"), ("Ab", "A_b")):
                    out[k_out + "_simo_" + law] = D[k_in]
            for law in LOG_LAWS:  # J2Log (finite strain, logarithmic) on the same two laws
                D = O.run_ref_dump(mesh, elevate, refine, "j2log_" + law, os.path.join(tmp, "log" + law), gf=gf)
                for k_out, k_in in (("r", "r"), ("r2", "r2"), ("A", "A"), ("state", "state"), ("rb", "r_b"), ("Ab", "A_b")):
                    out[k_out + "_log_" + law] = D[k_in]
            np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
            print(name, sorted(out)[:4])


def eig_matrices(dim):
    """Symmetric test matrices for CalcEigenvalues: generic, near-identity (like C_e = F_e^T F_e), tiny deviatoric
    (like the plastic increment), and the degenerate inputs that take the solver's special branches."""
    rng = np.random.default_rng(20 + dim)
    mats = []
    for _ in range(150):
        B = rng.standard_normal((dim, dim))
        mats.append(B + B.T)
    for s in (1e-2, 1e-5, 1e-9):
        for _ in range(40):
            B = s * rng.standard_normal((dim, dim))
            mats.append(np.eye(dim) + B + B.T)
    for _ in range(40):
        B = 1e-4 * rng.standard_normal((dim, dim))
        B = B + B.T
        mats.append(B - np.trace(B) / dim * np.eye(dim))
    for _ in range(30):      # repeated eigenvalue, rotated
        Q, _r = np.linalg.qr(rng.standard_normal((dim, dim)))
        mats.append(Q @ np.diag([1.0, 1.0, 3.0][:dim]) @ Q.T)
    for _ in range(30):      # coordinate-plane coupling only
        A = np.diag(rng.standard_normal(dim))
        A[0, dim - 1] = A[dim - 1, 0] = rng.standard_normal()
        mats.append(A)
    mats += [np.eye(dim) * 2.5, np.zeros((dim, dim)), np.diag([1.0, 1.0, 3.0][:dim]), np.diag([3.0, 1.0, 2.0][:dim]),
             1e300 * np.eye(dim), 1e-300 * (np.ones((dim, dim)) + np.eye(dim)), np.ones((dim, dim))]
    return np.array(mats)


def eig_main():
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        for dim in (2, 3):
            mats = eig_matrices(dim)
            lam, vec = O.run_ref_eig(mats, os.path.join(tmp, str(dim)))
            out["A%d" % dim], out["lambda%d" % dim], out["vec%d" % dim] = mats, lam, vec
    np.savez_compressed(os.path.join(OUT, "eig_sym.npz"), **out)
    print("eig_sym", out["A2"].shape, out["A3"].shape)


HISTORY_RUNS = {"neohook": ("neohookean", -5.0, 0.05), "j2": ("j2_balken", -3.0, 0.5),
                "j2_simo": ("j2simo_balken", -3.0, 0.5), "j2_log": ("j2log_balken", -3.0, 0.5)}


def histories_main():
    """The reference's own golden vectors for this path: tests/data/ref/<case>_h1_p2/x_<i>.txt, the
    displacement after each of 10 generalized-alpha steps of the 2D p=3 beam (tests/test_nonlinear_solid.py:54-112).
    Packed unchanged into one fixture; tests/test_reference_histories.py replays them."""
    out = {}
    for key, folder in (("neohook", "neohook_h1_p2"), ("j2", "j2_h1_p2"), ("j2_simo", "j2_simo_h1_p2"), ("j2_log", "j2_log_h1_p2")):
        out[key] = np.array([np.genfromtxt("/root/reference/tests/data/ref/%s/x_%d.txt" % (folder, i)) for i in range(10)])
    # and what the compiled reference itself does on that set-up here (harness mode 'steps', dense LU for UMFPACK):
    # Newton iteration counts per step, the north_star's "identical Newton iteration counts" reference
    with tempfile.TemporaryDirectory() as tmp:
        for key, (mat, force, dt) in HISTORY_RUNS.items():
            x, its = O.run_ref_steps("/root/reference/tests/data/balken.mesh", 2, 1, mat, force, dt, 10, os.path.join(tmp, key))
            assert np.allclose(x, out[key]), key
            out["ref_x_" + key], out["ref_its_" + key] = x, its
    np.savez_compressed(os.path.join(OUT, "balken_histories.npz"), **out)
    print("balken_histories", {k: v.shape for k, v in out.items()})


def sweeps_main():
    """Newton-iteration sweeps (harness mode 'sweeps') on the meshes of two base fixtures."""
    for name, base, fac0, fac1 in (("sweeps_cube_p2_n2", "cube_p2_n2", 0.25, 0.5), ("sweeps_quad_p2_n4", "quad_p2_n4", 0.0625, 0.25)):
        writer, elevate, refine, gf = CASES[base]
        with tempfile.TemporaryDirectory() as tmp:
            mesh = os.path.join(tmp, "m.mesh")
            writer(mesh)
            W = O.run_ref_sweeps(mesh, elevate, refine, "neohookean", os.path.join(tmp, "s"), fac0, fac1)
            W["base"] = np.array(base)
            # the linear system of that Newton iteration, solved by the reference's iterative option (harness 'gmres')
            W["gmres_x"], W["gmres_info"] = O.run_ref_gmres(W["I"], W["J"], W["Jac"], W["y"], os.path.join(tmp, "g"))
            np.savez_compressed(os.path.join(OUT, name + ".npz"), **W)
            print(name, W["y"].size, W["Jac"].size, W["ess"].size)


BODY_FORCE = (0.3, -1.7, 0.9)      # the constant force oracle/ref_harness.cpp puts into its body-force linear form


def bodyforce_main():
    """tests/golden/forms.npz: mfem::VectorDomainLFIntegrator of a constant body force (harness mode 'dump',
    bodyforce.f64) and the mfem::VectorDiffusionIntegrator viscosity matrix (viscform.f64) for every base fixture; the meshes / numbering are those of the fixtures of the same name."""
    out = {"f": np.array(BODY_FORCE)}
    for name, (writer, elevate, refine, gf) in CASES.items():
        with tempfile.TemporaryDirectory() as tmp:
            mesh = os.path.join(tmp, "m.mesh")
            writer(mesh)
            D = O.run_ref_dump(mesh, elevate, refine, "neohookean", os.path.join(tmp, "d"), gf=gf)
            base = np.load(os.path.join(OUT, name + ".npz"))
            assert np.array_equal(D["mass"], base["mass"]) and np.array_equal(D["x_ref"], base["x_ref"])
            out[name] = D["bodyforce"]
            out["visc_" + name] = D["viscform"]         # VectorDiffusionIntegrator(0.37) on the fixture's pattern
            # body force + traction in one LinearForm, and the face tables of the loaded boundary at the rule 2p
            NF, ndf, nqf = (int(x) for x in D["t_meta"])
            out["rhs_" + name] = D["rhs_bf_traction"]
            out["tmeta_" + name] = D["t_meta"]
            out["tvd_" + name] = D["t_v_dofs"].reshape(NF, -1)
            out["tN_" + name] = D["t_N"].reshape(NF, nqf, ndf)
            out["tw_" + name] = D["t_w"].reshape(NF, nqf)
            out["tdet_" + name] = D["t_detW"].reshape(NF, nqf)
    out["viscosity"] = np.float64(0.37)
    out["traction"] = np.array([0.5, -0.25, 1.5])
    np.savez_compressed(os.path.join(OUT, "forms.npz"), **out)
    print("bodyforce", {k: v.shape for k, v in out.items()})


# Rational / curved / non-affine geometry: exact quarter annulus (weights 1, sqrt(1/2), 1), thick
# quarter cylinder, and the reference's own non-affine tests/data/square-nurbs.mesh. Each file carries the patch as MFEM
# holds it after refinement (knots, el_ijk, node weights) and MFEM's setup forms, so the device generators can be pinned.
CURVED_CASES = {
    "annulus_p2_n2": (lambda f: meshgen.write_quarter_annulus(f), 0, 1, 0.8),
    "cylinder_p2_n2": (lambda f: meshgen.write_quarter_annulus(f, height=1.0), 0, 1, 0.8),
    "square_nurbs_p2_n4": (lambda f: meshgen.write_cell_p1(f, meshgen.SQUARE_NURBS_VERTS), 1, 2, 0.5),
}


def curved_main():
    for name, (writer, elevate, refine, gf) in CURVED_CASES.items():
        with tempfile.TemporaryDirectory() as tmp:
            mesh = os.path.join(tmp, "m.mesh")
            writer(mesh)
            out = {}
            for mat in ("neohookean", "stvk", "j2linear"):
                D = O.run_ref_dump(mesh, elevate, refine, mat, os.path.join(tmp, mat), gf=gf)
                if not out:
                    for k in ("I", "J", "v_dofs", "A_ids", "w", "detJ", "dN_dX", "N", "x_ref", "u", "mass", "ncp", "el_ijk",
                              "node_w", "bodyforce", "viscform", "rhs_bf_traction", "t_meta", "t_v_dofs", "t_N", "t_w", "t_detW"):
                        out[k] = D[k]
                    for d, kv in enumerate(D["knots"]):
                        out["knots%d" % d] = kv
                    out["p"] = np.int64(D["p"])
                    out["meta"] = np.array([D["dim"], D["nd"], D["nq"], D["NE"], D["NV"], D["nnz"]], np.int64)
                    out["gf"] = np.float64(gf)
                    out["lambda_mu"] = np.array([float.fromhex(D["lambda"]), float.fromhex(D["mu"])])
                out["r_" + mat], out["r2_" + mat], out["A_" + mat] = D["r"], D["r2"], D["A"]
                if mat == "j2linear":
                    out["state_j2linear"], out["rb_j2linear"], out["Ab_j2linear"] = D["state"], D["r_b"], D["A_b"]
            np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
            print(name, {k: int(v) for k, v in zip(("dim", "nd", "nq", "NE", "NV", "nnz"), out["meta"])},
                  "weights in [%.4f, %.4f]" % (out["node_w"].min(), out["node_w"].max()))


def config1_main(nsteps=20):
    """BASELINE config 1: examples/nonlinear_solid.py as shipped -- balken.mesh (= the 5 x 1 one-element box), elevate 1,
    subdivide 2, NeoHookean (2100, 0.3, rho 1), clamped boundary 2, body force -5 on component 1, dt 0.05, Newton
    (1e-12, 1e-8, 10), generalized-alpha with the default ode_coefficient 0.25, nl.setup(2) -- through harness mode `steps`
    on the CPU reference at nthreads = 2 (as shipped) and 1."""
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        meshgen.write_box_p1(mesh, (5.0, 1.0))
        out = {}
        for nt in (1, 2):
            x, its = O.run_ref_steps(mesh, 1, 2, "neohookean", -5.0, 0.05, nsteps, os.path.join(tmp, "o%d" % nt), nthreads=nt, rho_inf=0.25)
            out["x_nt%d" % nt], out["its_nt%d" % nt] = x, its
            print("config1 nthreads=%d Newton iterations %s max|u| %.4f" % (nt, its.tolist(), np.abs(x).max()))
        # the same run with the reference's iterative linear solver (use_iterative_solver: GMRESSolver + DSmoother,
        # py_nonlinear_solid.cpp:330-339) and the GMRES iteration count of every Newton iteration
        x, its, lin = O.run_ref_steps(mesh, 1, 2, "neohookean", -5.0, 0.05, nsteps, os.path.join(tmp, "og"), nthreads=1, rho_inf=0.25, lin="gmres")
        out["x_gmres"], out["its_gmres"], out["lin_its_gmres"] = x, its, lin
        print("config1 GMRES: Newton iterations %s, linear iterations %s" % (its.tolist(), lin.tolist()))
    np.savez_compressed(os.path.join(OUT, "config1_steps.npz"), **out)


# Multi-patch meshes (BASELINE config 4: patches glued along x, shared interface nodes numbered by MFEM's NURBSExtension,
# precomputed.cpp:39-218 run across the patch interfaces): 4 x (2 x 1 x 1) p = 3 and 3 x (2 x 2 x 2) p = 2.
MULTIPATCH_CASES = {
    "beam4_p3": (lambda f: meshgen.write_beam_patches(f, 3, (2, 1, 1), 4, (1.0, 0.5, 0.5)), 0.7),
    "beam3_p2": (lambda f: meshgen.write_beam_patches(f, 2, (2, 2, 2), 3, (1.0, 1.0, 1.0)), 0.9),
}


def multipatch_main():
    for name, (writer, gf) in MULTIPATCH_CASES.items():
        with tempfile.TemporaryDirectory() as tmp:
            mesh = os.path.join(tmp, "m.mesh")
            writer(mesh)
            out = {}
            for mat in ("neohookean", "stvk"):
                D = O.run_ref_dump(mesh, 0, 0, mat, os.path.join(tmp, mat), gf=gf)
                if not out:
                    for k in ("I", "J", "v_dofs", "A_ids", "w", "detJ", "dN_dX", "N", "x_ref", "u", "mass", "el_ijk", "el_patch", "node_w"):
                        out[k] = D[k]
                    out["p"], out["n_patches"] = np.int64(D["p"]), np.int64(D["n_patches"])
                    for q, (ncp, kv) in enumerate(zip(D["patch_ncp"], D["patch_knots"])):
                        out["ncp_%d" % q] = ncp
                        for d, k in enumerate(kv):
                            out["knots_%d_%d" % (q, d)] = k
                    out["meta"] = np.array([D["dim"], D["nd"], D["nq"], D["NE"], D["NV"], D["nnz"]], np.int64)
                    out["gf"] = np.float64(gf)
                    out["lambda_mu"] = np.array([float.fromhex(D["lambda"]), float.fromhex(D["mu"])])
                out["r_" + mat], out["r2_" + mat], out["A_" + mat] = D["r"], D["r2"], D["A"]
            D = O.run_ref_dump(mesh, 0, 0, "j2linear", os.path.join(tmp, "j2linear"), gf=gf)
            out["r_j2linear"], out["r2_j2linear"], out["A_j2linear"] = D["r"], D["r2"], D["A"]
            out["state_j2linear"], out["rb_j2linear"], out["Ab_j2linear"] = D["state"], D["r_b"], D["A_b"]
            np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
            print(name, {k: int(v) for k, v in zip(("dim", "nd", "nq", "NE", "NV", "nnz"), out["meta"])}, "patches", int(out["n_patches"]))


def main(only=None):
    assert O.have_ref(), "build oracle/_ref first: make -f oracle/Makefile.ref -j8"
    os.makedirs(OUT, exist_ok=True)
    for name, (writer, elevate, refine, gf) in CASES.items():
        if only and name not in only:
            continue
        with tempfile.TemporaryDirectory() as tmp:
            mesh = os.path.join(tmp, "m.mesh")
            writer(mesh)
            out = {}
            for mat in ("neohookean", "stvk"):
                D = O.run_ref_dump(mesh, elevate, refine, mat, os.path.join(tmp, mat), gf=gf)
                if not out:
                    for k in ("I", "J", "v_dofs", "A_ids", "w", "detJ", "dN_dX", "N", "dN_dxi", "x_ref", "u", "mass"):
                        out[k] = D[k]
                    out["meta"] = np.array([D["dim"], D["nd"], D["nq"], D["NE"], D["NV"], D["nnz"]], np.int64)
                    out["gf"] = np.float64(gf)
                    out["lambda_mu"] = np.array([float.fromhex(D["lambda"]), float.fromhex(D["mu"])])
                out["r_" + mat] = D["r"]
                out["r2_" + mat] = D["r2"]
                out["A_" + mat] = D["A"]
            # material with history variables: virgin-state assembly, accumulated state after DomainPostTimeAdvance(u),
            # and the assembly at 1.5 u on that state
            D = O.run_ref_dump(mesh, elevate, refine, "j2linear", os.path.join(tmp, "j2linear"), gf=gf)
            out["r_j2linear"], out["r2_j2linear"], out["A_j2linear"] = D["r"], D["r2"], D["A"]
            out["state_j2linear"], out["rb_j2linear"], out["Ab_j2linear"] = D["state"], D["r_b"], D["A_b"]
            np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
            print(name, {k: int(v) for k, v in zip(("dim", "nd", "nq", "NE", "NV", "nnz"), out["meta"])})


if __name__ == "__main__":
    if "config1" in sys.argv[1:]:
        config1_main()
    elif "multipatch" in sys.argv[1:]:
        multipatch_main()
    elif "curved" in sys.argv[1:]:
        curved_main()
    elif "contact_steps" in sys.argv[1:]:
        contact_steps_main()
    elif "contact" in sys.argv[1:]:
        contact_main()
    elif "bodyforce" in sys.argv[1:]:
        bodyforce_main()
    elif "j2" in sys.argv[1:]:
        j2_main()
    elif "sweeps" in sys.argv[1:]:
        sweeps_main()
    elif "eig" in sys.argv[1:]:
        eig_main()
    elif "histories" in sys.argv[1:]:
        histories_main()
    elif len(sys.argv) > 1:
        main(sys.argv[1:])
    else:
        main()
        contact_main()
        j2_main()
        sweeps_main()
        eig_main()
        histories_main()
        contact_steps_main()
        curved_main()
        multipatch_main()
        config1_main()
