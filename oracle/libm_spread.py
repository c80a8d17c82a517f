#!/usr/bin/env python
"""Reference-vs-reference spread of the J2 family: the SAME compiled reference (oracle/_ref/mimi_ref) run twice, once
with glibc's default libm variants (the FMA / AVX2 ifunc implementations of pow / exp / log / ... this CPU selects) and
once with GLIBC_TUNABLES=glibc.cpu.hwcaps=-AVX2,-FMA,-FMA4 (the generic variants). Both are "the reference"; they differ
in the last bit of some transcendental results. What that does to the residual, the committed history, the FD tangent and
the Newton iteration counts of the reference's own beam test is the floor any other libm (CUDA's libdevice) can be held to.
Writes tests/golden/j2_libm_spread.npz (test infrastructure; needs /root/reference compiled into oracle/_ref)."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mimi_b200 import meshgen  # noqa: E402
from oracle import make_golden as G  # noqa: E402
from oracle import oracle_py as O  # noqa: E402

NOFMA = {"GLIBC_TUNABLES": "glibc.cpu.hwcaps=-AVX2,-FMA,-FMA4,-AVX512F"}


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def main():
    out = {}
    base = "cube_p2_n2"
    writer, elevate, refine, gf = G.CASES[base]
    mats = ["j2_" + law for law in O.HARD] + ["j2simo_" + law for law in G.SIMO_LAWS] + ["j2log_" + law for law in G.LOG_LAWS]
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        writer(mesh)
        for mat in mats:
            A = O.run_ref_dump(mesh, elevate, refine, mat, os.path.join(tmp, mat + "_a"), gf=gf)
            B = O.run_ref_dump(mesh, elevate, refine, mat, os.path.join(tmp, mat + "_b"), gf=gf, env=NOFMA)
            row = [rel(A[k], B[k]) for k in ("r", "A", "state", "r_b", "A_b")]
            out["spread_" + mat] = np.array(row)
            print("%-18s rel spread  r %.1e  A %.1e  state %.1e  rb %.1e  Ab %.1e" % (mat, *row))
        # the reference's own beam test (tests/test_nonlinear_solid.py there): Newton iteration counts of the two variants
        bm = os.path.join(tmp, "b.mesh")
        meshgen.write_box_p1(bm, (5.0, 1.0))
        for case, (mat, force, dt) in {"j2": ("j2_balken", -3.0, 0.5), "j2_simo": ("j2simo_balken", -3.0, 0.5),
                                       "j2_log": ("j2log_balken", -3.0, 0.5)}.items():
            xa, ia = O.run_ref_steps(bm, 2, 1, mat, force, dt, 10, os.path.join(tmp, case + "_a"))
            os.environ.update(NOFMA)
            try:
                xb, ib = O.run_ref_steps(bm, 2, 1, mat, force, dt, 10, os.path.join(tmp, case + "_b"))
            finally:
                os.environ.pop("GLIBC_TUNABLES")
            out["its_default_" + case], out["its_generic_" + case] = ia, ib
            out["x_spread_" + case] = np.float64(np.abs(xa - xb).max())
            print("%-8s Newton iterations default libm %s\n         generic libm       %s   max |x_a - x_b| = %.2e" % (
                case, ia.tolist(), ib.tolist(), np.abs(xa - xb).max()))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "j2_libm_spread.npz"), **out)


if __name__ == "__main__":
    main()
