"""One Newton iteration of an implicit step at C3 scale (100^3 p=2 NeoHookean, NV = 3.18e6, nnz = 1.15e9) entirely on the
resident arrays: operator sweeps, GMRES(50) + Jacobi (rank 2), two line-search residuals (rank 4).
Mass (consistent, rho = 1) and body-force rhs are assembled on the device too (mb2_assemble_mass_device /
mb2_assemble_body_force_device), so J = M + fac0 K is what a generalized-alpha step with dt = 1e-3 sees and nothing but the
iterate leaves the host.
Prints one JSON line; timings are wall clock around synchronous C-ABI calls (each ends with a stream synchronize)."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import bench  # noqa: E402
from mimi_b200 import capi  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
ctx = capi.Context(0)
M = bench.build_problem(ctx, n, 2)
NV, nnz = M["NV"], M["nnz"]
x = M["u"]
ess = np.nonzero(np.repeat(M["x_ref"].reshape(-1, 3)[:, 0] == 0.0, 3))[0].astype(np.int32)
ctx.set_essential_dofs(ess)
# setup forms on the device (rank 4): consistent mass (rho = 1) with FormSystemMatrix's elimination, body-force rhs
t0 = time.perf_counter()
ctx.set_A(None)
ctx.assemble_mass_device(2, M["ncp"], M["knots"], M["el_ijk"], density=1.0)
ctx.newton_mass_from_A(True)
ctx.sync()
t1 = time.perf_counter()
ctx.set_r(None)
ctx.assemble_body_force_device(2, M["ncp"], M["knots"], M["el_ijk"], [0.0, 0.0, -1.0])
ctx.newton_rhs_from_r(True)
ctx.sync()
t2 = time.perf_counter()
a = np.zeros(NV)
dt = 1e-3
fac0 = 0.25 * dt * dt * 4.0 / 9.0               # beta alpha_f / alpha_m dt^2 for rho_inf = 0.5
out = {"n": n, "NV": NV, "nnz": nnz, "n_ess": int(ess.size), "fac0": fac0, "mass_setup_ms": (t1 - t0) * 1e3,
       "body_force_setup_ms": (t2 - t1) * 1e3}
for rep in range(2):                             # second pass = warm
    t0 = time.perf_counter(); ctx.newton_residual_and_grad_device(x, None, a, fac0, 0.0); ctx.sync(); t1 = time.perf_counter()
    it, fn, conv = ctx.newton_solve_gmres_device(); t2 = time.perf_counter()
    c = ctx.newton_get_solution(); t3 = time.perf_counter()
    nrm = ctx.newton_residual_device(x, None, a - c, fac0, 0.0); t4 = time.perf_counter()
    out.update(sweeps_ms=(t1 - t0) * 1e3, gmres_ms=(t2 - t1) * 1e3, gmres_iterations=it, gmres_converged=conv,
               gmres_final_norm=fn, line_search_residual_ms=(t4 - t3) * 1e3, residual_norm_after_full_step=nrm)
r0 = float(np.linalg.norm(ctx.get_r()))
out["note"] = "residual norm after one full Newton step printed above; timings of the second (warm) pass"
print(json.dumps(out))
