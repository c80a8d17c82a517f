"""Host-side replay of the reference's own end-to-end test (tests/test_nonlinear_solid.py:54-112 in the reference):
the 2D p=3 beam `balken.mesh`, clamped at x = 0, body force, 10 generalized-alpha steps, compared with
tests/data/ref/*_h1_p2/x_<i>.txt (fixture tests/golden/balken_histories.npz).

What is restated here, in numpy, is the part of the reference that is NOT on the hot path and that this repository
does not ship: the time integrator, the Newton loop with its three-point line search and the (50 x 50, dense) linear
solve. Residual, tangent and history commits come from a backend: the CPU oracle (tests/test_reference_histories.py,
CPU) or the CUDA library through its C ABI (same file, `-m gpu`).

  GeneralizedAlpha2::StepTime2          src/mimi/solvers/ode.cpp:16-79 (+ mfem GeneralizedAlpha2Solver::SetRhoInf)
  NonlinearSolid::{Mult, ImplicitSolve, Mult(a, y), ResidualAndGrad, PostTimeAdvance}
                                        src/mimi/operators/nonlinear_solid.cpp:124-294
  forms::Nonlinear::{AddMult, AddMultGrad} (essential dofs)   src/mimi/forms/nonlinear.hpp:53-116
  LineSearchNewton::Mult                src/mimi/solvers/newton.cpp:10-218
  PyNonlinearSolid::Setup (mass with eliminated essential dofs, body-force linear form)
                                        src/mimi/py/py_nonlinear_solid.cpp:155-283
"""
import numpy as np


def csr_to_dense(I, J, A, n):
    M = np.zeros((n, n))
    for i in range(n):
        M[i, J[I[i]:I[i + 1]]] = A[I[i]:I[i + 1]]
    return M


class Beam:
    def __init__(self, D, backend, body_force, dt, rho_inf=0.5, rel_tol=1e-12, abs_tol=1e-8, max_iter=10):
        self.D, self.be, self.dt = D, backend, dt
        self.NV, dim, nd, nq = D["NV"], D["dim"], D["nd"], D["nq"]
        self.I, self.J = D["I"], D["J"]
        # essential dofs: boundary id 2 = attribute 3 = the edge x = 0 (balken.mesh), both components
        xr = D["x_ref"].reshape(-1, dim)
        nodes = np.nonzero(xr[:, 0] == 0.0)[0]
        self.ess = np.sort(np.concatenate([dim * nodes + c for c in range(dim)])).astype(np.int32)
        # mass: VectorMassIntegrator(density = 1) as dumped from the reference, then FormSystemMatrix(zero_dofs):
        # essential rows / columns zeroed, diagonal one
        self.mass = D["mass"].copy()
        rows = np.repeat(np.arange(self.NV), np.diff(self.I))
        is_ess = np.zeros(self.NV, bool)
        is_ess[self.ess] = True
        self.mass[is_ess[rows] | is_ess[self.J]] = 0.0
        self.mass[(rows == self.J) & is_ess[rows]] = 1.0
        self.Md = csr_to_dense(self.I, self.J, self.mass, self.NV)
        # rhs: VectorDomainLFIntegrator of the constant body force, essential entries zeroed
        N = D["N"].reshape(D["NE"], nq, nd)
        wd = (D["w"] * D["detJ"]).reshape(D["NE"], nq)
        vd = D["v_dofs"].reshape(D["NE"], dim, nd)
        self.rhs = np.zeros(self.NV)
        for e in range(D["NE"]):
            fe = np.einsum("q,qa->a", wd[e], N[e])
            for c, val in body_force.items():
                np.add.at(self.rhs, vd[e, c], fe * val)
        self.rhs[self.ess] = 0.0
        # mfem::GeneralizedAlpha2Solver::SetRhoInf + GeneralizedAlpha2::ComputeFactors (ode.cpp:5-14)
        am = (2.0 - rho_inf) / (1.0 + rho_inf)
        af = 1.0 / (1.0 + rho_inf)
        beta = 0.25 * (1.0 + am - af) ** 2
        gamma = 0.5 + am - af
        self.f = (0.5 - beta / am, af, af * (1.0 - gamma / am), beta * af / am, gamma * af / am, am)
        self.rel_tol, self.abs_tol, self.max_iter = rel_tol, abs_tol, max_iter
        self.x = np.zeros(self.NV)
        self.v = np.zeros(self.NV)
        self.a = None
        self.newton_iterations = []

    # NonlinearSolid::Mult(a, y) (operators/nonlinear_solid.cpp:172-205)
    def _residual(self, xa, fac0, a):
        y = self.Md @ a + self.be.residual(np.ascontiguousarray(xa + fac0 * a))
        y[self.ess] = 0.0
        y -= self.rhs
        y[self.ess] = 0.0
        return y

    # NonlinearSolid::ResidualAndGrad (operators/nonlinear_solid.cpp:240-283)
    def _residual_and_grad(self, xa, fac0, a):
        r, A = self.be.residual_and_grad(np.ascontiguousarray(xa + fac0 * a), fac0, self.mass)
        y = self.Md @ a + r
        Jd = csr_to_dense(self.I, self.J, A, self.NV)
        Jd[self.ess, :] = 0.0          # SparseMatrix::EliminateRowCol, DIAG_ONE
        Jd[:, self.ess] = 0.0
        Jd[self.ess, self.ess] = 1.0
        y[self.ess] = 0.0
        y -= self.rhs
        y[self.ess] = 0.0
        return y, Jd

    def _solve(self, Jd, r):      # the linear solver of the Newton step (UMFPACK in the reference's test)
        return np.linalg.solve(Jd, r)

    # LineSearchNewton::Mult with iterative_mode == false (solvers/newton.cpp:10-218)
    def _newton(self, xa, fac0):
        a = np.zeros(self.NV)
        r, Jd = self._residual_and_grad(xa, fac0, a)
        norm = np.sqrt(r @ r)
        goal = max(self.rel_tol * norm, self.abs_tol)
        improved = [True] * 5
        n_imp, best, best_a, it = 0, np.finfo(float).max, a.copy(), 0
        while True:
            if norm <= goal:
                break
            if it >= self.max_iter:
                if it != 0:
                    a = best_a
                break
            if not any(improved):
                a = best_a
                break
            c = self._solve(Jd, r)
            q1 = norm
            r3 = self._residual(xa, fac0, a - c)
            q3 = np.sqrt(r3 @ r3)
            r2 = self._residual(xa, fac0, a - 0.5 * c)
            q2 = np.sqrt(r2 @ r2)
            den = q1 - 2.0 * q2 + q3
            eps = (3.0 * q1 - 4.0 * q2 + q3) / (4.0 * den) if den != 0.0 else np.inf
            if den > 0 and 0 < eps < 1:
                scale = eps
            elif q3 < q1:
                scale = 1.0
            else:
                scale = 0.05
            if abs(scale) < 1e-12:
                break
            a = a - scale * c
            if it == self.max_iter - 1:
                r = self._residual(xa, fac0, a)
            else:
                r, Jd = self._residual_and_grad(xa, fac0, a)
            norm = np.sqrt(r @ r)
            if norm < best:
                best_a, best = a.copy(), norm
                improved[n_imp % 5] = True
            else:
                improved[n_imp % 5] = False
            n_imp += 1
            it += 1
        self.newton_iterations.append(it)
        return a

    # GeneralizedAlpha2::StepTime2 (solvers/ode.cpp:16-79)
    def step(self):
        f0, f1, f2, f3, f4, f5 = self.f
        dt = self.dt
        self.be.set_time_step(dt)
        if self.a is None:  # NonlinearSolid::Mult(x, v, a): a = M^-1 (rhs - K(x))   (mass_inv_: CG in the reference)
            z = -self.be.residual(self.x) + self.rhs
            z[self.ess] = 0.0
            self.a = np.linalg.solve(self.Md, z)
        xa = self.x + (self.v + f0 * dt * self.a) * (f1 * dt)
        va = self.v + f2 * dt * self.a
        aa = self._newton(xa, f3 * dt * dt)
        xa = xa + f3 * dt * dt * aa
        va = va + f4 * dt * aa
        prev = 1.0 - 1.0 / f1
        self.x = self.x * prev + xa / f1
        self.v = self.v * prev + va / f1
        self.a = self.a * prev + aa / f5
        self.be.post_time_advance(np.ascontiguousarray(self.x))
        return self.x


class DeviceNewtonBeam(Beam):
    """The same time loop with every O(N) / O(nnz) operation of the Newton iteration on the GPU (SURVEY 8(f) ranks 1, 2
    and the line-search part of 4): mb2_newton_residual_and_grad_device (operator composition + elimination),
    mb2_newton_solve_gmres_device (the reference's GMRES + Jacobi option instead of UMFPACK) and
    mb2_newton_residual_device (line-search residuals). Only the 50-entry iterate travels."""

    def __init__(self, D, backend, ctx, *args, **kw):
        super().__init__(D, backend, *args, **kw)
        self.ctx = ctx
        ctx.set_mass(self.mass)
        ctx.set_rhs(self.rhs)
        ctx.set_essential_dofs(self.ess)
        self.gmres_iterations = []

    def _residual(self, xa, fac0, a):
        self.ctx.newton_residual_device(np.ascontiguousarray(xa), None, np.ascontiguousarray(a), fac0, 0.0, want_norm=False)
        return self.ctx.get_r()

    def _residual_and_grad(self, xa, fac0, a):
        self.ctx.newton_residual_and_grad_device(np.ascontiguousarray(xa), None, np.ascontiguousarray(a), fac0, 0.0)
        return self.ctx.get_r(), None

    def _solve(self, Jd, r):      # J and y are resident: nothing is uploaded
        it, fn, conv = self.ctx.newton_solve_gmres_device()
        assert conv
        self.gmres_iterations.append(it)
        return self.ctx.newton_get_solution()
