"""The reference's OWN golden vectors for this path: tests/data/ref/{neohook,j2,j2_simo,j2_log}_h1_p2/
x_<i>.txt, checked by the reference with np.allclose after each of 10 generalized-alpha steps of the clamped 2D p=3 beam
(tests/test_nonlinear_solid.py:54-112 there). They pin the composition assembly + Newton + linear solve + time
integration; tests/beam_driver.py restates the non-hot-path part in numpy and takes residual / tangent / history from

  * the CPU oracle        -> pins the oracle against vectors the reference's authors committed (CPU test), and
  * the CUDA library through its C ABI (`-m gpu`) -> the same assertions the reference's test makes, on the B200 path.

The assertions are the reference's: np.allclose(u, ref) (rtol 1e-5, atol 1e-8) for every step.
"""
import os

import numpy as np
import pytest

from oracle import oracle_py as O
from tests.beam_driver import Beam
from tests.cases import GOLDEN, j2_params, load_golden

CASES = ["neohook", "j2", "j2_simo", "j2_log"]
# balken_plasticity (tests/test_nonlinear_solid.py:16-33): JohnsonCookTemperatureAndRateDependentHardening, A 70, B 140,
# n 0.2835, m 1.3558, eps0_dot 0.004, reference temperature 20; C is left value-initialised (0); melting 1500
JC_HP = np.array([70.0, 140.0, 0.2835, 0.0, 0.004, 20.0, 1500.0, 1.3558, 0.0, 0.0])
SETUP = {  # body force on component 1, time step
    "neohook": (-5.0, 0.05),
    "j2": (-3.0, 0.5),
    "j2_simo": (-3.0, 0.5),
    "j2_log": (-3.0, 0.5),
}


def histories():
    z = np.load(os.path.join(GOLDEN, "balken_histories.npz"))
    return {k: z[k] for k in z.files}


class OracleBackend:
    def __init__(self, D, case):
        self.D = D
        NE, nq, dim = D["NE"], D["nq"], D["dim"]
        if case == "neohook":
            self.S, self.m = None, O.mat(0)
        else:
            self.S = {"j2": lambda: O.j2_initial_state(NE, nq), "j2_simo": lambda: O.j2simo_initial_state(NE, nq, dim),
                      "j2_log": lambda: O.j2log_initial_state(NE, nq, dim)}[case]()
            self.m = O.mat_j2(self.S, "jcthermo", simo=case == "j2_simo", log=case == "j2_log", hp=JC_HP)
        self.args = (D["dim"], D["nd"], D["nq"], self.m, D["v_dofs"])

    def set_time_step(self, dt):
        self.m.dt = dt

    def residual(self, x):
        D = self.D
        return O.assemble_residual(*self.args, D["dN_dX"], D["w"], D["detJ"], x)

    def residual_and_grad(self, x, gf, A0):
        D = self.D
        return O.assemble_residual_and_grad(*self.args, D["A_ids"], D["dN_dX"], D["w"], D["detJ"], x, gf, A0)

    def post_time_advance(self, x):
        if self.S is not None:
            D = self.D
            O.post_time_advance(D["dim"], D["nd"], D["nq"], self.m, D["v_dofs"], D["dN_dX"], x)


class DeviceBackend:
    """mb2_add_domain_residual / _and_grad / mb2_post_time_advance / mb2_set_time_step with host buffers."""

    def __init__(self, D, case, ctx):
        self.D, self.ctx = D, ctx
        ctx.set_connectivity(D["dim"], D["nd"], D["nq"], D["NV"], D["v_dofs"])
        ctx.upload_quad_tables(D["dN_dX"], D["w"], D["detJ"])
        ctx.build_pattern()
        if case == "neohook":
            ctx.set_material(0, list(O.lame()))
        else:
            ctx.set_material({"j2": 3, "j2_simo": 4, "j2_log": 5}[case], j2_params("jcthermo", hp=JC_HP))

    def set_time_step(self, dt):
        self.ctx.set_time_step(dt)

    def residual(self, x):
        r = np.zeros(self.D["NV"])
        self.ctx.add_domain_residual(x, r)
        return r

    def residual_and_grad(self, x, gf, A0):
        r = np.zeros(self.D["NV"])
        A = A0.copy()
        self.ctx.add_domain_residual_and_grad(x, gf, r, A)
        return r, A

    def post_time_advance(self, x):
        self.ctx.post_time_advance(x)


def replay(case, backend_factory):
    D = load_golden("balken_p3_n2")
    force, dt = SETUP[case]
    beam = Beam(D, backend_factory(D, case), {1: force}, dt)
    ref = histories()[case]
    worst = 0.0
    for i in range(10):
        u = beam.step()
        worst = max(worst, float(np.max(np.abs(u - ref[i]) / (1e-8 + 1e-5 * np.abs(ref[i])))))
        assert np.allclose(u, ref[i]), (case, i)      # the reference's own assertion
    return beam, worst


@pytest.mark.parametrize("case", CASES)
def test_oracle_replays_reference_histories(case):
    beam, worst = replay(case, OracleBackend)
    print("\n[%s] oracle: worst |u - ref| / (atol + rtol |ref|) = %.2e, Newton iterations %s" % (case, worst, beam.newton_iterations))
    assert np.abs(beam.x).max() > 1e-3      # the beam actually bends
    if case != "neohook":
        assert np.any(beam.be.S[:, :, -2] > 0)      # plastic flow happened (accumulated plastic strain)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_device_replays_reference_histories(ctx_factory, case):
    beam, worst = replay(case, lambda D, c: DeviceBackend(D, c, ctx_factory()))
    print("\n[%s] device: worst |u - ref| / (atol + rtol |ref|) = %.2e, Newton iterations %s" % (case, worst, beam.newton_iterations))
    ob, _ = replay(case, OracleBackend)
    assert beam.newton_iterations == ob.newton_iterations      # identical Newton iteration counts (north_star)
    assert np.allclose(beam.x, ob.x, rtol=1e-9, atol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_device_newton_loop_replays_reference_histories(ctx_factory, case):
    """The whole Newton iteration on the device (tests/beam_driver.py: DeviceNewtonBeam): operator composition and
    elimination, GMRES(50) + Jacobi in place of UMFPACK, line-search residuals; ten generalized-alpha steps against the
    reference's committed vectors with the reference's own assertion."""
    from tests.beam_driver import DeviceNewtonBeam
    D = load_golden("balken_p3_n2")
    force, dt = SETUP[case]
    ctx = ctx_factory()
    beam = DeviceNewtonBeam(D, DeviceBackend(D, case, ctx), ctx, {1: force}, dt)
    ref = histories()[case]
    worst = 0.0
    for i in range(10):
        u = beam.step()
        worst = max(worst, float(np.max(np.abs(u - ref[i]) / (1e-8 + 1e-5 * np.abs(ref[i])))))
        assert np.allclose(u, ref[i]), (case, i)
    print("\n[%s] device Newton loop: worst |u - ref| / (atol + rtol |ref|) = %.2e, Newton iterations %s, GMRES iterations "
          "per solve %d..%d" % (case, worst, beam.newton_iterations, min(beam.gmres_iterations), max(beam.gmres_iterations)))
