"""CPU: the Newton-iteration sweeps restated on the oracle's pieces against dumps of the
unmodified reference: mfem::add, SparseMatrix::Mult/AddMult, the domain integrator, EliminateRowCol, SparseMatrix::Add,
Vector::Add in the order of operators/nonlinear_solid.cpp:240-283. Bit-exact."""
import numpy as np
import pytest

from oracle import oracle_py as O
from tests.cases import SWEEP_CASES, load_sweeps_golden


@pytest.mark.parametrize("name", SWEEP_CASES)
def test_sweeps_bit_exact(name):
    D, W = load_sweeps_golden(name)
    assert np.array_equal(W["I"], D["I"]) and np.array_equal(W["J"], D["J"]) and np.array_equal(W["mass"], D["mass"])
    y, Jm, tx = O.newton_residual_and_grad(D["dim"], D["nd"], D["nq"], O.mat(0), D, W)
    assert np.array_equal(tx, W["tx"])
    assert np.array_equal(y, W["y"]) and np.array_equal(Jm, W["Jac"])
    # as in the reference, the viscosity is added AFTER the elimination: eliminated rows hold e_rc + fac1 S(rc, :)
    I, J = W["I"], W["J"]
    rc = int(W["ess"][0])
    row = Jm[I[rc]:I[rc + 1]]
    assert np.array_equal(row, np.where(J[I[rc]:I[rc + 1]] == rc, 1.0, 0.0) + W["fac1"] * W["visc"][I[rc]:I[rc + 1]])


@pytest.mark.parametrize("name", SWEEP_CASES)
def test_gmres_jacobi_bit_exact(name):
    """the reference's iterative linear solver (GMRESSolver + DSmoother, py_nonlinear_solid.cpp:330-339)
    on the Newton system J c = y of the sweeps fixture; the C restatement reproduces the compiled reference's solution,
    iteration count and final norm to the bit."""
    D, W = load_sweeps_golden(name)
    x, it, fn, conv = O.gmres_jacobi(W["I"], W["J"], W["Jac"], W["y"])
    assert conv and W["gmres_info"][2] == 1.0
    assert it == int(W["gmres_info"][0]) and fn == W["gmres_info"][1]
    assert np.array_equal(x, W["gmres_x"])
    # and it solves the system: preconditioned residual below the reference's tolerance
    r = O.spmv(W["I"], W["J"], W["Jac"], x) - W["y"]
    assert np.linalg.norm(r) <= 1e-6 * np.linalg.norm(W["y"])
    print("\n[gmres %s] %d iterations, final norm %.3e" % (name, it, fn))
