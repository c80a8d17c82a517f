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
