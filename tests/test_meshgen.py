"""CPU: the synthetic single-patch connectivity reproduces MFEM's dof numbering (checked against the
reference-dumped fixtures and, where oracle/_ref is built, a fresh dump with three different span counts)."""
import os
import tempfile

import numpy as np
import pytest

from mimi_b200 import meshgen
from oracle import oracle_py as O
from tests.cases import load_golden


@pytest.mark.parametrize("name,p,spans,L", [("cube_p2_n2", 2, (2, 2, 2), (1.0, 1.0, 1.0)), ("cube_p1_n2", 1, (2, 2, 2), (1.0, 1.0, 1.0)),
                                            ("cube_p2_n3_patch", 2, (3, 2, 2), (1.5, 1.0, 1.0))])
def test_numbering_matches_fixture(name, p, spans, L):
    D = load_golden(name)
    M = meshgen.patch_connectivity(p, spans, L)
    assert np.array_equal(M["v_dofs"], D["v_dofs"]) and np.array_equal(M["x_ref"], D["x_ref"])
    assert np.array_equal(meshgen.synth_u(D["NV"]), D["u"])


@pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("p,spans", [(2, (3, 4, 5)), (3, (2, 3, 2)), (1, (4, 2, 3))])
def test_numbering_matches_live_reference(p, spans):
    tmp = tempfile.mkdtemp(prefix="mb2mg_")
    mesh = os.path.join(tmp, "m.mesh")
    meshgen.write_box_patch(mesh, p, spans, (1.0, 2.0, 0.5))
    D = O.run_ref_dump(mesh, 0, 0, "stvk", os.path.join(tmp, "d"), lite=True)
    M = meshgen.patch_connectivity(p, spans, (1.0, 2.0, 0.5))
    assert np.array_equal(M["v_dofs"], D["v_dofs"]) and np.array_equal(M["x_ref"], D["x_ref"])


def test_node_grid_is_a_permutation():
    g = meshgen.patch_node_grid((7, 5, 9))
    assert sorted(g.ravel().tolist()) == list(range(7 * 5 * 9))
