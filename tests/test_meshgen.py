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


def test_top_face_tables_match_the_reference_tables():
    """meshgen.top_face_tables (synthetic contact workloads of bench.py) against the boundary tables the unmodified
    reference precomputed for the same 4^3 p=2 cube (tests/golden/contact_cube_p2_n4_sphere.npz): same faces in the same
    order, same node order, values to rounding."""
    import numpy as np
    from mimi_b200 import meshgen
    from tests.cases import load_contact_golden
    D = load_contact_golden("contact_cube_p2_n4_sphere")
    M = meshgen.patch_connectivity(2, (4, 4, 4))
    T = meshgen.top_face_tables(M, (4, 4, 4))
    assert np.array_equal(M["v_dofs"], D["v_dofs"]) and np.array_equal(T["v_dofs"], D["c_v_dofs"])
    assert np.abs(T["N"] - D["c_N"]).max() < 1e-14 and np.abs(T["dN_dxi"] - D["c_dN_dxi"]).max() < 1e-14
    assert np.abs(T["w"] - D["c_w"]).max() < 1e-15


def test_top_face_tables_from_grid_equals_the_whole_mesh_version():
    import numpy as np
    from mimi_b200 import meshgen
    spans = (3, 2, 4)
    M = meshgen.patch_connectivity(2, spans)
    A = meshgen.top_face_tables(M, spans)
    g = meshgen.patch_node_grid(M["ncp"])
    B = meshgen.top_face_tables_from_grid(2, spans[:2], M["knots"], g[:, :, -1])
    for k in ("v_dofs", "N", "dN_dxi", "w"):
        assert np.array_equal(A[k], B[k]), k
