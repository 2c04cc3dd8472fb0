"""The synthetic unit-cube generator (disco4est_b200/brick.py) must describe
the same mesh the reference builds: checked through the CPU oracle against the
reference's own outputs on the fixtures."""
import numpy as np
import pytest

import oracle_lib
from disco4est_b200.brick import brick_desc, brick_xyz


@pytest.mark.parametrize("case,p", [("brick_l2_p1", 1), ("brick_l2_p3", 3), ("brick_l2_p5", 5)])
def test_brick_generator_matches_reference_mesh(golden, case, p):
    A = golden(case)
    B = brick_desc(2, p)
    assert B["n_local"] == A["n_local"][0] and B["local_nodes"] == A["local_nodes"][0]
    assert np.allclose(brick_xyz(2, p).reshape(-1), A["xyz"], atol=1e-14)
    assert np.array_equal(B["nodal_stride"], A["nodal_stride"])
    assert np.allclose(B["J_quad"], A["J_quad"], rtol=1e-14)
    assert np.allclose(B["rst_xyz_quad"], A["rst_xyz_quad"], rtol=1e-14, atol=1e-14)
    Au = oracle_lib.Oracle(B).apply(A["u"])
    assert np.linalg.norm(Au - A["Au_plain"]) / np.linalg.norm(A["Au_plain"]) < 1e-12
