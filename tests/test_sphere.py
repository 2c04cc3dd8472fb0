"""Synthetic cubed-sphere description (disco4est_b200/sphere.py): topology against the
reference's own sphere fixture, the operator built from it (geometric factors generated on the
device) against the reference's Au and the CPU oracle, size-independent properties at bench size."""
import numpy as np
import pytest

import d4bio
import geom_ref as R
from test_geometry import sidecar

TOL = 1e-12


def _records(D):
    S = int(np.asarray(D["n_sides"]).reshape(-1)[0])
    r = {}
    for s in range(S):
        fp = int(D["side_faces_p"][s])
        r[(int(D["side_e_m"][4 * s]), int(D["side_f_m"][s]))] = (
            fp, int(D["side_e_p"][4 * s]) if fp else -1, int(D["side_f_p"][s]) if fp else -1,
            int(D["side_perm"][s]) if fp else 0)
    return r


def test_topology_matches_reference_fixture(golden):
    """same elements in the same order, same neighbour / face / permutation on every side as the
    reference's p4est_iterate produced for the level-1 sphere"""
    from disco4est_b200.sphere import sphere_desc

    A, G = golden("sphere_p2"), sidecar("sphere_p2")
    B, g = sphere_desc(1, p=2)
    assert np.array_equal(G["elem_tree"], g["elem_tree"])
    assert np.array_equal(G["elem_q"], g["elem_q"]) and np.array_equal(G["elem_dq"], g["elem_dq"])
    assert np.array_equal(A["deg"], B["deg"]) and np.array_equal(A["nodal_stride"], B["nodal_stride"])
    assert _records(A) == _records(B)


def test_topology_is_consistent_at_depth():
    """every interior side has a reverse side that points back with the inverse permutation"""
    from disco4est_b200.sphere import _perm_index, sphere_desc

    B, _ = sphere_desc(3, p=1)
    S = B["n_sides"]
    e_m, e_p = B["side_e_m"].reshape(S, 4)[:, 0], B["side_e_p"].reshape(S, 4)[:, 0]
    f_m, f_p, perm = B["side_f_m"], B["side_f_p"], B["side_perm"]
    interior = B["side_faces_p"] == 1
    back = 6 * e_p[interior] + f_p[interior]
    assert np.array_equal(e_p[back], e_m[interior]) and np.array_equal(f_p[back], f_m[interior])
    n = 3
    for s in np.nonzero(interior)[0][::37]:
        r = 6 * e_p[s] + f_p[s]
        for k in range(n * n):
            kp = _perm_index(n, int(perm[s]), k % n, k // n)
            assert _perm_index(n, int(perm[r]), kp % n, kp // n) == k
    # boundary = outer face (f = 5) of the six outer-shell trees only
    nb = np.nonzero(~interior)[0]
    assert np.all(f_m[nb] == 5) and np.all(e_m[nb] < 6 * 8 ** 3) and nb.size == 6 * 8 * 8


def _full_arrays(B, g):
    """topology + geometric factors from the CPU restatement (independent of the device path)"""
    gg = dict(g)
    gg["geom_type"], gg["face_h_type"] = 1, {"j_div_sj_quad": 0, "tree_h": 4, "volume_div_area": 5}[g["face_h_type"]]
    A = dict(B)
    A["J_quad"], A["rst_xyz_quad"] = R.volume_factors(B, gg)
    A.update(R.mortar_factors(B, gg))
    return A


@pytest.mark.gpu
def test_operator_matches_reference_fixture(golden):
    from disco4est_b200 import Mesh
    from disco4est_b200.sphere import sphere_desc

    A = golden("sphere_p2")
    B, g = sphere_desc(1, p=2, prefactor=1.1)
    m = Mesh(B, geom=g)
    u = np.ascontiguousarray(A["u"])
    Au = np.zeros_like(u)
    m.apply(u, Au)
    m.close()
    assert np.linalg.norm(Au - A["Au_plain"]) / np.linalg.norm(A["Au_plain"]) < TOL


@pytest.mark.gpu
@pytest.mark.parametrize("kw", [dict(p=3), dict(pmax=4), dict(pmax=3, compactify=(1, 1), face_h_type="tree_h")])
def test_operator_matches_oracle(kw):
    import oracle_lib
    from disco4est_b200 import Mesh
    from disco4est_b200.sphere import sphere_desc

    B, g = sphere_desc(1, prefactor=2.0, **kw)
    m = Mesh(B, geom=g)
    u = np.random.default_rng(11).standard_normal(m.n)
    Au = np.zeros_like(u)
    m.apply(u, Au)
    m.close()
    ref = oracle_lib.Oracle(_full_arrays(B, g)).apply(u)
    assert np.linalg.norm(Au - ref) / np.linalg.norm(ref) < TOL


@pytest.mark.gpu
@pytest.mark.parametrize("kw", [dict(p=3), dict(pmax=4), dict(pmax=2)])
def test_properties_at_depth(kw):
    """level 3 (6656 elements): symmetry of the operator, linearity, the kernels agree"""
    import os

    from disco4est_b200 import Mesh
    from disco4est_b200.sphere import sphere_desc

    B, g = sphere_desc(3, **kw)
    m = Mesh(B, geom=g)
    rng = np.random.default_rng(5)
    u, v = rng.standard_normal(m.n), rng.standard_normal(m.n)
    Au, Av, Auv = np.zeros(m.n), np.zeros(m.n), np.zeros(m.n)
    m.apply(u, Au); m.apply(v, Av); m.apply(u + 2.0 * v, Auv)
    assert abs(v @ Au - u @ Av) < 1e-11 * abs(v @ Au)
    assert np.linalg.norm(Auv - Au - 2.0 * Av) < 1e-12 * np.linalg.norm(Auv)
    assert u @ Au > 0
    m.close()
    os.environ["D4B_FORCE_GENERIC"] = "1"
    try:
        mg = Mesh(B, geom=g)
        Ag = np.zeros(m.n)
        mg.apply(u, Ag)
        mg.close()
    finally:
        del os.environ["D4B_FORCE_GENERIC"]
    assert np.linalg.norm(Ag - Au) < 1e-12 * np.linalg.norm(Au)


@pytest.mark.gpu
def test_random_hp_sphere_runs_in_the_hp_kernel():
    from disco4est_b200 import Mesh
    from disco4est_b200.sphere import sphere_desc

    B, g = sphere_desc(2, pmax=5)
    m = Mesh(B, geom=g)
    assert m.fast_elements() == B["n_local"] and m.hp_elements() > 0.9 * B["n_local"]
    m.close()
