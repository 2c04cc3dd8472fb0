"""Synthetic AMR brick (disco4est_b200.brick.brick_amr_desc: unit cube with one octant refined once, hanging
faces on three interface planes -- what one AMR step of the ConstantDensityStar driver produces, BASELINE.json
configs[0]).  CPU: the description through the pinned oracle (every element face covered once, symmetry, the
reference's polynomial known answer across hanging faces).  GPU: the device path against that oracle, the
face split against the element-wise mortar kernel, properties at bench size."""
import numpy as np
import pytest

import geom_ref as R


def _full(A, g):
    gg = dict(g)
    gg.update(geom_type=0, face_h_type={"volume_div_area": 5, "tree_h": 4}[g["face_h_type"]],
              compactify_inner_shell=0, compactify_outer_shell=0)
    F = dict(A)
    F["J_quad"], F["rst_xyz_quad"] = R.volume_factors(A, gg)
    F.update(R.mortar_factors(A, gg))
    return F, gg


def test_description_through_the_oracle():
    import oracle_lib
    from disco4est_b200.brick import brick_amr_desc

    A, g = brick_amr_desc(2, 2)
    S = A["n_sides"]
    em = A["side_e_m"].reshape(S, 4)
    cnt = np.zeros((A["n_local"], 6), int)
    for s in range(S):
        for i in range(A["side_faces_m"][s]):
            cnt[em[s, i], A["side_f_m"][s]] += 1
    assert (cnt == 1).all()
    assert (A["side_faces_p"] == 4).sum() == (A["side_faces_m"] == 4).sum() == 3 * 2 * 2
    F, gg = _full(A, g)
    O = oracle_lib.Oracle(F)
    rng = np.random.default_rng(0)
    u, v = rng.standard_normal(O.n), rng.standard_normal(O.n)
    Au, Av = O.apply(u), O.apply(v)
    assert abs(v @ Au - u @ Av) < 1e-13 * abs(v @ Au) and u @ Au > 0
    # A(x^2+y^2+z^2) = M(-6) away from the boundary: a continuous polynomial has no jumps, hanging faces included
    xyz = R.nodal_xyz(A, gg).reshape(3, -1)
    Ap = O.apply((xyz ** 2).sum(0))
    M1 = O.apply_mass(np.ones(O.n), np.ascontiguousarray(F["J_quad"]))
    lo, hi = xyz.reshape(3, -1, 27).min(2), xyz.reshape(3, -1, 27).max(2)
    inter = ((lo > 1e-12) & (hi < 1 - 1e-12)).all(0)
    assert inter.sum() > 20
    assert np.abs((Ap + 6 * M1).reshape(-1, 27)[inter]).max() < 1e-13


@pytest.mark.gpu
@pytest.mark.parametrize("level,p,fh", [(2, 2, "volume_div_area"), (2, 3, "tree_h"), (3, 1, "volume_div_area")])
def test_device_matches_oracle(level, p, fh, monkeypatch):
    import oracle_lib
    from disco4est_b200 import Mesh
    from disco4est_b200.brick import brick_amr_desc

    A, g = brick_amr_desc(level, p, face_h_type=fh)
    F, _ = _full(A, g)
    u = np.random.default_rng(4).standard_normal(int(A["local_nodes"]))
    ref = oracle_lib.Oracle(F).apply(u)
    for split in ("1", "0"):
        monkeypatch.setenv("D4B_SPLIT", split)
        m = Mesh(A, geom=g)
        assert m.hp_elements() > 0 and m.fast_elements() == A["n_local"]
        Au = np.zeros_like(u)
        m.apply(u, Au)
        m.close()
        assert np.linalg.norm(Au - ref) / np.linalg.norm(ref) < 1e-12, split
    monkeypatch.delenv("D4B_SPLIT")


@pytest.mark.gpu
def test_properties_at_size():
    """32^3 base + refined octant (57 344 elements, p=3): polynomial known answer, symmetry, split == mortar kernel"""
    import os

    from disco4est_b200 import Mesh
    from disco4est_b200.brick import brick_amr_desc
    from disco4est_b200.runtime import geom_compute

    A, g = brick_amr_desc(5, 3)
    m = Mesh(A, geom=g)
    out = geom_compute(A, g, want=["xyz", "J_quad"])
    xyz = out["xyz"].reshape(3, -1)
    up = np.ascontiguousarray((xyz ** 2).sum(0))
    Ap, M1 = np.zeros(m.n), np.zeros(m.n)
    m.apply(up, Ap)
    m.apply_mass(np.ones(m.n), out["J_quad"], M1)
    lo, hi = xyz.reshape(3, -1, 64).min(2), xyz.reshape(3, -1, 64).max(2)
    inter = ((lo > 1e-12) & (hi < 1 - 1e-12)).all(0)
    assert np.abs((Ap + 6 * M1).reshape(-1, 64)[inter]).max() < 1e-10 * np.abs(M1).max() * 6
    rng = np.random.default_rng(6)
    u, v = rng.standard_normal(m.n), rng.standard_normal(m.n)
    Au, Av = np.zeros(m.n), np.zeros(m.n)
    m.apply(u, Au); m.apply(v, Av)
    assert abs(v @ Au - u @ Av) < 1e-11 * abs(v @ Au)
    m.close()
    os.environ["D4B_SPLIT"] = "0"
    try:
        m2 = Mesh(A, geom=g)
        A2 = np.zeros(m.n)
        m2.apply(u, A2)
        m2.close()
    finally:
        del os.environ["D4B_SPLIT"]
    assert np.linalg.norm(A2 - Au) < 1e-13 * np.linalg.norm(Au)
