"""Geometric factors generated on the device (d4b_geom_compute) against the
arrays the unmodified reference computed (d4est_mesh_data_compute,
d4est_mesh_compute_mortar_quadrature_quantities) and against the CPU restatement
tests/geom_ref.py.  Tolerance: 1e-12 relative to the largest entry of each array.

The *.geom.d4b.gz sidecars (ref_driver --geom) hold every element's position in the forest,
the parameters of the map, and the reference's hm / hp under each face_h_type."""
import os

import numpy as np
import pytest

import d4bio
import geom_ref as R
from conftest import GOLDEN

TOL = 1e-12
CASES = ["brick_l2_p3", "sphere_p2", "sphere_hp_amr1"]
MORTAR = ["sj_m_mortar_quad", "n_m_mortar_quad", "drst_dxyz_m_mortar_quad", "drst_dxyz_p_mortar_quad_porder",
          "hm_mortar_quad", "hp_mortar_quad"]


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def sidecar(case):
    return d4bio.load(os.path.join(GOLDEN, case + ".geom.d4b.gz"))


def interior_masks(A):
    """True at the mortar points of interior sides: the reference never writes hp and
    drst_dxyz_p on boundary sides (d4est_mesh.c:481-640), so only those are compared"""
    MP = int(np.asarray(A["local_mortar_nodes_quad"]).reshape(-1)[0])
    stride = A["side_mortar_stride"].astype(np.int64)
    order = np.argsort(stride)
    ends = np.empty_like(stride)
    ends[order] = np.concatenate([stride[order][1:], [MP]])
    m1, m9 = np.zeros(MP, bool), np.zeros(9 * MP, bool)
    for s in np.nonzero(A["side_faces_p"] > 0)[0]:
        m1[stride[s]:ends[s]] = True
        m9[9 * stride[s]:9 * ends[s]] = True
    return m1, m9


def check_mortar(M, A, m1, m9):
    for k in MORTAR:
        mask = m9 if k == "drst_dxyz_p_mortar_quad_porder" else m1 if k == "hp_mortar_quad" else slice(None)
        assert rel(M[k][mask], A[k][mask]) < TOL, k


# ------------------------------------------------------------------ CPU: pin the restatement
@pytest.mark.parametrize("case", CASES)
def test_geom_ref_matches_reference(golden, case):
    A, G = golden(case), sidecar(case)
    g = R.geom_from_sidecar(G)
    J, rx = R.volume_factors(A, g)
    assert rel(J, A["J_quad"]) < TOL and rel(rx, A["rst_xyz_quad"]) < TOL
    assert rel(R.nodal_xyz(A, g), A["xyz"]) < TOL
    m1, m9 = interior_masks(A)
    check_mortar(R.mortar_factors(A, g), A, m1, m9)


@pytest.mark.parametrize("case", ["brick_l2_p3", "sphere_p2"])
def test_geom_ref_face_h_types(golden, case):
    A, G = golden(case), sidecar(case)
    g = R.geom_from_sidecar(G)
    for t in range(6):
        M = R.mortar_factors(A, g, face_h_type=t)
        assert rel(M["hm_mortar_quad"], G["hm_type%d" % t]) < TOL
        assert rel(M["hp_mortar_quad"], G["hp_type%d" % t]) < TOL


# ------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_device_geometry_matches_reference(golden, case):
    from disco4est_b200.runtime import geom_compute

    A, G = golden(case), sidecar(case)
    g = R.geom_from_sidecar(G)
    out = geom_compute(A, g)
    assert rel(out["J_quad"], A["J_quad"]) < TOL
    assert rel(out["rst_xyz_quad"], A["rst_xyz_quad"]) < TOL
    assert rel(out["xyz"], A["xyz"]) < TOL
    m1, m9 = interior_masks(A)
    check_mortar(out, A, m1, m9)
    assert np.all(out["hp_mortar_quad"][~m1] == 0.0)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_device_face_h_types(golden, case):
    from disco4est_b200.runtime import geom_compute

    A, G = golden(case), sidecar(case)
    for t in range(6):
        out = geom_compute(A, R.geom_from_sidecar(G, face_h_type=t), want=["hm_mortar_quad", "hp_mortar_quad"])
        assert rel(out["hm_mortar_quad"], G["hm_type%d" % t]) < TOL, t
        assert rel(out["hp_mortar_quad"], G["hp_type%d" % t]) < TOL, t


@pytest.mark.gpu
def test_device_geometry_compactified_shells(golden):
    """maps the fixtures do not cover (compactified radius laws, other radii): hand-derived
    device derivatives against complex-step differentiation of the restated maps"""
    from disco4est_b200.runtime import geom_compute

    A, G = golden("sphere_hp_amr1"), sidecar("sphere_hp_amr1")
    for ci, co, radii in [(1, 1, (1.0, 2.0, 3.0)), (0, 1, (0.7, 5.0, 1000.0)), (1, 0, (2.0, 3.0, 7.5))]:
        g = R.geom_from_sidecar(G)
        g.update(compactify_inner_shell=ci, compactify_outer_shell=co, R0=radii[0], R1=radii[1], R2=radii[2])
        out = geom_compute(A, g)
        J, rx = R.volume_factors(A, g)
        assert rel(out["J_quad"], J) < TOL and rel(out["rst_xyz_quad"], rx) < TOL
        assert rel(out["xyz"], R.nodal_xyz(A, g)) < TOL
        M = R.mortar_factors(A, g)
        for k in MORTAR:
            assert rel(out[k], M[k]) < TOL, (ci, co, k)


@pytest.mark.gpu
def test_device_outputs_device_pointers(golden):
    import torch

    from disco4est_b200.runtime import geom_compute

    A, G = golden("sphere_p2"), sidecar("sphere_p2")
    g = R.geom_from_sidecar(G)
    dev = torch.device("cuda:0")
    out = {"J_quad": torch.zeros(int(np.asarray(A["local_nodes_quad"]).reshape(-1)[0]), dtype=torch.float64, device=dev),
           "sj_m_mortar_quad": torch.zeros(int(np.asarray(A["local_mortar_nodes_quad"]).reshape(-1)[0]), dtype=torch.float64, device=dev)}
    geom_compute(A, g, out=out)
    assert rel(out["J_quad"].cpu().numpy(), A["J_quad"]) < TOL
    assert rel(out["sj_m_mortar_quad"].cpu().numpy(), A["sj_m_mortar_quad"]) < TOL


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_mesh_with_device_geometry_applies_like_reference(golden, case):
    """d4b_mesh_create_with_geometry: the operator built from generated factors == the reference's Au"""
    from disco4est_b200 import Mesh

    A, G = golden(case), sidecar(case)
    m = Mesh(A, geom=R.geom_from_sidecar(G))
    u = np.ascontiguousarray(A["u"])
    Au = np.zeros_like(u)
    m.apply(u, Au)
    m.close()
    assert np.linalg.norm(Au - A["Au_plain"]) / np.linalg.norm(A["Au_plain"]) < TOL
