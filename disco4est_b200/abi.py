"""ctypes mirror of include/d4est_b200.h (the C ABI of libd4est_b200.so).

Only the test / bench harness uses this: the product is the shared library and
the C adapter (disco4est_b200/adapter); Python is plumbing.
"""
import ctypes as C

import numpy as np

D4B_ABI_VERSION = 1

PENALTY = {
    "maxp_sqr_over_minh": 0,
    "maxpp1_sqr_over_minh": 1,
    "meanp_sqr_over_meanh": 2,
    "mean_p_sqr_over_h": 3,
}

_I32P = C.POINTER(C.c_int32)
_I64P = C.POINTER(C.c_int64)
_F64P = C.POINTER(C.c_double)


class MeshDesc(C.Structure):
    """d4b_mesh_desc_t, field for field (include/d4est_b200.h)."""

    _fields_ = [
        ("abi_version", C.c_int32),
        ("n_local", C.c_int32),
        ("n_ghost", C.c_int32),
        ("n_sides", C.c_int32),
        ("local_nodes", C.c_int64),
        ("local_nodes_quad", C.c_int64),
        ("local_mortar_nodes_quad", C.c_int64),
        ("ghost_nodes", C.c_int64),
        ("deg", _I32P),
        ("deg_quad", _I32P),
        ("nodal_stride", _I64P),
        ("quad_stride", _I64P),
        ("ghost_deg", _I32P),
        ("ghost_deg_quad", _I32P),
        ("ghost_nodal_stride", _I64P),
        ("J_quad", _F64P),
        ("rst_xyz_quad", _F64P),
        ("sj_m_mortar_quad", _F64P),
        ("n_m_mortar_quad", _F64P),
        ("drst_dxyz_m_mortar_quad", _F64P),
        ("drst_dxyz_p_mortar_quad_porder", _F64P),
        ("hm_mortar_quad", _F64P),
        ("hp_mortar_quad", _F64P),
        ("dirichlet_g_mortar_quad", _F64P),
        ("side_faces_m", _I32P),
        ("side_faces_p", _I32P),
        ("side_f_m", _I32P),
        ("side_f_p", _I32P),
        ("side_e_m", _I32P),
        ("side_e_p", _I32P),
        ("side_perm", _I32P),
        ("side_sub_order", _I32P),
        ("side_mortar_stride", _I64P),
        ("side_penalty_mult", _F64P),
        ("sipg_penalty_prefactor", C.c_double),
        ("sipg_penalty_fcn", C.c_int32),
        ("bc_type", C.c_int32),
        ("mpirank", C.c_int32),
        ("mpisize", C.c_int32),
        ("n_mirrors", C.c_int32),
        ("mirror_elem", _I32P),
        ("mirror_proc_offsets", _I32P),
        ("mirror_proc_mirrors", _I32P),
        ("ghost_proc_offsets", _I32P),
        ("robin_coeff_mortar_quad", _F64P),
        ("robin_rhs_mortar_quad", _F64P),
    ]


class GeomDesc(C.Structure):
    """d4b_geom_desc_t"""

    _fields_ = [
        ("geom_type", C.c_int32),
        ("face_h_type", C.c_int32),
        ("compactify_inner_shell", C.c_int32),
        ("compactify_outer_shell", C.c_int32),
        ("X0", C.c_double), ("X1", C.c_double), ("Y0", C.c_double), ("Y1", C.c_double),
        ("Z0", C.c_double), ("Z1", C.c_double),
        ("R0", C.c_double), ("R1", C.c_double), ("R2", C.c_double),
        ("elem_tree", _I32P), ("elem_q", _I32P), ("elem_dq", _I32P),
        ("ghost_tree", _I32P), ("ghost_q", _I32P), ("ghost_dq", _I32P),
    ]


GEOM_ARRAY_FIELDS = ["xyz", "J_quad", "rst_xyz_quad", "sj_m_mortar_quad", "n_m_mortar_quad",
                     "drst_dxyz_m_mortar_quad", "drst_dxyz_p_mortar_quad_porder", "hm_mortar_quad", "hp_mortar_quad"]


class GeomArrays(C.Structure):
    """d4b_geom_arrays_t (addresses; host or device)"""

    _fields_ = [(name, C.c_void_p) for name in GEOM_ARRAY_FIELDS + ["xyz_quad"]]


NL_KIND = {"punctures": 1, "constant_density_star": 2, "table": 3}
MAX_PUNCTURES = 10


class NonlinearDesc(C.Structure):
    """d4b_nonlinear_desc_t"""

    _fields_ = [("kind", C.c_int32), ("n_punctures", C.c_int32),
                ("punctures", (C.c_double * 10) * MAX_PUNCTURES),
                ("star_rho0", C.c_double), ("star_R", C.c_double), ("star_c", C.c_double * 3),
                ("exponent", C.c_int32),
                ("a_quad", C.c_void_p), ("b_quad", C.c_void_p), ("xyz_quad", C.c_void_p), ("jac_quad", C.c_void_p)]


class NewtonParams(C.Structure):
    """d4b_newton_params_t"""

    _fields_ = [("atol", C.c_double), ("rtol", C.c_double), ("imin", C.c_int32), ("imax", C.c_int32),
                ("fcg_atol", C.c_double), ("fcg_rtol", C.c_double), ("fcg_imax", C.c_int32), ("fcg_flags", C.c_int32)]


class NewtonStats(C.Structure):
    """d4b_newton_stats_t"""

    _fields_ = [("steps", C.c_int32), ("fcg_iterations", C.c_int32 * 64), ("residual_norm", C.c_double * 65),
                ("jacobian_applies", C.c_int64)]


GEOM_TYPE = {"brick": 0, "cubed_sphere": 1}
FACE_H = {"j_div_sj_quad": 0, "j_div_sj_min_lobatto": 1, "j_div_sj_mean_lobatto": 2, "j_div_sj_max_lobatto": 3,
          "tree_h": 4, "volume_div_area": 5}


def make_geom_desc(g):
    """GeomDesc from a dict: geom_type / face_h_type (names or numbers), the map parameters and the
    element positions elem_tree, elem_q, elem_dq (+ ghost_*).  Returns (desc, keepalive)."""
    d = GeomDesc()
    keep = []
    gt, fh = g["geom_type"], g["face_h_type"]
    d.geom_type = GEOM_TYPE[gt] if isinstance(gt, str) else int(np.asarray(gt).reshape(-1)[0])
    d.face_h_type = FACE_H[fh] if isinstance(fh, str) else int(np.asarray(fh).reshape(-1)[0])
    for name, typ in GeomDesc._fields_:
        if name in ("geom_type", "face_h_type") or name not in g or g[name] is None:
            continue
        if typ is _I32P:
            a = np.ascontiguousarray(g[name], dtype=np.int32).reshape(-1)
            if a.size == 0:
                a = np.zeros(1, np.int32)
            keep.append(a)
            setattr(d, name, a.ctypes.data_as(_I32P))
        else:
            v = np.asarray(g[name]).reshape(-1)[0]
            setattr(d, name, float(v) if typ is C.c_double else int(v))
    return d, keep


class EstimatorDesc(C.Structure):
    """d4b_estimator_desc_t (addresses; host or device)"""

    _fields_ = [(n, C.c_void_p) for n in ("u", "residual", "jac_quad", "diam_volume", "sj_m_mortar_quad",
                                          "hm_mortar_quad", "hp_mortar_quad", "dirichlet_g_mortar_quad")] + [
        ("penalty_family", C.c_int32), ("penalty_prefactor", C.c_double)]


EST_FAMILY = {"houston_maxp_minh": 0, "houston_max_p2_over_h": 1, "bi_maxp_minh": 2, "bi_max_p2_over_h": 3}


class ChebyParams(C.Structure):
    _fields_ = [
        ("cheby_imax", C.c_int32),
        ("cheby_eigs_cg_imax", C.c_int32),
        ("cheby_eigs_lmax_lmin_ratio", C.c_double),
        ("cheby_eigs_max_multiplier", C.c_double),
        ("cheby_use_new_cg_eigs", C.c_int32),
        ("cheby_use_zero_guess_for_eigs", C.c_int32),
        ("cheby_eigs_reuse_fromdownvcycle", C.c_int32),
        ("cheby_eigs_reuse_fromlastvcycle", C.c_int32),
    ]


class MgDesc(C.Structure):
    """d4b_mg_desc_t"""

    _fields_ = [
        ("n_levels", C.c_int32),
        ("meshes", C.POINTER(C.c_void_p)),
        ("child_first", C.POINTER(_I32P)),
        ("child_count", C.POINTER(_I32P)),
        ("vcycle_imax", C.c_int32),
        ("vcycle_rtol", C.c_double),
        ("vcycle_atol", C.c_double),
        ("smoother", ChebyParams),
        ("bottom_solver", C.c_int32),
        ("bottom_iter", C.c_int32),
        ("bottom_rtol", C.c_double),
        ("bottom_atol", C.c_double),
        ("bottom_cheby", ChebyParams),
    ]


_PTR_FIELDS = {name: typ for name, typ in MeshDesc._fields_ if typ in (_I32P, _I64P, _F64P)}
_NP_OF = {_I32P: np.int32, _I64P: np.int64, _F64P: np.float64}


def make_desc(arrays):
    """Build a MeshDesc from a dict of numpy arrays / scalars.

    Keys follow the struct field names.  Missing pointer fields become NULL,
    missing scalars 0.  Returns (desc, keepalive) -- keepalive holds the
    contiguous arrays the struct points into.
    """
    d = MeshDesc()
    keep = []
    d.abi_version = D4B_ABI_VERSION
    for name, typ in MeshDesc._fields_:
        if name == "abi_version" or name not in arrays:
            continue
        val = arrays[name]
        if name in _PTR_FIELDS:
            if val is None:
                continue
            a = np.ascontiguousarray(val, dtype=_NP_OF[typ]).reshape(-1)
            if a.size == 0:
                a = np.zeros(1, dtype=_NP_OF[typ])
            keep.append(a)
            setattr(d, name, a.ctypes.data_as(typ))
        else:
            v = np.asarray(val).reshape(-1)[0]
            setattr(d, name, float(v) if typ is C.c_double else int(v))
    if not d.mpisize:
        d.mpisize = 1
    return d, keep
