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
    ]


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
