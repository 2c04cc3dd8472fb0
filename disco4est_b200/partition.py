"""Split a one-rank flattened mesh description into per-rank descriptions.

The reference keeps p4est's partition: elements in tree / Morton order, cut into `world`
contiguous ranges of (almost) equal element count (p4est_partition without weights,
src/driver.c:137,148); each rank sees its own elements, a ghost layer of off-rank neighbours
and the flux side records whose (-) side holds at least one local element
(d4est_mortars_compute_flux_on_local_elements_aux, src/Mesh/d4est_mortars.c:640-760: "unless every
hanging face is a ghost, we calculate the flux").  The serial reference build in oracle/_ref
cannot produce partitioned meshes (no MPI), so the multi-rank tests and bench derive them here from
the one-rank description of ANY mesh (hp, hanging faces, curved): same arrays, same layouts,
restricted to what rank `rank` would hold, plus the mirror / ghost exchange plan of
d4est_ghost_data_exchange (src/Mesh/d4est_ghost_data.c:142-255).

Ghost elements are numbered by global element id, i.e. by owner rank and then position, like
p4est's ghost layer; only face-connected ghosts are kept (the flux needs no others).
"""
import numpy as np

from .brick import partition_offsets

_SIDE_FIELDS = ("side_faces_m", "side_faces_p", "side_f_m", "side_f_p", "side_perm", "side_penalty_mult")
_MORTAR_FIELDS = (("sj_m_mortar_quad", 1), ("n_m_mortar_quad", 3), ("drst_dxyz_m_mortar_quad", 9),
                  ("drst_dxyz_p_mortar_quad_porder", 9), ("hm_mortar_quad", 1), ("hp_mortar_quad", 1),
                  ("dirichlet_g_mortar_quad", 1), ("robin_coeff_mortar_quad", 1), ("robin_rhs_mortar_quad", 1))


def _i(x):
    """scalars may arrive as 1-element arrays (fixtures)"""
    return int(np.asarray(x).reshape(-1)[0])


def _side_masks(A):
    S = _i(A["n_sides"])
    e_m = np.asarray(A["side_e_m"]).reshape(S, 4).astype(np.int64)
    e_p = np.asarray(A["side_e_p"]).reshape(S, 4).astype(np.int64)
    lane = np.arange(4)[None, :]
    valid_m = lane < np.asarray(A["side_faces_m"])[:, None]
    valid_p = lane < np.asarray(A["side_faces_p"])[:, None]
    return e_m, e_p, valid_m, valid_p


def _kept_sides(e_m, valid_m, start, end):
    return (((e_m >= start) & (e_m < end)) & valid_m).any(axis=1)


def _side_totals(A):
    """mortar quadrature nodes of every side, from the strides (sides tile the mortar arrays)"""
    stride = np.asarray(A["side_mortar_stride"]).astype(np.int64)
    order = np.argsort(stride, kind="stable")
    ends = np.concatenate([stride[order][1:], [_i(A["local_mortar_nodes_quad"])]])
    total = np.empty_like(stride)
    total[order] = ends - stride[order]
    assert (total > 0).all(), "sides must tile the mortar arrays"
    return total


def _gather_blocks(src, starts, lengths):
    """concatenate src[starts[i] : starts[i] + lengths[i]]"""
    if len(starts) == 0:
        return np.zeros(0, src.dtype)
    offs = np.concatenate([[0], np.cumsum(lengths)])
    idx = np.arange(offs[-1]) - np.repeat(offs[:-1], lengths) + np.repeat(starts, lengths)
    return np.ascontiguousarray(src[idx])


def partition_desc(A, rank, world, offsets=None):
    """dict of arrays (make_desc()/Mesh()/Oracle() input) for rank `rank` of `world`."""
    assert _i(A.get("n_ghost", 0)) == 0, "partition_desc expects a one-rank description"
    E = _i(A["n_local"])
    offs = np.asarray(offsets if offsets is not None else partition_offsets(E, world)).astype(np.int64)
    start, end = int(offs[rank]), int(offs[rank + 1])
    deg = np.asarray(A["deg"]).astype(np.int64)
    deg_quad = np.asarray(A["deg_quad"]).astype(np.int64)
    nodal = np.asarray(A["nodal_stride"]).astype(np.int64)
    quad = np.asarray(A["quad_stride"]).astype(np.int64)
    n3 = (deg + 1) ** 3
    q3 = (deg_quad + 1) ** 3
    # elements are laid out back to back in element order (d4est_mesh.c:2394-2490)
    assert np.array_equal(nodal, np.concatenate([[0], np.cumsum(n3)[:-1]]))
    assert np.array_equal(quad, np.concatenate([[0], np.cumsum(q3)[:-1]]))
    node0 = int(nodal[start]) if start < E else _i(A["local_nodes"])
    node1 = int(nodal[end]) if end < E else _i(A["local_nodes"])
    q0 = int(quad[start]) if start < E else _i(A["local_nodes_quad"])
    q1 = int(quad[end]) if end < E else _i(A["local_nodes_quad"])
    LQ = _i(A["local_nodes_quad"])

    e_m, e_p, valid_m, valid_p = _side_masks(A)
    keep = _kept_sides(e_m, valid_m, start, end)
    ks = np.nonzero(keep)[0]
    refs = np.concatenate([e_m[keep][valid_m[keep]], e_p[keep][valid_p[keep]]])
    ghost_gid = np.unique(refs[(refs < start) | (refs >= end)])
    n_ghost = ghost_gid.size

    def remap(x, valid):
        out = np.full(x.shape, np.iinfo(np.int32).min, np.int64)
        loc = valid & (x >= start) & (x < end)
        gh = valid & ~loc
        out[loc] = x[loc] - start
        out[gh] = -(np.searchsorted(ghost_gid, x[gh]) + 1)
        return out.astype(np.int32)

    B = {}
    B["n_local"] = end - start
    B["n_ghost"] = n_ghost
    B["n_sides"] = ks.size
    B["local_nodes"] = node1 - node0
    B["local_nodes_quad"] = q1 - q0
    B["deg"] = np.ascontiguousarray(A["deg"][start:end]).astype(np.int32)
    B["deg_quad"] = np.ascontiguousarray(A["deg_quad"][start:end]).astype(np.int32)
    B["nodal_stride"] = np.ascontiguousarray(nodal[start:end] - node0)
    B["quad_stride"] = np.ascontiguousarray(quad[start:end] - q0)
    if A.get("J_quad") is not None:   # absent when the factors are generated on the device (sphere.py)
        B["J_quad"] = np.ascontiguousarray(A["J_quad"][q0:q1])
        B["rst_xyz_quad"] = np.ascontiguousarray(np.asarray(A["rst_xyz_quad"]).reshape(9, LQ)[:, q0:q1]).reshape(-1)

    B["ghost_deg"] = deg[ghost_gid].astype(np.int32)
    B["ghost_deg_quad"] = deg_quad[ghost_gid].astype(np.int32)
    gn3 = n3[ghost_gid]
    B["ghost_nodal_stride"] = np.concatenate([[0], np.cumsum(gn3)[:-1]]).astype(np.int64) if n_ghost else np.zeros(0, np.int64)
    B["ghost_nodes"] = int(gn3.sum())
    owner = np.searchsorted(offs, ghost_gid, side="right") - 1
    B["ghost_proc_offsets"] = np.searchsorted(owner, np.arange(world + 1)).astype(np.int32)
    B["ghost_global_id"] = ghost_gid               # not part of the ABI; used by tests
    B["global_node_range"] = (node0, node1)        # likewise

    for k in _SIDE_FIELDS:
        B[k] = np.ascontiguousarray(np.asarray(A[k])[ks])
    B["side_e_m"] = remap(e_m[ks], valid_m[ks]).reshape(-1)
    B["side_e_p"] = remap(e_p[ks], valid_p[ks]).reshape(-1)
    B["side_sub_order"] = np.ascontiguousarray(np.asarray(A["side_sub_order"]).reshape(-1, 4)[ks]).reshape(-1)

    total = _side_totals(A)
    stride = np.asarray(A["side_mortar_stride"]).astype(np.int64)
    kt, kstride = total[ks], stride[ks]
    new_stride = np.concatenate([[0], np.cumsum(kt)[:-1]]).astype(np.int64) if ks.size else np.zeros(0, np.int64)
    B["side_mortar_stride"] = new_stride
    B["local_mortar_nodes_quad"] = int(kt.sum())
    for name, mult in _MORTAR_FIELDS:
        if A.get(name) is None:
            continue
        B[name] = _gather_blocks(np.asarray(A[name]), mult * kstride, mult * kt)

    # mirrors: local elements that some other rank references from the sides it keeps
    per_peer = []
    for q in range(world):
        if q == rank:
            per_peer.append(np.zeros(0, np.int64))
            continue
        qk = _kept_sides(e_m, valid_m, int(offs[q]), int(offs[q + 1]))
        r = np.concatenate([e_m[qk][valid_m[qk]], e_p[qk][valid_p[qk]]])
        per_peer.append(np.unique(r[(r >= start) & (r < end)]) - start)
    mirror_elem = np.unique(np.concatenate(per_peer)) if world > 1 else np.zeros(0, np.int64)
    B["mpirank"], B["mpisize"] = rank, world
    B["n_mirrors"] = mirror_elem.size
    B["mirror_elem"] = mirror_elem.astype(np.int32)
    B["mirror_proc_offsets"] = np.concatenate([[0], np.cumsum([a.size for a in per_peer])]).astype(np.int32)
    B["mirror_proc_mirrors"] = (np.searchsorted(mirror_elem, np.concatenate(per_peer)).astype(np.int32)
                                if world > 1 else np.zeros(0, np.int32))
    B["sipg_penalty_prefactor"] = float(np.asarray(A["sipg_penalty_prefactor"]).reshape(-1)[0])
    B["sipg_penalty_fcn"] = _i(A["sipg_penalty_fcn"])
    B["bc_type"] = _i(A["bc_type"])
    return B
