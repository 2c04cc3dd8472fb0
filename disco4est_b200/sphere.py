"""Synthetic 13-tree cubed-sphere mesh description (BASELINE.json configs[3]).

Topology of `[geometry] name = cubed_sphere` (p8est_connectivity_new_sphere: trees 0-5 outer
shell, 6-11 inner shell, 12 centre cube; d4est_geometry_cubed_sphere.c:315-400) with every tree
refined uniformly to `level`, elements numbered tree by tree in Morton order like p4est, uniform
degree or a seeded random degree per element ("hp-adapted variable degree": conforming faces
between elements of unequal degree; hanging faces come from the reference's own AMR fixtures).

Only the topology is built here (numpy); the geometric factors are generated on the device
(d4b_mesh_create_with_geometry, csrc/d4b_geom.cu).  Inter-tree face connections are found
geometrically: the corners of every tree-boundary face are pushed through the tree maps and
matched, which also yields the face-node permutation of include/d4est_b200.h (side_perm) without
restating p4est's orientation tables.  tests/test_sphere.py checks the result against the
reference's own sphere fixture (same side records, same operator).
"""
import numpy as np

from .brick import morton_decode, morton_encode, partition_offsets

ROOT_LEN = 1 << 30
_TANGENT = {0: (1, 2), 1: (0, 2), 2: (0, 1)}   # face normal direction -> (a, b), a fastest


def _unit_sphere_X(tree, tc):
    """physical coordinates of tree coordinates tc (n, 3) in [0,1]^3 on the R = 1, 2, 3 sphere
    (connectivity matching only; the real radii go to the device generator)"""
    tree = np.asarray(tree)
    a, b, c = 2.0 * tc[:, 0] - 1.0, 2.0 * tc[:, 1] - 1.0, tc[:, 2] + 1.0
    T, U = np.tan(a * np.pi / 4), np.tan(b * np.pi / 4)
    outer = tree < 6
    w = c - 1.0
    x = np.where(outer, T, (1.0 - w) * a + w * T)
    y = np.where(outer, U, (1.0 - w) * b + w * U)
    R = np.where(outer, 2.0 * (2.0 - c) + 3.0 * (c - 1.0), 1.0 * (2.0 - c) + 2.0 * (c - 1.0))
    q = R / np.sqrt(np.where(outer, T * T + U * U + 1.0, 3.0 - 2.0 * w + w * (T * T + U * U)))
    top = np.stack([q * x, q * y, q], axis=1)
    out = np.empty_like(top)
    src = {0: (0, 2, 1), 1: (0, 1, 2), 2: (0, 2, 1), 3: (2, 0, 1), 4: (1, 0, 2), 5: (2, 0, 1)}
    sgn = {0: (1, -1, 1), 1: (1, 1, 1), 2: (1, 1, -1), 3: (1, -1, -1), 4: (-1, -1, -1), 5: (-1, -1, 1)}
    for k in range(6):
        m = (tree % 6 == k) & (tree < 12)
        for i in range(3):
            out[m, i] = sgn[k][i] * top[m, src[k][i]]
    cube = tree == 12
    out[cube] = (2.0 * tc[cube] - 1.0) / np.sqrt(3.0)
    return out


def _perm_index(n, perm, a, b):
    row, col = (a, b) if perm & 4 else (b, a)
    if perm & 2:
        row = n - 1 - row
    if perm & 1:
        col = n - 1 - col
    return row * n + col


def sphere_desc(level, p=3, pmax=None, seed=102321, prefactor=2.0, penalty_fcn=0, radii=(1.0, 2.0, 3.0),
                compactify=(0, 0), face_h_type="j_div_sj_quad", graded=False):
    """(topology, geom): dicts for Mesh(topology, geom=geom) describing the whole 13-tree sphere on
    one rank (disco4est_b200.partition splits it).  pmax: draw every element's degree uniformly
    from 1..pmax (seeded) instead of the uniform degree p.  graded: degree p in the centre cube and
    the inner shell, then decreasing outwards in radial layers of the outer shell down to
    max(p - 2, 1) -- the kind of mesh p-adaptation of a centrally concentrated source produces
    (ConstantDensityStar, two punctures): p-jumps only on the layer interfaces."""
    from scipy.spatial import cKDTree

    n = 1 << level
    n3e = n ** 3
    E = 13 * n3e
    e = np.arange(E, dtype=np.int64)
    tree = e // n3e
    ix, iy, iz = morton_decode(e % n3e)
    ijk = np.stack([ix, iy, iz], axis=1)
    if pmax is not None:
        deg = np.random.default_rng(seed).integers(1, pmax + 1, size=E).astype(np.int64)
    else:
        deg = np.full(E, p, np.int64)
    if graded:
        outer = tree < 6
        layer = (3 * iz) // n            # radial third of the outer shell (iz is the radial index there)
        deg = np.where(outer, np.maximum(p - layer, max(p - 2, 1)), p).astype(np.int64)

    S = 6 * E
    f = np.tile(np.arange(6), E)
    es = np.repeat(e, 6)
    d = f // 2
    sg = np.where(f % 2 == 1, 1, -1)
    nb = np.repeat(ijk, 6, axis=0)
    nb[np.arange(S), d] += sg
    inside = (nb[np.arange(S), d] >= 0) & (nb[np.arange(S), d] < n)
    nbc = np.clip(nb, 0, n - 1)
    e_p = np.where(inside, np.repeat(tree, 6) * n3e + morton_encode(nbc[:, 0], nbc[:, 1], nbc[:, 2]), -1)
    f_p = np.where(inside, f ^ 1, -1)
    perm = np.zeros(S, np.int64)

    # tree-boundary faces: corners through the tree maps, matched by face centre
    bs = np.nonzero(~inside)[0]
    eb, fb, db = es[bs], f[bs], d[bs]
    corners = np.zeros((bs.size, 4, 3))
    h = 1.0 / n
    for cidx in range(4):
        ca, cb = cidx & 1, cidx >> 1
        off = np.zeros((bs.size, 3))
        for nd in range(3):
            m = db == nd
            ta, tb = _TANGENT[nd]
            off[m, ta], off[m, tb] = ca, cb
            off[m, nd] = (fb[m] % 2)
        corners[:, cidx, :] = _unit_sphere_X(tree[eb], (ijk[eb] + off) * h)
    centre = corners.mean(axis=1)
    dist, idx = cKDTree(centre).query(centre, k=2)
    me = np.arange(bs.size)
    other = np.where(idx[:, 0] == me, idx[:, 1], idx[:, 0])   # coincident faces come back in either order
    partner = np.where(dist[:, 1] < 1e-9, other, -1)
    has = partner >= 0
    assert np.array_equal(partner[partner[has]], np.nonzero(has)[0]), "face matching is not symmetric"
    pi, pj = np.nonzero(has)[0], partner[has]
    err = np.zeros((8, pi.size))
    for code in range(8):
        for cidx in range(4):
            kp = _perm_index(2, code, cidx & 1, cidx >> 1)
            err[code] = np.maximum(err[code], np.abs(corners[pj, kp, :] - corners[pi, cidx, :]).max(axis=1))
    code = err.argmin(axis=0)
    assert (err[code, np.arange(pi.size)] < 1e-9).all(), "no face permutation matches"
    e_p[bs[pi]], f_p[bs[pi]], perm[bs[pi]] = eb[pj], fb[pj], code
    interior = e_p >= 0

    A = {}
    n3 = (deg + 1) ** 3
    A["n_local"], A["n_ghost"], A["n_sides"] = E, 0, S
    A["local_nodes"] = A["local_nodes_quad"] = int(n3.sum())
    A["ghost_nodes"] = 0
    A["deg"] = deg.astype(np.int32)
    A["deg_quad"] = deg.astype(np.int32)
    A["nodal_stride"] = np.concatenate([[0], np.cumsum(n3)[:-1]]).astype(np.int64)
    A["quad_stride"] = A["nodal_stride"]
    degq_p = np.where(interior, deg[np.clip(e_p, 0, E - 1)], 0)
    nq2 = (np.maximum(np.repeat(deg, 6), degq_p) + 1) ** 2
    A["side_mortar_stride"] = np.concatenate([[0], np.cumsum(nq2)[:-1]]).astype(np.int64)
    A["local_mortar_nodes_quad"] = int(nq2.sum())
    A["side_faces_m"] = np.ones(S, np.int32)
    A["side_faces_p"] = interior.astype(np.int32)
    A["side_f_m"] = f.astype(np.int32)
    A["side_f_p"] = f_p.astype(np.int32)
    pad = np.iinfo(np.int32).min
    em4 = np.full((S, 4), pad, np.int32)
    ep4 = np.full((S, 4), pad, np.int32)
    em4[:, 0] = es
    ep4[interior, 0] = e_p[interior]
    A["side_e_m"], A["side_e_p"] = em4.reshape(-1), ep4.reshape(-1)
    A["side_perm"] = perm.astype(np.int32)
    A["side_sub_order"] = np.tile(np.arange(4, dtype=np.int32), S)
    A["side_penalty_mult"] = np.ones(S)
    A["sipg_penalty_prefactor"], A["sipg_penalty_fcn"], A["bc_type"] = prefactor, penalty_fcn, 0
    A["mpirank"], A["mpisize"], A["n_mirrors"] = 0, 1, 0

    dq = ROOT_LEN >> level
    geom = {"geom_type": "cubed_sphere", "face_h_type": face_h_type,
            "R0": radii[0], "R1": radii[1], "R2": radii[2],
            "compactify_inner_shell": compactify[0], "compactify_outer_shell": compactify[1],
            "elem_tree": tree.astype(np.int32), "elem_q": (ijk * dq).astype(np.int32).reshape(-1),
            "elem_dq": np.full(E, dq, np.int32)}
    return A, geom


def partition_geom(geom, B, rank, world, n_global, offsets=None):
    """geometry dict of rank `rank` matching B = partition_desc(A, rank, world)"""
    offs = np.asarray(offsets if offsets is not None else partition_offsets(n_global, world)).astype(np.int64)
    start, end = int(offs[rank]), int(offs[rank + 1])
    gid = np.asarray(B["ghost_global_id"]).astype(np.int64)
    g = dict(geom)
    q = np.asarray(geom["elem_q"]).reshape(-1, 3)
    g["elem_tree"], g["elem_q"], g["elem_dq"] = geom["elem_tree"][start:end], q[start:end].reshape(-1), geom["elem_dq"][start:end]
    g["ghost_tree"], g["ghost_q"], g["ghost_dq"] = geom["elem_tree"][gid], q[gid].reshape(-1), geom["elem_dq"][gid]
    return g
