"""Synthetic unit-cube mesh description (BASELINE.json configs[1]/[2]).

Produces, with numpy, exactly the arrays the reference computes for
`[geometry] name=brick`, X,Y,Z in [0,1], one tree refined uniformly to
`level` (n = 2**level elements per direction, Morton order as p4est stores
them), uniform degree p, deg_quad_inc = 0, Gauss-Legendre quadrature,
face_h_type = FACE_H_EQ_VOLUME_DIV_AREA, homogeneous Dirichlet on all six
sides (src/Problems/Poisson/options.input).  tests/test_brick.py checks the
output against the fixtures the reference itself wrote (tests/golden/brick_*).

Everything is in the reference's layouts (include/d4est_b200.h,
d4b_mesh_desc_t); face "side" records are ordered (element, face) instead of
p4est_iterate order, which the library does not depend on.
"""
import numpy as np


def _part1by2(x):
    x = x.astype(np.int64) & 0x1FFFFF
    x = (x | (x << 32)) & 0x1F00000000FFFF
    x = (x | (x << 16)) & 0x1F0000FF0000FF
    x = (x | (x << 8)) & 0x100F00F00F00F00F
    x = (x | (x << 4)) & 0x10C30C30C30C30C3
    x = (x | (x << 2)) & 0x1249249249249249
    return x


def morton_encode(ix, iy, iz):
    return _part1by2(ix) | (_part1by2(iy) << 1) | (_part1by2(iz) << 2)


def _compact1by2(x):
    x = x & 0x1249249249249249
    x = (x | (x >> 2)) & 0x10C30C30C30C30C3
    x = (x | (x >> 4)) & 0x100F00F00F00F00F
    x = (x | (x >> 8)) & 0x1F0000FF0000FF
    x = (x | (x >> 16)) & 0x1F00000000FFFF
    x = (x | (x >> 32)) & 0x1FFFFF
    return x


def morton_decode(m):
    m = m.astype(np.int64)
    return _compact1by2(m), _compact1by2(m >> 1), _compact1by2(m >> 2)


def lgl_nodes(N):
    """LGL nodes on [-1,1] (numpy, for building test vectors only)."""
    p = N - 1
    if p == 1:
        return np.array([-1.0, 1.0])
    x = -np.cos(np.pi * np.arange(N) / p)
    for _ in range(100):
        P = np.polynomial.legendre.Legendre.basis(p)
        dP = P.deriv()
        ddP = dP.deriv()
        xi = x[1:-1]
        xi = xi - dP(xi) / ddP(xi)
        x[1:-1] = xi
    return x


def partition_offsets(E, world):
    """Equal-count Morton-curve partition, the same rule as p4est_partition with
    uniform weights: rank r owns global elements [E*r//world, E*(r+1)//world)."""
    return np.array([E * r // world for r in range(world + 1)], dtype=np.int64)


def trees_for_world(world):
    """Brick of unit-cube trees used for weak scaling: one tree per rank."""
    return {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}.get(world, (world, 1, 1))


def _global_id(X, Y, Z, n, trees):
    a, b, _ = trees
    t = (X // n) + a * ((Y // n) + b * (Z // n))
    return t * n ** 3 + morton_encode(X % n, Y % n, Z % n)


def _global_coords(g, n, trees):
    a, b, _ = trees
    t, m = g // n ** 3, g % n ** 3
    ix, iy, iz = morton_decode(m)
    return (t % a) * n + ix, ((t // a) % b) * n + iy, (t // (a * b)) * n + iz


def brick_desc(level, p, prefactor=2.0, penalty_fcn=0, rank=0, world=1, trees=(1, 1, 1)):
    """dict of arrays for make_desc()/Mesh() describing rank `rank`'s part of a
    brick of trees[0] x trees[1] x trees[2] unit-cube trees (p4est_connectivity_new_brick),
    each refined uniformly to `level` (element size h = 2**-level everywhere), elements
    numbered tree by tree in Morton order and split into `world` contiguous ranges.

    Ghost elements are the face neighbours owned by other ranks, numbered by
    (owner, global id) like p4est's ghost layer; mirrors are the local elements
    other ranks see as ghosts (src/Mesh/d4est_ghost.c:47-90,
    d4est_ghost_data.c:142-255 for the exchange the plan describes)."""
    n = 1 << level
    Eg = n ** 3 * trees[0] * trees[1] * trees[2]
    N = p + 1
    n3, n2 = N ** 3, N ** 2
    h = 1.0 / n
    offs = partition_offsets(Eg, world)
    start, end = int(offs[rank]), int(offs[rank + 1])
    E = end - start
    e = np.arange(E, dtype=np.int64)
    ix, iy, iz = _global_coords(e + start, n, trees)
    ext = np.array([trees[0] * n, trees[1] * n, trees[2] * n], dtype=np.int64)
    A = {}
    A["n_local"] = E
    A["n_sides"] = 6 * E
    A["local_nodes"] = E * n3
    A["local_nodes_quad"] = E * n3
    A["local_mortar_nodes_quad"] = 6 * E * n2
    A["deg"] = np.full(E, p, np.int32)
    A["deg_quad"] = np.full(E, p, np.int32)
    A["nodal_stride"] = e * n3
    A["quad_stride"] = e * n3
    LQ = E * n3
    A["J_quad"] = np.full(LQ, (0.5 * h) ** 3)
    rst = np.zeros((9, LQ))
    for i in range(3):
        rst[i * 3 + i, :] = 2.0 / h
    A["rst_xyz_quad"] = rst.reshape(-1)
    # sides: s = 6*e + f
    S = 6 * E
    f = np.tile(np.arange(6, dtype=np.int32), E)
    es = np.repeat(e, 6)
    coords = np.stack([np.repeat(ix, 6), np.repeat(iy, 6), np.repeat(iz, 6)])
    d = (f // 2).astype(np.int64)
    sgn = np.where(f % 2 == 1, 1, -1).astype(np.int64)
    nb = coords.copy()
    nb[d, np.arange(S)] += sgn
    interior = (nb[d, np.arange(S)] >= 0) & (nb[d, np.arange(S)] < ext[d])
    nbc = np.clip(nb, 0, (ext - 1)[:, None])
    nbr = _global_id(nbc[0], nbc[1], nbc[2], n, trees)   # global ids
    local_nbr = interior & (nbr >= start) & (nbr < end)
    ghost_nbr = interior & ~local_nbr
    # ghost layer: off-rank face neighbours, sorted by global id (= by owner, then id)
    ghost_gid = np.unique(nbr[ghost_nbr])
    n_ghost = ghost_gid.size
    ghost_owner = np.searchsorted(offs, ghost_gid, side="right") - 1
    A["n_ghost"] = n_ghost
    A["ghost_nodes"] = n_ghost * n3
    A["ghost_deg"] = np.full(n_ghost, p, np.int32)
    A["ghost_deg_quad"] = np.full(n_ghost, p, np.int32)
    A["ghost_nodal_stride"] = np.arange(n_ghost, dtype=np.int64) * n3
    A["ghost_proc_offsets"] = np.searchsorted(ghost_owner, np.arange(world + 1)).astype(np.int32)
    A["ghost_global_id"] = ghost_gid          # not part of the ABI; used by tests
    A["side_faces_m"] = np.ones(S, np.int32)
    A["side_faces_p"] = interior.astype(np.int32)
    A["side_f_m"] = f
    A["side_f_p"] = np.where(interior, f ^ 1, -1).astype(np.int32)
    e_m = np.full((S, 4), np.iinfo(np.int32).min, np.int32)
    e_p = np.full((S, 4), np.iinfo(np.int32).min, np.int32)
    e_m[:, 0] = es
    e_p[local_nbr, 0] = nbr[local_nbr] - start
    e_p[ghost_nbr, 0] = -(np.searchsorted(ghost_gid, nbr[ghost_nbr]) + 1)
    A["side_e_m"] = e_m.reshape(-1)
    A["side_e_p"] = e_p.reshape(-1)
    A["side_perm"] = np.zeros(S, np.int32)
    A["side_sub_order"] = np.tile(np.arange(4, dtype=np.int32), S)
    A["side_mortar_stride"] = np.arange(S, dtype=np.int64) * n2
    A["side_penalty_mult"] = np.ones(S)
    # mirrors: per peer q the local elements adjacent to q's elements, in local (= global) order
    side_owner = np.searchsorted(offs, nbr, side="right") - 1
    per_peer = [np.unique(es[ghost_nbr & (side_owner == q)]) if q != rank else np.zeros(0, np.int64)
                for q in range(world)]
    mirror_elem = np.unique(np.concatenate(per_peer)) if world > 1 else np.zeros(0, np.int64)
    A["mpirank"] = rank
    A["mpisize"] = world
    A["n_mirrors"] = mirror_elem.size
    A["mirror_elem"] = mirror_elem.astype(np.int32)
    A["mirror_proc_offsets"] = np.concatenate([[0], np.cumsum([a.size for a in per_peer])]).astype(np.int32)
    A["mirror_proc_mirrors"] = (np.searchsorted(mirror_elem, np.concatenate(per_peer)).astype(np.int32)
                                if world > 1 else np.zeros(0, np.int32))
    MP = S * n2
    A["sj_m_mortar_quad"] = np.full(MP, (0.5 * h) ** 2)
    nvec = np.zeros((S, 3, n2))
    nvec[np.arange(S), d, :] = sgn[:, None].astype(np.float64)
    A["n_m_mortar_quad"] = nvec.reshape(-1)
    del nvec
    dr = np.zeros((S, 9, n2))
    for i in range(3):
        dr[:, i + 3 * i, :] = 2.0 / h
    A["drst_dxyz_m_mortar_quad"] = dr.reshape(-1)
    A["drst_dxyz_p_mortar_quad_porder"] = A["drst_dxyz_m_mortar_quad"]
    A["hm_mortar_quad"] = np.full(MP, h)
    A["hp_mortar_quad"] = np.repeat(np.where(interior, h, 0.0), n2)
    A["sipg_penalty_prefactor"] = prefactor
    A["sipg_penalty_fcn"] = penalty_fcn
    A["bc_type"] = 0
    return A


def brick_xyz(level, p, rank=0, world=1, trees=(1, 1, 1)):
    """nodal coordinates (3, local_nodes) in the reference's node order."""
    n = 1 << level
    offs = partition_offsets(n ** 3 * trees[0] * trees[1] * trees[2], world)
    N = p + 1
    h = 1.0 / n
    ix, iy, iz = _global_coords(np.arange(offs[rank], offs[rank + 1], dtype=np.int64), n, trees)
    x1 = 0.5 * (lgl_nodes(N) + 1.0) * h
    X = ix[:, None, None, None] * h + x1[None, None, None, :]
    Y = iy[:, None, None, None] * h + x1[None, None, :, None]
    Z = iz[:, None, None, None] * h + x1[None, :, None, None]
    X, Y, Z = np.broadcast_arrays(X, Y, Z)
    return np.stack([X.reshape(-1), Y.reshape(-1), Z.reshape(-1)])


def brick_hierarchy(level, p, rank=0, world=1, trees=(1, 1, 1), **kw):
    """h-multigrid hierarchy of the brick, coarsest (one element per tree) first, as the
    reference builds it with p4est_coarsen_ext (d4est_solver_multigrid.c:984): every
    coarse element is the parent of the 8 consecutive (z-order) fine elements
    8e..8e+7; the degree stays p (coarse degree = min of the children,
    callbacks.h:51-76).  Needs one tree per rank when world > 1 so that families and
    their parents share a rank at every level.  Returns (descs, transfers) with
    transfers[l] = (child_first, child_count) between level l and l+1."""
    assert world == 1 or world == trees[0] * trees[1] * trees[2], "multi-rank hierarchy: one tree per rank"
    descs = [brick_desc(l, p, rank=rank, world=world, trees=trees, **kw) for l in range(level + 1)]
    transfers = []
    for l in range(level):
        E = descs[l]["n_local"]
        transfers.append((8 * np.arange(E, dtype=np.int32), np.full(E, 8, np.int32)))
    return descs, transfers
