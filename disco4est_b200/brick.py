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


ROOT_LEN = 1 << 30   # P4EST_ROOT_LEN


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


def brick_desc(level, p, prefactor=2.0, penalty_fcn=0, rank=0, world=1, trees=(1, 1, 1), geometry=True):
    """dict of arrays for make_desc()/Mesh() describing rank `rank`'s part of a
    brick of trees[0] x trees[1] x trees[2] unit-cube trees (p4est_connectivity_new_brick),
    each refined uniformly to `level` (element size h = 2**-level everywhere), elements
    numbered tree by tree in Morton order and split into `world` contiguous ranges.

    Ghost elements are the face neighbours owned by other ranks, numbered by
    (owner, global id) like p4est's ghost layer; mirrors are the local elements
    other ranks see as ghosts (src/Mesh/d4est_ghost.c:47-90,
    d4est_ghost_data.c:142-255 for the exchange the plan describes).

    geometry=False: returns (topology, geom) for Mesh(topology, geom=geom) instead -- no per-point array is
    built on the host, the factors are generated and folded on the device (face_h = tree h, which is the
    element width 2**-level exactly, like the hm = hp = h of the full description)."""
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
    if geometry:
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
    A["sipg_penalty_prefactor"] = prefactor
    A["sipg_penalty_fcn"] = penalty_fcn
    A["bc_type"] = 0
    if not geometry:
        dq = ROOT_LEN >> level
        gx, gy, gz = _global_coords(ghost_gid, n, trees)
        geom = {"geom_type": "brick", "face_h_type": "tree_h", "X0": 0.0, "X1": 1.0, "Y0": 0.0, "Y1": 1.0,
                "Z0": 0.0, "Z1": 1.0,
                "elem_tree": ((e + start) // n ** 3).astype(np.int32),
                "elem_q": (np.stack([ix % n, iy % n, iz % n], axis=1) * dq).astype(np.int32).reshape(-1),
                "elem_dq": np.full(E, dq, np.int32),
                "ghost_tree": (ghost_gid // n ** 3).astype(np.int32),
                "ghost_q": (np.stack([gx % n, gy % n, gz % n], axis=1) * dq).astype(np.int32).reshape(-1),
                "ghost_dq": np.full(n_ghost, dq, np.int32)}
        return A, geom
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
        E = (descs[l][0] if isinstance(descs[l], tuple) else descs[l])["n_local"]   # geometry=False: (topology, geom)
        transfers.append((8 * np.arange(E, dtype=np.int32), np.full(E, 8, np.int32)))
    return descs, transfers


def brick_amr_desc(level, p, prefactor=2.0, penalty_fcn=0, face_h_type="volume_div_area"):
    """(topology, geom) of the unit cube refined uniformly to `level` with the octant [0,1/2]^3 refined once
    more: a 2:1-balanced forest with hanging faces on the three interface planes, the kind of mesh one AMR
    step of the reference produces (ConstantDensityStar driver, BASELINE.json configs[0]).  Elements in
    Morton order like p4est (the refined octant is Morton-contiguous and comes first); flux side records as
    the reference's face iteration emits them (src/Mesh/d4est_mortars.c:601-802): one per (element, face)
    on conforming faces, one (1 -> 4) per coarse element and one (4 -> 1) per group of four fine elements
    on hanging faces.  Geometric factors come from the device generator (Mesh(topology, geom=geom))."""
    n = 1 << level                      # coarse cells per direction
    nh = n // 2
    root = 1 << 30
    N2 = (p + 1) ** 2
    n3 = (p + 1) ** 3
    # --- elements: fine block (level + 1, coordinates in [0, n)^3 of the 2n grid), then the coarse rest
    nf = n ** 3
    fx, fy, fz = morton_decode(np.arange(nf, dtype=np.int64))
    mc = np.arange(n ** 3 // 8, n ** 3, dtype=np.int64)
    cx, cy, cz = morton_decode(mc)
    ncz = mc.size
    E = nf + ncz
    fine_id = lambda x, y, z: morton_encode(x, y, z)
    coarse_id = lambda x, y, z: nf + morton_encode(x, y, z) - n ** 3 // 8
    sides = {k: [] for k in ("faces_m", "faces_p", "f_m", "f_p", "e_m", "e_p")}

    def add(faces_m, faces_p, f_m, f_p, e_m, e_p):
        k = len(f_m)
        pad = np.full((k, 4), np.iinfo(np.int32).min, np.int64)
        em, ep = pad.copy(), pad.copy()
        em[:, :e_m.shape[1]] = e_m
        if e_p is not None:
            ep[:, :e_p.shape[1]] = e_p
        sides["faces_m"].append(np.full(k, faces_m)); sides["faces_p"].append(np.full(k, faces_p))
        sides["f_m"].append(f_m); sides["f_p"].append(f_p)
        sides["e_m"].append(em); sides["e_p"].append(ep)

    tang = {0: (1, 2), 1: (0, 2), 2: (0, 1)}
    for f in range(6):
        d, sg = f // 2, (1 if f % 2 else -1)
        ta, tb = tang[d]
        # fine cells: neighbour at the same level inside the block, domain boundary, or the coarse cell behind
        c = np.stack([fx, fy, fz]); nb = c.copy(); nb[d] += sg
        ids = np.arange(nf)
        inside = (nb[d] >= 0) & (nb[d] < n)
        add(1, 1, np.full(inside.sum(), f), np.full(inside.sum(), f ^ 1), ids[inside][:, None],
            fine_id(nb[0][inside], nb[1][inside], nb[2][inside])[:, None])
        bnd = nb[d] < 0
        add(1, 0, np.full(bnd.sum(), f), np.full(bnd.sum(), -1), ids[bnd][:, None], None)
        if sg > 0:   # + face on the interface plane: groups of four fine cells against one coarse cell
            lead = (nb[d] >= n) & (c[ta] % 2 == 0) & (c[tb] % 2 == 0)
            g = c[:, lead]
            grp = []
            for i in range(4):
                q = g.copy(); q[ta] += i & 1; q[tb] += i >> 1
                grp.append(fine_id(q[0], q[1], q[2]))
            cc = g // 2; cc[d] = nh
            add(4, 1, np.full(lead.sum(), f), np.full(lead.sum(), f ^ 1), np.stack(grp, axis=1),
                coarse_id(cc[0], cc[1], cc[2])[:, None])
        # coarse cells
        c = np.stack([cx, cy, cz]); nb = c.copy(); nb[d] += sg
        ids = nf + np.arange(ncz)
        bnd = (nb[d] < 0) | (nb[d] >= n)
        add(1, 0, np.full(bnd.sum(), f), np.full(bnd.sum(), -1), ids[bnd][:, None], None)
        into_fine = ~bnd & (nb[0] < nh) & (nb[1] < nh) & (nb[2] < nh)
        same = ~bnd & ~into_fine
        add(1, 1, np.full(same.sum(), f), np.full(same.sum(), f ^ 1), ids[same][:, None],
            coarse_id(nb[0][same], nb[1][same], nb[2][same])[:, None])
        g = 2 * nb[:, into_fine]
        g[d] = n - 1                                  # only reached with sg < 0: the fine layer below the plane
        grp = []
        for i in range(4):
            q = g.copy(); q[ta] += i & 1; q[tb] += i >> 1
            grp.append(fine_id(q[0], q[1], q[2]))
        add(1, 4, np.full(into_fine.sum(), f), np.full(into_fine.sum(), f ^ 1), ids[into_fine][:, None],
            np.stack(grp, axis=1) if into_fine.any() else np.zeros((0, 4), np.int64))
    cat = {k: np.concatenate(v) for k, v in sides.items()}
    S = cat["f_m"].size
    mort = np.maximum(cat["faces_m"], cat["faces_p"]).astype(np.int64)
    A = {"n_local": E, "n_ghost": 0, "n_sides": S, "local_nodes": E * n3, "local_nodes_quad": E * n3, "ghost_nodes": 0,
         "deg": np.full(E, p, np.int32), "deg_quad": np.full(E, p, np.int32),
         "nodal_stride": np.arange(E, dtype=np.int64) * n3, "quad_stride": np.arange(E, dtype=np.int64) * n3,
         "side_faces_m": cat["faces_m"].astype(np.int32), "side_faces_p": cat["faces_p"].astype(np.int32),
         "side_f_m": cat["f_m"].astype(np.int32), "side_f_p": cat["f_p"].astype(np.int32),
         "side_e_m": cat["e_m"].astype(np.int32).reshape(-1), "side_e_p": cat["e_p"].astype(np.int32).reshape(-1),
         "side_perm": np.zeros(S, np.int32), "side_sub_order": np.tile(np.arange(4, dtype=np.int32), S),
         "side_mortar_stride": np.concatenate([[0], np.cumsum(mort * N2)[:-1]]).astype(np.int64),
         "local_mortar_nodes_quad": int((mort * N2).sum()), "side_penalty_mult": np.ones(S),
         "sipg_penalty_prefactor": prefactor, "sipg_penalty_fcn": penalty_fcn, "bc_type": 0,
         "mpirank": 0, "mpisize": 1, "n_mirrors": 0}
    dqf, dqc = root >> (level + 1), root >> level
    geom = {"geom_type": "brick", "face_h_type": face_h_type, "X0": 0.0, "X1": 1.0, "Y0": 0.0, "Y1": 1.0, "Z0": 0.0, "Z1": 1.0,
            "elem_tree": np.zeros(E, np.int32),
            "elem_q": np.concatenate([np.stack([fx, fy, fz], axis=1) * dqf, np.stack([cx, cy, cz], axis=1) * dqc]).astype(np.int32).reshape(-1),
            "elem_dq": np.concatenate([np.full(nf, dqf), np.full(ncz, dqc)]).astype(np.int32)}
    return A, geom
