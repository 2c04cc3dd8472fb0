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


def brick_desc(level, p, prefactor=2.0, penalty_fcn=0):
    """dict of arrays for make_desc()/Mesh() describing the n^3 unit cube."""
    n = 1 << level
    E = n ** 3
    N = p + 1
    n3, n2 = N ** 3, N ** 2
    h = 1.0 / n
    e = np.arange(E, dtype=np.int64)
    ix, iy, iz = morton_decode(e)
    A = {}
    A["n_local"] = E
    A["n_ghost"] = 0
    A["n_sides"] = 6 * E
    A["local_nodes"] = E * n3
    A["local_nodes_quad"] = E * n3
    A["local_mortar_nodes_quad"] = 6 * E * n2
    A["ghost_nodes"] = 0
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
    interior = (nb[d, np.arange(S)] >= 0) & (nb[d, np.arange(S)] < n)
    nbc = np.clip(nb, 0, n - 1)
    nbr = morton_encode(nbc[0], nbc[1], nbc[2])
    A["side_faces_m"] = np.ones(S, np.int32)
    A["side_faces_p"] = interior.astype(np.int32)
    A["side_f_m"] = f
    A["side_f_p"] = np.where(interior, f ^ 1, -1).astype(np.int32)
    e_m = np.full((S, 4), np.iinfo(np.int32).min, np.int32)
    e_p = np.full((S, 4), np.iinfo(np.int32).min, np.int32)
    e_m[:, 0] = es
    e_p[interior, 0] = nbr[interior]
    A["side_e_m"] = e_m.reshape(-1)
    A["side_e_p"] = e_p.reshape(-1)
    A["side_perm"] = np.zeros(S, np.int32)
    A["side_sub_order"] = np.tile(np.arange(4, dtype=np.int32), S)
    A["side_mortar_stride"] = np.arange(S, dtype=np.int64) * n2
    A["side_penalty_mult"] = np.ones(S)
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


def brick_xyz(level, p):
    """nodal coordinates (3, local_nodes) in the reference's node order."""
    n = 1 << level
    E = n ** 3
    N = p + 1
    h = 1.0 / n
    ix, iy, iz = morton_decode(np.arange(E, dtype=np.int64))
    x1 = 0.5 * (lgl_nodes(N) + 1.0) * h
    X = ix[:, None, None, None] * h + x1[None, None, None, :]
    Y = iy[:, None, None, None] * h + x1[None, None, :, None]
    Z = iz[:, None, None, None] * h + x1[None, :, None, None]
    X, Y, Z = np.broadcast_arrays(X, Y, Z)
    return np.stack([X.reshape(-1), Y.reshape(-1), Z.reshape(-1)])
