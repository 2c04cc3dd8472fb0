"""CPU restatement of the reference's geometric-factor computation (TEST INFRASTRUCTURE).

Follows d4est_mesh_data_compute (src/Mesh/d4est_mesh.c:2543-2755),
d4est_mesh_compute_mortar_quadrature_quantities (:481-1140) and
d4est_mortars_compute_geometric_data_on_mortar_aux (src/Mesh/d4est_mortars.c:10-277)
for [geometry] DX_compute_method = analytic, JAC_compute_method = numerical.

Only the maps X are restated (d4est_geometry_brick.c, d4est_geometry_cubed_sphere.c:
127-167, :315-400); dX/d(rst) comes from complex-step differentiation of X, which is
exact to round-off for these analytic maps and independent of the hand-derived
derivatives in disco4est_b200/csrc/d4b_geom.cu.  Pinned against the arrays the
unmodified reference wrote into tests/golden (tests/test_geometry.py).
"""
import numpy as np

ROOT = float(1 << 30)
FACE_CORNERS = [[0, 2, 4, 6], [1, 3, 5, 7], [0, 1, 4, 5], [2, 3, 6, 7], [0, 1, 2, 3], [4, 5, 6, 7]]


def gauss(n):
    x, w = np.polynomial.legendre.leggauss(n)
    return x, w


def lobatto(n):
    p = n - 1
    if p == 1:
        return np.array([-1.0, 1.0]), np.array([1.0, 1.0])
    P = np.polynomial.legendre.Legendre.basis(p)
    x = np.concatenate([[-1.0], np.sort(P.deriv().roots().real), [1.0]])
    w = 2.0 / (p * (p + 1) * P(x) ** 2)
    return x, w


def shell_radius(Ra, Rb, compactify, c):
    if compactify:
        m = 1.0 / ((1.0 / Rb) - (1.0 / Ra))
        t = (Ra - 2.0 * Rb) / (Ra - Rb)
        return m / (c - t)
    return Ra * (2.0 - c) + Rb * (c - 1.0)


def X(g, tree, q, dq, rst):
    """physical coordinates (3, n) of reference points rst (3, n) (may be complex)"""
    tc = [(q[d] + dq * 0.5 * (rst[d] + 1.0)) / ROOT for d in range(3)]
    if g["geom_type"] == 0:
        return np.stack([(g["X1"] - g["X0"]) * tc[0] + g["X0"], (g["Y1"] - g["Y0"]) * tc[1] + g["Y0"],
                         (g["Z1"] - g["Z0"]) * tc[2] + g["Z0"]])
    if tree == 12:
        cl = g["R0"] / np.sqrt(3.0)
        return np.stack([(2.0 * t - 1.0) * cl for t in tc])
    a, b, c = 2.0 * tc[0] - 1.0, 2.0 * tc[1] - 1.0, tc[2] + 1.0
    if tree < 6:
        x, y = np.tan(a * np.pi / 4), np.tan(b * np.pi / 4)
        R = shell_radius(g["R1"], g["R2"], g["compactify_outer_shell"], c)
        qq = R / np.sqrt(x * x + y * y + 1.0)
    else:
        p = 2.0 - c
        tx, ty = np.tan(a * np.pi / 4), np.tan(b * np.pi / 4)
        x, y = p * a + (1.0 - p) * tx, p * b + (1.0 - p) * ty
        R = shell_radius(g["R0"], g["R1"], g["compactify_inner_shell"], c)
        qq = R / np.sqrt(1.0 + (1.0 - p) * (tx * tx + ty * ty) + 2.0 * p)
    k = tree % 6
    if k == 0:
        return np.stack([qq * x, -qq, qq * y])
    if k == 1:
        return np.stack([qq * x, qq * y, qq])
    if k == 2:
        return np.stack([qq * x, qq, -qq * y])
    if k == 3:
        return np.stack([qq, -qq * x, -qq * y])
    if k == 4:
        return np.stack([-qq * y, -qq * x, -qq])
    return np.stack([-qq, -qq * x, qq * y])


def DX(g, tree, q, dq, rst):
    """dxyz_drst[i][j] = d x_i / d rst_j, shape (3, 3, n), by complex step"""
    h = 1e-30
    out = np.zeros((3, 3, rst.shape[1]))
    for j in range(3):
        z = rst.astype(complex)
        z[j] += 1j * h
        out[:, j, :] = X(g, tree, q, dq, z).imag / h
    return out


def jac_and_inverse(dx):
    """J (n,), cof[i][j] = J d rst_i / d xyz_j (3,3,n); d4est_geometry.c:877-1030"""
    xr, xs, xt = dx[0]
    yr, ys, yt = dx[1]
    zr, zs, zt = dx[2]
    J = xr * (ys * zt - zs * yt) - yr * (xs * zt - zs * xt) + zr * (xs * yt - ys * xt)
    cof = np.array([[ys * zt - zs * yt, -(xs * zt - zs * xt), xs * yt - ys * xt],
                    [-(yr * zt - zr * yt), xr * zt - zr * xt, -(xr * yt - yr * xt)],
                    [yr * zs - zr * ys, -(xr * zs - zr * xs), xr * ys - yr * xs]])
    return J, cof


def tensor3(x1):
    n = x1.size
    k = np.arange(n ** 3)
    return np.stack([x1[k % n], x1[(k // n) % n], x1[k // (n * n)]])


def face_points(f, x1):
    """rst (3, n^2) of the tensor points x1 (x) x1 on face f, lower tangential direction fastest
    (d4est_geometry_get_face_info, d4est_geometry.c:1378-1415)"""
    n = x1.size
    k = np.arange(n * n)
    c = f >> 1
    a = 1 if c == 0 else 0
    b = 1 if c == 2 else 2
    r = np.zeros((3, n * n))
    r[a], r[b], r[c] = x1[k % n], x1[k // n], (1.0 if f & 1 else -1.0)
    return r


def perm_index(n, perm, a, b):
    row, col = (a, b) if perm & 4 else (b, a)
    if perm & 2:
        row = n - 1 - row
    if perm & 1:
        col = n - 1 - col
    return row * n + col


def mortar_elem(q, dq, faces_side, faces_mortar, f, c):
    """(q, dq) of the element whose face carries mortar sub-face c
    (d4est_mortars_compute_qcoords_on_mortar, d4est_mortars.c:419-470)"""
    q = list(q)
    if faces_mortar == 1:
        return q, dq
    nd = f >> 1
    ta = 1 if nd == 0 else 0
    tb = 1 if nd == 2 else 2
    hanging = faces_side != faces_mortar
    step = dq // 2 if hanging else dq
    if hanging:
        if f & 1:
            q[nd] += dq // 2
        dq = dq // 2
    q[ta] += (c & 1) * step
    q[tb] += ((c >> 1) & 1) * step
    return q, dq


def volume_factors(A, g):
    LQ = int(np.asarray(A["local_nodes_quad"]).reshape(-1)[0])
    J = np.zeros(LQ)
    rst_xyz = np.zeros(9 * LQ)
    for e in range(int(np.asarray(A["n_local"]).reshape(-1)[0])):
        nq = int(A["deg_quad"][e]) + 1
        pts = tensor3(gauss(nq)[0])
        dx = DX(g, int(g["elem_tree"][e]), g["elem_q"][3 * e:3 * e + 3].astype(float), float(g["elem_dq"][e]), pts)
        Je, cof = jac_and_inverse(dx)
        qs = int(A["quad_stride"][e])
        J[qs:qs + nq ** 3] = Je
        for i in range(3):
            for j in range(3):
                rst_xyz[(i * 3 + j) * LQ + qs:(i * 3 + j) * LQ + qs + nq ** 3] = cof[i, j] / Je
    return J, rst_xyz


def nodal_xyz(A, g):
    LN = int(np.asarray(A["local_nodes"]).reshape(-1)[0])
    xyz = np.zeros(3 * LN)
    for e in range(int(np.asarray(A["n_local"]).reshape(-1)[0])):
        n = int(A["deg"][e]) + 1
        pts = tensor3(lobatto(n)[0])
        x = X(g, int(g["elem_tree"][e]), g["elem_q"][3 * e:3 * e + 3].astype(float), float(g["elem_dq"][e]), pts)
        ns = int(A["nodal_stride"][e])
        for d in range(3):
            xyz[d * LN + ns:d * LN + ns + n ** 3] = x[d]
    return xyz


def element_sizes(A, g):
    """volume [E], and per face area / min / mean / max of J/sj on the LGL nodes
    (d4est_mesh.c:1620-1830); local elements only"""
    E = int(np.asarray(A["n_local"]).reshape(-1)[0])
    vol, area = np.zeros(E), np.zeros((E, 6))
    jmin, jmean, jmax = np.zeros((E, 6)), np.zeros((E, 6)), np.zeros((E, 6))
    for e in range(E):
        n = int(A["deg"][e]) + 1
        x1, w1 = lobatto(n)
        tree, q, dq = int(g["elem_tree"][e]), g["elem_q"][3 * e:3 * e + 3].astype(float), float(g["elem_dq"][e])
        w3 = tensor3(w1)
        Jv, _ = jac_and_inverse(DX(g, tree, q, dq, tensor3(x1)))
        vol[e] = np.sum(w3[0] * w3[1] * w3[2] * Jv)
        k = np.arange(n * n)
        for f in range(6):
            Jf, cof = jac_and_inverse(DX(g, tree, q, dq, face_points(f, x1)))
            sj = np.sqrt(np.sum(cof[f >> 1] ** 2, axis=0))
            area[e, f] = np.sum(w1[k % n] * w1[k // n] * sj)
            r = Jf / sj
            # d4est_util_mean_dbl_array (src/Utilities/d4est_util.c:678-685) starts its sum at index 1
            jmin[e, f], jmean[e, f], jmax[e, f] = r.min(), r[1:].sum() / r.size, r.max()
    return vol, area, jmin, jmean, jmax


def mortar_factors(A, g, face_h_type=None):
    """sj, n, drst_dxyz_m, drst_dxyz_p_porder, hm, hp in the reference's layouts (one rank)"""
    fh = int(g["face_h_type"]) if face_h_type is None else face_h_type
    MP = int(np.asarray(A["local_mortar_nodes_quad"]).reshape(-1)[0])
    sj_o, n_o = np.zeros(MP), np.zeros(3 * MP)
    drm, drp, hm, hp = np.zeros(9 * MP), np.zeros(9 * MP), np.zeros(MP), np.zeros(MP)
    sizes = element_sizes(A, g) if fh in (1, 2, 3, 5) else None

    def pos(e):
        return int(g["elem_tree"][e]), g["elem_q"][3 * e:3 * e + 3].astype(np.int64), int(g["elem_dq"][e])

    def h_of(e, f, jdsj, dq0):
        if fh == 0:
            return jdsj
        if fh == 4:
            return np.full_like(jdsj, dq0 / ROOT)
        if fh == 5:
            return np.full_like(jdsj, sizes[0][e] / sizes[1][e, f])
        return np.full_like(jdsj, sizes[fh + 1][e, f])

    for s in range(int(np.asarray(A["n_sides"]).reshape(-1)[0])):
        fm_n, fp_n = int(A["side_faces_m"][s]), int(A["side_faces_p"][s])
        f_m, f_p, perm = int(A["side_f_m"][s]), int(A["side_f_p"][s]), int(A["side_perm"][s])
        M = max(fm_n, fp_n)
        e_m, e_p, sub = A["side_e_m"][4 * s:4 * s + 4], A["side_e_p"][4 * s:4 * s + 4], A["side_sub_order"][4 * s:4 * s + 4]
        S = int(A["side_mortar_stride"][s])
        em = [int(e_m[i if fm_n == 4 else 0]) for i in range(M)]
        ep = [int(e_p[sub[i] if fp_n == 4 else 0]) if fp_n else 0 for i in range(M)]
        nq = []
        for i in range(M):
            dq_ = int(A["deg_quad"][em[i]])
            if fp_n:
                dq_ = max(dq_, int(A["deg_quad"][ep[i]]))
            nq.append(dq_ + 1)
        off = np.concatenate([[0], np.cumsum([n * n for n in nq])]).astype(int)
        T = int(off[-1])
        nqp = [0] * M
        for i in range(M):
            nqp[int(sub[i]) if M == 4 else i] = nq[i] ** 2
        offp = np.concatenate([[0], np.cumsum(nqp)]).astype(int)
        tm, qm, dqm = pos(int(e_m[0]))
        if fp_n:
            tp, qp, dqp = pos(int(e_p[0]))
        sg = 1.0 if f_m & 1 else -1.0
        for i in range(M):
            n1 = nq[i]
            x1 = gauss(n1)[0]
            qi, dqi = mortar_elem(qm, dqm, fm_n, M, f_m, i)
            Jm, cof = jac_and_inverse(DX(g, tm, np.array(qi, float), float(dqi), face_points(f_m, x1)))
            nsj = sg * cof[f_m >> 1]
            sj = np.sqrt(np.sum(nsj ** 2, axis=0))
            sl = slice(S + off[i], S + off[i] + n1 * n1)
            sj_o[sl] = sj
            for d in range(3):
                n_o[3 * S + d * T + off[i]:3 * S + d * T + off[i] + n1 * n1] = nsj[d] / sj
            for a in range(3):
                for l in range(3):
                    b0 = 9 * S + (l + 3 * a) * T + off[i]
                    drm[b0:b0 + n1 * n1] = cof[l, a] / Jm
            hm[sl] = h_of(em[i], f_m, Jm / sj, dqm)
            if fp_n:
                j = int(sub[i]) if M == 4 else 0
                qj, dqj = mortar_elem(qp, dqp, fp_n, M, f_p, j)
                Jp, cofp = jac_and_inverse(DX(g, tp, np.array(qj, float), float(dqj), face_points(f_p, x1)))
                sjp = np.sqrt(np.sum(cofp[f_p >> 1] ** 2, axis=0))
                for a in range(3):
                    for l in range(3):
                        b0 = 9 * S + (l + 3 * a) * T + offp[j]
                        drp[b0:b0 + n1 * n1] = cofp[l, a] / Jp
                k = np.arange(n1 * n1)
                kp = np.array([perm_index(n1, perm, int(kk % n1), int(kk // n1)) for kk in k])
                hp[sl] = h_of(ep[i], f_p, (Jp / sjp)[kp], dqp)
    return {"sj_m_mortar_quad": sj_o, "n_m_mortar_quad": n_o, "drst_dxyz_m_mortar_quad": drm,
            "drst_dxyz_p_mortar_quad_porder": drp, "hm_mortar_quad": hm, "hp_mortar_quad": hp}


def geom_from_sidecar(G, face_h_type=None):
    """geometry dict (abi.make_geom_desc keys) from a *.geom.d4b sidecar written by ref_driver --geom"""
    p = G["geom_params"]
    g = {"geom_type": int(G["geom_type"][0]), "face_h_type": int(G["face_h_type"][0]) if face_h_type is None else face_h_type,
         "elem_tree": G["elem_tree"], "elem_q": G["elem_q"], "elem_dq": G["elem_dq"],
         "compactify_inner_shell": 0, "compactify_outer_shell": 0}
    if g["geom_type"] == 0:
        g.update(X0=p[0], X1=p[1], Y0=p[2], Y1=p[3], Z0=p[4], Z1=p[5])
    else:
        g.update(R0=p[0], R1=p[1], R2=p[2], compactify_inner_shell=int(p[3]), compactify_outer_shell=int(p[4]))
    return g
