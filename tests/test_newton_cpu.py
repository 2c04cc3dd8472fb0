"""CPU restatement of BASELINE.json configs[4] in small (two punctures, Newton -> FCG -> multigrid with the
matrix operator) on top of the pinned oracle apply, checked against the fixture the UNMODIFIED reference wrote
(tests/golden/sphere_punctures_newton, oracle/ref_driver.c --newton).  Pins, without a GPU:
  * the puncture functor f = -K^2/8 psi^-7 and its derivative (multi_punctures_fcns_with_opt.h:292-423),
  * what the reference's multigrid matrix operator EXECUTES on the coarse level -- not the Galerkin product its
    comments describe: d4est_operators_compute_PT_mat_P slices the transpose of the STACKED prolongation
    (src/dGMath/d4est_operators.c:638, :650), and the fine matrices are always linearised about u = 0
    (argument shift, src/Solver/d4est_solver_multigrid_matrix_operator.c:215-240),
  * Newton steps, FCG iterations per step, residual norms."""
import numpy as np
import pytest

import geom_ref
import oracle_lib
import solver_ref
from disco4est_b200.sphere import sphere_desc

TOL = 1e-12


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def puncture_terms(P, x, y, z):
    """K^2 = A_ij A^ij and s = sum m_n / (2 r_n) at points (x, y, z); P: [n][10] = M, X, Y, Z, P, S"""
    Aij = np.zeros((3, 3) + x.shape)
    s = np.zeros_like(x)
    for h in P:
        d = np.stack([x - h[1], y - h[2], z - h[3]])
        r = np.sqrt((d ** 2).sum(0))
        n = d / r
        Pm, S = h[4:7], h[7:10]
        sxn = np.stack([-n[2] * S[1] + n[1] * S[2], n[2] * S[0] - n[0] * S[2], -n[1] * S[0] + n[0] * S[1]])
        pdn = Pm[0] * n[0] + Pm[1] * n[1] + Pm[2] * n[2]
        for a in range(3):
            for b in range(3):
                g = 1.0 if a == b else 0.0
                Aij[a, b] += 1.5 / r ** 2 * (Pm[a] * n[b] + Pm[b] * n[a] - (g - n[a] * n[b]) * pdn) \
                    + 3.0 / r ** 3 * (sxn[a] * n[b] + sxn[b] * n[a])
        s += h[0] / (2.0 * r)
    return (Aij ** 2).sum((0, 1)), s


class Problem:
    def __init__(self, A):
        self.A = A
        LQ = int(A["local_nodes_quad"][0])
        self.K2, self.s = puncture_terms(A["punctures"].reshape(-1, 10), *A["xyz_quad"].reshape(3, LQ))
        self.O1 = oracle_lib.Oracle(A)
        self.J = A["J_quad"]
        # coarse level: one element per tree, factors from the CPU restatement of the geometry
        T0, g0 = sphere_desc(0, p=2, prefactor=2.0, radii=(1.0, 2.0, 3.0))
        g0 = dict(g0, geom_type=1, face_h_type=0)
        A0 = dict(T0)
        J0, rst0 = geom_ref.volume_factors(T0, g0)
        A0.update(J_quad=J0, rst_xyz_quad=rst0)
        A0.update(geom_ref.mortar_factors(T0, g0))
        mp = int(A0["local_mortar_nodes_quad"])
        A0.update(bc_type=1, robin_coeff_mortar_quad=np.full(mp, 1.0 / 3.0), robin_rhs_mortar_quad=np.zeros(mp))
        self.A0, self.O0 = A0, oracle_lib.Oracle(A0)
        self.tr = solver_ref.Transfer(A0, A, 8 * np.arange(13), np.full(13, 8))
        # LGL(2) -> Gauss(2) interpolation in 3-D (row = quadrature point), from the reference's own table
        V1 = A["tab_lgl_to_gauss_2_2"].reshape(3, 3) if "tab_lgl_to_gauss_2_2" in A else None
        if V1 is None:
            from disco4est_b200.brick import lgl_nodes
            xl = lgl_nodes(3)
            xg = np.polynomial.legendre.leggauss(3)[0]
            V1 = np.ones((3, 3))
            for j in range(3):
                for k in range(3):
                    if k != j:
                        V1[:, j] *= (xg - xl[k]) / (xl[j] - xl[k])
        self.V = np.kron(V1, np.kron(V1, V1))
        self.wq = np.polynomial.legendre.leggauss(3)[1]
        self.w3 = np.kron(self.wq, np.kron(self.wq, self.wq))
        self.P = [solver_ref.prolong_3d(2, 2, c, 8) for c in range(8)]

    def f(self, u):
        return -self.K2 / 8.0 / (1.0 + self.O1.interp_to_quad(u) + self.s) ** 7

    def fu(self, u0):
        return 7.0 / 8.0 * self.K2 / (1.0 + self.O1.interp_to_quad(u0) + self.s) ** 8

    def residual(self, u):
        return self.O1.apply(u) + self.O1.apply_galerkin(self.f(u), self.J)

    def coarse_mass_matrices(self, W, executed):
        """dense 27 x 27 mass-type matrix of every coarse element from the fine weights W = J f_u:
        executed = True: sum_i Q_i M_i P_i with Q_i the slice the reference takes; False: sum_i P_i^T M_i P_i"""
        n = H = 27
        Pstack = np.vstack(self.P)
        PTflat = np.ascontiguousarray(Pstack.T).reshape(-1)
        out = []
        for e in range(13):
            M = np.zeros((H, H))
            for i in range(8):
                qs = int(self.A["quad_stride"][8 * e + i])
                Mi = self.V.T @ ((self.w3 * W[qs:qs + 27])[:, None] * self.V)
                Qi = PTflat[i * n * H:(i + 1) * n * H].reshape(H, n) if executed else self.P[i].T
                M += Qi @ (Mi @ self.P[i])
            out.append(M)
        return out

    def linearize(self, u0, executed=True):
        W = self.J * self.fu(u0)
        # executed: the reference evaluates f_u at u = 0 for its matrix operator whatever u0 is (argument shift in
        # d4est_solver_multigrid_matrix_setup_fofufofvlilj_operator, d4est_solver_multigrid_matrix_operator.c:215-240)
        Ms = self.coarse_mass_matrices(self.J * self.fu(np.zeros_like(u0)) if executed else W, executed)
        ap1 = lambda x: self.O1.apply(x) + self.O1.apply_mass(x, W)

        def ap0(x):
            y = self.O0.apply(x)
            for e, M in enumerate(Ms):
                y[27 * e:27 * e + 27] += M @ x[27 * e:27 * e + 27]
            return y

        ch = self.A["mg_cheby_params"]
        pc = lambda r: solver_ref.vcycle2(ap1, ap0, self.tr, r, int(ch[0]), int(ch[1]), ch[2], int(ch[4]))
        return ap1, pc


@pytest.fixture(scope="module")
def prob(golden):
    return Problem(golden("sphere_punctures_newton"))


def test_puncture_functor_restatement(prob):
    A = prob.A
    assert rel(prob.residual(A["nw_us"]), A["nw_F_us"]) < TOL
    Jv = prob.O1.apply(A["nw_v"]) + prob.O1.apply_mass(A["nw_v"], prob.J * prob.fu(A["nw_us"]))
    assert rel(Jv, A["nw_J_us_v"]) < TOL


def test_first_preconditioner_application(prob):
    """the V-cycle with the coarse mass term as the reference EXECUTES it reproduces its output; the Galerkin
    product P^T M P its comments describe does not"""
    A = prob.A
    _, pc = prob.linearize(np.zeros(prob.O1.n), executed=True)
    assert rel(pc(A["nw_pc_x0"]), A["nw_pc_y0"]) < 1e-12
    _, pc = prob.linearize(np.zeros(prob.O1.n), executed=False)
    assert rel(pc(A["nw_pc_x0"]), A["nw_pc_y0"]) > 1e-3


def test_newton_iteration_counts(prob):
    A = prob.A
    p = A["nw_params"]
    u, its, norms = solver_ref.newton(prob.residual, prob.linearize, np.zeros(prob.O1.n), p[0], p[1], int(p[2]),
                                      int(p[3]), p[5], p[4], int(p[6]))
    assert its == A["nw_fcg_iterations"].astype(int).tolist()
    assert np.allclose(norms, A["nw_residual_norms"], rtol=1e-6, atol=1e-18)
    assert rel(u, A["nw_u"]) < 1e-9
