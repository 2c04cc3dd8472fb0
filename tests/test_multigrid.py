"""V-cycle / inter-grid transfer against the unmodified reference's multigrid
(tests/golden/brick_mg_l2_p2: d4est_solver_multigrid_solve called 4 times with
vcycle_imax = 1 by oracle/ref_driver.c --mg; residual and iterate after each)."""
import numpy as np
import pytest

from oracle_lib import Oracle
from solver_ref import Transfer, prolong_1d, vcycle

CASE = "brick_mg_l2_p2"
# same mesh and right-hand side with mg_bottom_solver_cheby (cheby_imax 15, cheby_eigs_cg_imax 15, ratio 30): the
# configuration of the bench V-cycle and of the shipped nonlinear problems' input files
CASES = [("brick_mg_l2_p2", None), ("brick_mg_l2_p2_cheby", (15, 15, 30.0))]


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def hierarchy(A):
    from disco4est_b200.brick import brick_hierarchy

    p = int(A["deg"][0])
    descs, transfers = brick_hierarchy(2, p)
    return descs, transfers


def test_synthetic_fine_level_is_the_reference_mesh(golden):
    """the hierarchy's top level must be the mesh the reference solved on"""
    A = golden(CASE)
    descs, _ = hierarchy(A)
    u = np.ascontiguousarray(A["u"])
    assert rel(Oracle(descs[-1]).apply(u), A["Au_plain"]) < 1e-13


def test_prolong_tables_match_reference(golden):
    A = golden("brick_l2_p5")
    for p in range(1, 6):
        for q in range(p, 6):
            ref = np.asarray(A["tab_p_prolong_%d_%d" % (p, q)]).reshape(q + 1, p + 1)
            assert np.abs(prolong_1d(p, q, 0) - ref).max() < 1e-13


@pytest.mark.parametrize("case,bottom_cheby", CASES)
def test_restatement_matches_reference_multigrid(golden, case, bottom_cheby):
    """pins tests/solver_ref.py:vcycle (and the synthetic hierarchy) to the reference"""
    A = golden(case)
    descs, transfers = hierarchy(A)
    assert len(descs) == int(A["mg_num_levels"][0])
    orcs = [Oracle(d) for d in descs]
    trs = [Transfer(descs[l], descs[l + 1], *transfers[l]) for l in range(len(descs) - 1)]
    imax, eigs_imax, ratio, mult, use_new = (float(x) for x in A["mg_cheby_params"])
    rhs = np.ascontiguousarray(A["mg_rhs"])
    n = rhs.size
    u = np.zeros(n)
    r2_ref = np.asarray(A["mg_r2"])
    assert abs(float(rhs @ rhs) - r2_ref[0]) <= 1e-12 * r2_ref[0]
    hist = np.asarray(A["mg_u_hist"]).reshape(-1, n)
    for k in range(hist.shape[0]):
        u, r2, eigs = vcycle([o.apply for o in orcs], trs, u, rhs, int(imax), int(eigs_imax), ratio, mult, int(use_new),
                             bottom_cheby=bottom_cheby)
        # relative to the initial residual: the last cycles sit at round-off level
        assert abs(r2 - r2_ref[k + 1]) <= 1e-6 * r2_ref[k + 1] + 1e-25 * r2_ref[0]
        assert rel(u, hist[k]) < 1e-9
    assert np.allclose(eigs[1:], A["mg_eigs_last"][1:], rtol=1e-5)   # level 0 has no smoother (bottom CG)


@pytest.fixture()
def gpu_mg(golden):
    from disco4est_b200 import Mesh
    from disco4est_b200.runtime import Multigrid, cheby_params

    A = golden(CASE)
    descs, transfers = hierarchy(A)
    meshes = [Mesh(d) for d in descs]
    imax, eigs_imax, ratio, mult, use_new = (float(x) for x in A["mg_cheby_params"])

    def make(**kw):
        return Multigrid(meshes, transfers, cheby_params(int(imax), int(eigs_imax), ratio, mult, int(use_new)), **kw)

    yield A, descs, transfers, make
    for m in meshes:
        m.close()


@pytest.mark.gpu
def test_gpu_transfers_match_restatement(gpu_mg):
    A, descs, transfers, make = gpu_mg
    mg = make()
    rng = np.random.default_rng(3)
    for l in range(len(descs) - 1):
        tr = Transfer(descs[l], descs[l + 1], *transfers[l])
        fine = rng.standard_normal(descs[l + 1]["local_nodes"])
        coarse = np.zeros(descs[l]["local_nodes"])
        mg.restrict(l, fine, coarse)
        assert rel(coarse, tr.restrict(fine)) < 1e-13
        c = rng.standard_normal(descs[l]["local_nodes"])
        f = np.zeros(descs[l + 1]["local_nodes"])
        mg.prolong(l, c, f)
        assert rel(f, tr.prolong(c)) < 1e-13
        # restriction is the transpose of prolongation
        assert abs(float(fine @ f) - float(coarse @ c)) <= 1e-12 * abs(float(fine @ f))
    mg.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case,bottom_cheby", CASES)
def test_gpu_vcycles_match_reference(gpu_mg, golden, case, bottom_cheby):
    from disco4est_b200.runtime import cheby_params

    _, descs, transfers, make = gpu_mg
    A = golden(case)
    mg = make() if bottom_cheby is None else make(bottom_solver="cheby",
                                                  bottom_cheby=cheby_params(bottom_cheby[0], bottom_cheby[1], bottom_cheby[2]))
    rhs = np.ascontiguousarray(A["mg_rhs"])
    n = rhs.size
    u = np.zeros(n)
    r2_ref = np.asarray(A["mg_r2"])
    hist = np.asarray(A["mg_u_hist"]).reshape(-1, n)
    for k in range(hist.shape[0]):
        r2 = mg.vcycle(u, rhs)
        assert abs(r2 - r2_ref[k + 1]) <= 1e-6 * r2_ref[k + 1] + 1e-25 * r2_ref[0]
        assert rel(u, hist[k]) < 1e-9
    assert np.allclose(mg.eigs()[1:], A["mg_eigs_last"][1:], rtol=1e-5)   # 4th cycle: residual at round-off
    mg.close()


@pytest.mark.gpu
def test_gpu_solve_stops_like_the_reference(gpu_mg):
    """d4est_solver_multigrid_solve: same number of V-cycles to rtol as the restatement
    built on the oracle, same residual history"""
    A, descs, transfers, make = gpu_mg
    rhs = np.ascontiguousarray(A["mg_rhs"])
    r2_ref = np.asarray(A["mg_r2"])
    rtol = 1e-6
    want = next(k for k in range(len(r2_ref)) if r2_ref[k] <= rtol * rtol * r2_ref[0])
    mg = make(vcycle_imax=10, vcycle_rtol=rtol)
    u = np.zeros_like(rhs)
    n, hist = mg.solve(u, rhs)
    assert n == want
    assert np.allclose(hist, r2_ref[: n + 1], rtol=1e-6)
    # device-pointer call gives the same iterate
    import torch

    du = torch.zeros(rhs.size, dtype=torch.float64, device="cuda")
    n2, hist2 = mg.solve(du, torch.from_numpy(rhs).cuda())
    assert n2 == n and np.array_equal(du.cpu().numpy(), u)
    mg.close()


@pytest.mark.gpu
def test_graph_replay_equals_eager_cycles(monkeypatch):
    """the recorded V-cycle (CUDA graph, Chebyshev bottom solver) reproduces the eagerly enqueued one
    bit for bit, and a graph recorded before d4b_mesh_set_mass_diag is not replayed after it"""
    import torch

    from disco4est_b200 import Mesh
    from disco4est_b200.brick import brick_hierarchy
    from disco4est_b200.runtime import Multigrid, cheby_params

    descs, transfers = brick_hierarchy(3, 2)
    dev = torch.device("cuda:0")

    def run(graph):
        monkeypatch.setenv("D4B_MG_GRAPH", "1" if graph else "0")
        meshes = [Mesh(d) for d in descs]
        mg = Multigrid(meshes, transfers, cheby_params(6, 6), vcycle_imax=1, bottom_solver="cheby",
                       bottom_cheby=cheby_params(6, 8))
        n = meshes[-1].n
        rhs = torch.rand(n, dtype=torch.float64, device=dev, generator=torch.Generator(device=dev).manual_seed(3))
        u = torch.zeros_like(rhs)
        r2 = [mg.vcycle(u, rhs) for _ in range(4)]       # eager, capture, replay, replay
        jac = torch.full((int(descs[-1]["local_nodes_quad"]),), 0.7, dtype=torch.float64, device=dev)
        meshes[-1].set_mass_diag(jac)                    # the operator changes: A + 0.7 M-like term
        r2 += [mg.vcycle(u, rhs) for _ in range(3)]
        # new values in the same array (a Newton solve re-linearises every step): the recorded graphs stay valid and
        # must see them
        jac2 = torch.full((int(descs[-1]["local_nodes_quad"]),), 0.2, dtype=torch.float64, device=dev)
        replays = mg.graph_replays()
        meshes[-1].set_mass_diag(jac2)
        r2 += [mg.vcycle(u, rhs) for _ in range(2)]
        assert (mg.graph_replays() - replays == 2) == graph
        out = u.cpu().numpy().copy()
        mg.close()
        for m in meshes:
            m.close()
        return r2, out

    r2_g, u_g = run(True)
    r2_e, u_e = run(False)
    assert r2_g == r2_e
    assert np.array_equal(u_g, u_e)
    assert r2_g[4] != r2_g[3]
