"""GPU parity tests proper: the CUDA path (through the C ABI) against the
fixtures written by the unmodified reference and against the CPU oracle.
Tolerance: relative L2 <= 1e-12 (BASELINE.json north_star)."""
import numpy as np
import pytest

from conftest import GOLDEN_CASES

pytestmark = pytest.mark.gpu
TOL = 1e-12


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.fixture(scope="module")
def meshes():
    from disco4est_b200 import Mesh

    cache = {}

    def get(name, A):
        if name not in cache:
            cache[name] = Mesh(A)
        return cache[name]

    yield get
    for m in cache.values():
        m.close()


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_apply_matches_reference(golden, meshes, case):
    A = golden(case)
    m = meshes(case, A)
    u = np.ascontiguousarray(A["u"])
    Au = np.full_like(u, np.nan)
    m.apply(u, Au)
    assert rel(Au, A["Au_plain"]) < TOL
    assert rel(Au, A["Au_with_opt"]) < TOL


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_apply_matches_oracle_random_vector(golden, meshes, case):
    import oracle_lib

    A = golden(case)
    m = meshes(case, A)
    rng = np.random.default_rng(7)
    u = rng.standard_normal(m.n)
    Au = np.zeros_like(u)
    m.apply(u, Au)
    assert rel(Au, oracle_lib.Oracle(A).apply(u)) < TOL


def test_stiffness_only(golden, meshes):
    A = golden("brick_l2_p3")
    m = meshes("brick_l2_p3", A)
    u = np.ascontiguousarray(A["u"])
    Ku = np.zeros_like(u)
    m.apply_stiffness(u, Ku)
    assert rel(Ku, A["Ku"]) < TOL


def test_device_pointers_and_linearity(golden, meshes):
    import torch

    A = golden("sphere_hp_amr1")
    m = meshes("sphere_hp_amr1", A)
    A0 = dict(A)
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(3)
    x = torch.randn(m.n, dtype=torch.float64, generator=g)
    y = torch.randn(m.n, dtype=torch.float64, generator=g)
    xd, yd = x.to(dev), y.to(dev)
    Ax, Ay, Axy = (torch.empty(m.n, dtype=torch.float64, device=dev) for _ in range(3))
    torch.cuda.synchronize()
    # inhomogeneous Dirichlet data makes the map affine: A(x) - A(0) is linear
    z = torch.zeros_like(xd)
    A0v = torch.empty_like(xd)
    m.apply(z, A0v); m.apply(xd, Ax); m.apply(yd, Ay)
    s = (2.0 * xd - 3.0 * yd).contiguous()
    m.apply(s, Axy)
    m.sync()
    lin = 2.0 * (Ax - A0v) - 3.0 * (Ay - A0v)
    assert rel((Axy - A0v).cpu().numpy(), lin.cpu().numpy()) < 1e-11
    # host path == device path bit for bit
    Ah = np.zeros(m.n)
    m.apply(x.numpy(), Ah)
    assert np.array_equal(Ah, Ax.cpu().numpy())


@pytest.mark.parametrize("case", ["brick_l2_p1", "brick_l2_p3", "brick_l2_p7", "sphere_p2"])
def test_uniform_meshes_take_the_fast_kernel(golden, meshes, case):
    A = golden(case)
    m = meshes(case, A)
    assert m.fast_elements() == int(A["n_local"][0])


def test_hp_mortar_kernel_takes_the_collocated_hp_mesh(golden, monkeypatch):
    """random hp refinement with deg_quad = deg (hanging faces both ways, unequal degrees, all
    orientations): every element runs in fast_apply_kernel or hp_apply_kernel, the general kernel
    gets nothing, and the result equals the general kernel's and the reference's"""
    from disco4est_b200 import Mesh

    A = golden("sphere_p_amr1")
    u = np.ascontiguousarray(A["u"])
    m1 = Mesh(A)
    assert m1.hp_elements() > 0 and m1.fast_elements() == int(A["n_local"][0])
    monkeypatch.setenv("D4B_NO_HP", "1")
    m2 = Mesh(A)
    monkeypatch.delenv("D4B_NO_HP")
    assert m2.hp_elements() == 0 and m2.fast_elements() < m1.fast_elements()
    a1, a2 = np.zeros_like(u), np.zeros_like(u)
    m1.apply(u, a1); m2.apply(u, a2)
    m1.close(); m2.close()
    assert rel(a1, a2) < 1e-13
    assert rel(a1, A["Au_plain"]) < TOL


@pytest.mark.parametrize("case", ["brick_l2_p3", "sphere_p2", "sphere_hp_amr1", "sphere_p_amr1"])
def test_fast_and_general_kernels_agree(golden, case, monkeypatch):
    from disco4est_b200 import Mesh

    A = golden(case)
    u = np.ascontiguousarray(A["u"])
    m1 = Mesh(A)
    monkeypatch.setenv("D4B_FORCE_GENERIC", "1")
    m2 = Mesh(A)
    monkeypatch.delenv("D4B_FORCE_GENERIC")
    assert m2.fast_elements() == 0
    a1, a2 = np.zeros_like(u), np.zeros_like(u)
    m1.apply(u, a1); m2.apply(u, a2)
    assert rel(a1, a2) < 1e-13
    m1.close(); m2.close()


@pytest.mark.parametrize("case", ["brick_l2_p1", "brick_l2_p2_bc", "brick_l2_p3", "brick_l2_p5", "brick_l2_p7"])
def test_brick_takes_the_tangential_free_kernel(golden, meshes, case):
    """Axis-aligned faces have no tangential metric terms: every element of
    the reference's brick fixtures must run in flat_apply_kernel."""
    A = golden(case)
    m = meshes(case, A)
    assert m.flat_elements() == int(A["n_local"][0])


def test_curved_mesh_does_not_take_the_tangential_free_kernel(golden, meshes):
    A = golden("sphere_p2")
    m = meshes("sphere_p2", A)
    assert m.flat_elements() < int(A["n_local"][0])


@pytest.mark.parametrize("case", ["brick_l2_p1", "brick_l2_p2_bc", "brick_l2_p3", "brick_l2_p5", "brick_l2_p7"])
def test_flat_and_general_fast_kernels_agree(golden, case, monkeypatch):
    from disco4est_b200 import Mesh

    A = golden(case)
    u = np.ascontiguousarray(A["u"])
    m1 = Mesh(A)
    monkeypatch.setenv("D4B_NO_FLAT", "1")
    m2 = Mesh(A)
    monkeypatch.delenv("D4B_NO_FLAT")
    assert m2.flat_elements() == 0 and m2.fast_elements() == m1.fast_elements()
    a1, a2 = np.zeros_like(u), np.zeros_like(u)
    m1.apply(u, a1); m2.apply(u, a2)
    assert rel(a1, a2) < 1e-13
    assert rel(a1, A["Au_with_opt"]) < 1e-12
    m1.close(); m2.close()


@pytest.mark.parametrize("p", [1, 2, 3, 4, 5, 6, 7])
def test_flat_kernel_all_degrees_against_oracle(p):
    """Synthetic 4^3 brick (disco4est_b200.brick) at every supported degree."""
    from disco4est_b200 import Mesh
    from disco4est_b200.brick import brick_desc
    from oracle_lib import Oracle

    A = brick_desc(2, p)
    rng = np.random.default_rng(p)
    u = rng.standard_normal(A["local_nodes"])
    m = Mesh(A)
    assert m.flat_elements() == A["n_local"]
    au = np.zeros_like(u)
    m.apply(u, au)
    ref = Oracle(A).apply(u)
    assert rel(au, ref) < 1e-12
    m.close()


@pytest.mark.parametrize("case", ["brick_l2_p3", "sphere_p2", "sphere_p_amr1", "sphere_p_amr1_robin"])
def test_pipelined_host_apply_equals_device_apply(golden, case, monkeypatch):
    """host-pointer applies on one-rank meshes go up / down in chunks overlapped with the kernels
    (apply_host_pipelined); forced onto the small fixtures here: bit-identical to the device-pointer apply"""
    import torch

    from disco4est_b200 import Mesh

    A = golden(case)
    monkeypatch.setenv("D4B_PIPELINE_MIN_ELEMS", "1")
    m = Mesh(A)
    monkeypatch.delenv("D4B_PIPELINE_MIN_ELEMS")
    u = np.ascontiguousarray(A["u"])
    a_host = np.full_like(u, np.nan)
    m.apply(u, a_host)
    m.apply(u, a_host)                       # events and staging are re-used
    du = torch.from_numpy(u).cuda()
    dAu = torch.empty_like(du)
    m.apply(du, dAu)
    m.sync()
    assert np.array_equal(a_host, dAu.cpu().numpy())
    assert rel(a_host, A["Au_plain"]) < TOL
    m.close()


def test_pipelined_host_apply_at_size():
    """16^3 p=2 brick (above the default threshold): pinned and pageable host buffers"""
    import torch

    from disco4est_b200 import Mesh
    from disco4est_b200.brick import brick_desc

    m = Mesh(brick_desc(4, 2))
    u = np.random.default_rng(1).standard_normal(m.n)
    du = torch.from_numpy(u).cuda()
    dAu = torch.empty_like(du)
    m.apply(du, dAu)
    m.sync()
    ref = dAu.cpu().numpy()
    a1 = np.zeros_like(u)
    m.apply(u, a1)
    pu, pa = torch.from_numpy(u).pin_memory(), torch.empty(m.n, dtype=torch.float64).pin_memory()
    m.apply(pu.numpy(), pa.numpy())
    assert np.array_equal(a1, ref) and np.array_equal(pa.numpy(), ref)
    # pageable vectors page-locked by the library (what d4est_b200_pin_vectors does for the reference's vectors)
    u2, a2 = u.copy(), np.zeros_like(u)
    assert m.L.d4b_host_pin(u2.ctypes.data, u2.nbytes) == 0 and m.L.d4b_host_pin(a2.ctypes.data, a2.nbytes) == 0
    m.apply(u2, a2)
    assert np.array_equal(a2, ref)
    assert m.L.d4b_host_unpin(u2.ctypes.data) == 0 and m.L.d4b_host_unpin(a2.ctypes.data) == 0
    assert m.L.d4b_host_unpin(a2.ctypes.data) != 0   # (not pinned any more: an error code, no abort)
    m.close()


@pytest.mark.parametrize("case", ["sphere_p_amr1", "sphere_p_amr1_robin"])
def test_face_split_equals_mortar_kernel(golden, case, monkeypatch):
    """face split: irregular elements run in the conforming kernels with their irregular faces supplied by a
    FACE_ONLY launch of the mortar kernel; same operator as the element-wise mortar kernel and the reference"""
    from disco4est_b200 import Mesh

    A = golden(case)
    u = np.ascontiguousarray(A["u"])
    monkeypatch.setenv("D4B_SPLIT", "1")
    m1 = Mesh(A)
    monkeypatch.setenv("D4B_SPLIT", "0")
    m2 = Mesh(A)
    monkeypatch.delenv("D4B_SPLIT")
    assert m1.hp_elements() == m2.hp_elements() > 0
    a1, a2 = np.zeros_like(u), np.zeros_like(u)
    m1.apply(u, a1); m2.apply(u, a2)
    m1.close(); m2.close()
    assert rel(a1, a2) < 1e-13
    assert rel(a1, A["Au_plain"]) < TOL


@pytest.mark.parametrize("kw", [dict(pmax=2), dict(pmax=4), dict(p=3, graded=True)])
def test_face_split_on_synthetic_spheres(kw, monkeypatch):
    from disco4est_b200 import Mesh
    from disco4est_b200.sphere import sphere_desc

    B, g = sphere_desc(3, **kw)
    u = np.random.default_rng(2).standard_normal(int(B["local_nodes"]))
    out = []
    for flag in ("1", "0"):
        monkeypatch.setenv("D4B_SPLIT", flag)
        m = Mesh(B, geom=g)
        a = np.zeros_like(u)
        m.apply(u, a)
        m.close()
        out.append(a)
    monkeypatch.delenv("D4B_SPLIT")
    assert rel(out[0], out[1]) < 1e-13


@pytest.mark.parametrize("case", ["brick_l2_p1", "brick_l2_p3", "brick_l2_p5", "brick_l2_p7", "brick_l2_p2_bc"])
def test_affine_path_matches_per_point_metric(golden, case, monkeypatch):
    """bricks: every element is affine, the tangential-free kernel then reads the volume metric once per element
    (flat_apply_kernel<N, MASS, CG = true>); D4B_AFFINE=0 keeps the per-point metric.  Same operator."""
    from disco4est_b200 import Mesh

    A = golden(case)
    u = np.ascontiguousarray(A["u"])
    out = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("D4B_AFFINE", flag)
        m = Mesh(A)
        assert m.flat_elements() == int(A["n_local"][0])
        out[flag] = np.zeros_like(u)
        m.apply(u, out[flag])
        m.close()
    assert rel(out["1"], A["Au_plain"]) < TOL
    assert rel(out["0"], A["Au_plain"]) < TOL
    assert rel(out["1"], out["0"]) < 1e-13


@pytest.mark.parametrize("level,p", [(2, 1), (2, 3), (1, 4), (1, 7)])
def test_device_generated_and_folded_brick_matches_host_arrays(level, p):
    """brick_desc(geometry=False): topology only, factors generated AND folded on the device
    (d4b_mesh_create_with_geometry -> fold_volume_kernel / fold_mortar_kernel, face records and class metrics
    gathered by kernels) against the same mesh described with host arrays (host fold)"""
    from disco4est_b200 import Mesh
    from disco4est_b200.brick import brick_desc

    A = brick_desc(level, p)
    T, g = brick_desc(level, p, geometry=False)
    ma, mt = Mesh(A), Mesh(T, geom=g)
    assert mt.fast_elements() == ma.fast_elements() == int(A["n_local"])
    assert mt.flat_elements() == ma.flat_elements()
    u = np.random.default_rng(5).standard_normal(ma.n)
    a, b = np.zeros_like(u), np.zeros_like(u)
    ma.apply(u, a)
    mt.apply(u, b)
    ma.close()
    mt.close()
    assert rel(b, a) < 1e-13
