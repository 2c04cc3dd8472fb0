"""Host-side harness over the C ABI (include/d4est_b200.h) via ctypes.

Mirrors the reference's calling pattern for the hot path: a mesh object built
once per mesh (d4est_mesh_update), then `apply` = eqns->apply_lhs
(src/EllipticSystem/d4est_elliptic_eqns.c:3-29), `cg_eigs` and `cheby_smooth`
(src/Solver/d4est_solver_cg_eigs.c:115-276,
src/Solver/d4est_solver_multigrid_smoother_cheby.c:81-176).

There is no CPU fallback: if libd4est_b200.so is missing or no CUDA device is
visible, construction fails loudly.
"""
import ctypes as C
import os

import numpy as np

from .abi import (EST_FAMILY, GEOM_ARRAY_FIELDS, NL_KIND, ChebyParams, EstimatorDesc, GeomArrays, GeomDesc, MeshDesc, MgDesc,
                  NewtonParams, NewtonStats, NonlinearDesc, make_desc, make_geom_desc)

_HERE = os.path.dirname(os.path.abspath(__file__))
# D4B_LIBRARY_PATH lets profiling scripts A/B differently compiled builds of the same library
LIB_PATH = os.environ.get("D4B_LIBRARY_PATH") or os.path.join(_HERE, "libd4est_b200.so")
_F64P = C.POINTER(C.c_double)
_I32P_T = C.POINTER(C.c_int32)

EXPORTS = [
    "d4b_set_abort_hook", "d4b_abi_version", "d4b_device_count", "d4b_init",
    "d4b_geom_compute", "d4b_mesh_create_with_geometry",
    "d4b_mesh_create", "d4b_mesh_destroy", "d4b_mesh_local_nodes", "d4b_mesh_device_bytes", "d4b_mesh_fast_elements", "d4b_mesh_flat_elements", "d4b_mesh_hp_elements", "d4b_mesh_launches_per_apply", "d4b_mesh_halo_mode", "d4b_host_pin", "d4b_host_unpin",
    "d4b_apply", "d4b_apply_stiffness", "d4b_sync", "d4b_set_stream",
    "d4b_interp_to_quad", "d4b_apply_galerkin", "d4b_apply_mass", "d4b_mesh_set_mass_diag", "d4b_build_rhs", "d4b_estimator_bi",
    "d4b_vec_dot", "d4b_vec_axpy", "d4b_vec_xpby", "d4b_vec_scale", "d4b_vec_dot_noalloc",
    "d4b_cg_eigs", "d4b_cheby_smooth",
    "d4b_mg_create", "d4b_mg_destroy", "d4b_mg_vcycle", "d4b_mg_solve", "d4b_mg_get_eigs", "d4b_mg_graph_replays", "d4b_mg_restrict",
    "d4b_mg_prolong",
    "d4b_cg_solve", "d4b_fcg_solve",
    "d4b_mesh_set_nonlinear_term", "d4b_mesh_linearize", "d4b_build_residual", "d4b_mg_set_coarse_mass", "d4b_newton_solve",
    "d4b_comm_get_unique_id", "d4b_comm_init", "d4b_comm_destroy", "d4b_comm_size", "d4b_comm_rank",
]

_lib = None


def load_library():
    """dlopen the in-tree shared library and declare the C signatures."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libd4est_b200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
            "there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    L.d4b_abi_version.restype = C.c_int
    L.d4b_device_count.restype = C.c_int
    L.d4b_init.argtypes = [C.c_int]
    L.d4b_mesh_create.restype = C.c_void_p
    L.d4b_mesh_create.argtypes = [C.POINTER(MeshDesc)]
    L.d4b_mesh_destroy.argtypes = [C.c_void_p]
    L.d4b_geom_compute.argtypes = [C.POINTER(MeshDesc), C.POINTER(GeomDesc), C.POINTER(GeomArrays)]
    L.d4b_geom_compute.restype = None
    L.d4b_mesh_create_with_geometry.restype = C.c_void_p
    L.d4b_mesh_create_with_geometry.argtypes = [C.POINTER(MeshDesc), C.POINTER(GeomDesc)]
    L.d4b_mesh_local_nodes.restype = C.c_int64
    L.d4b_mesh_local_nodes.argtypes = [C.c_void_p]
    L.d4b_mesh_device_bytes.restype = C.c_int64
    L.d4b_mesh_device_bytes.argtypes = [C.c_void_p]
    L.d4b_mesh_fast_elements.restype = C.c_int64
    L.d4b_mesh_fast_elements.argtypes = [C.c_void_p]
    L.d4b_mesh_flat_elements.restype = C.c_int64
    L.d4b_mesh_flat_elements.argtypes = [C.c_void_p]
    L.d4b_mesh_hp_elements.restype = C.c_int64
    L.d4b_mesh_hp_elements.argtypes = [C.c_void_p]
    L.d4b_mesh_launches_per_apply.restype = C.c_int32
    L.d4b_mesh_launches_per_apply.argtypes = [C.c_void_p]
    if hasattr(L, "d4b_host_pin"):
        L.d4b_host_pin.restype = C.c_int32
        L.d4b_host_pin.argtypes = [C.c_void_p, C.c_int64]
        L.d4b_host_unpin.restype = C.c_int32
        L.d4b_host_unpin.argtypes = [C.c_void_p]
    if hasattr(L, "d4b_mesh_halo_mode"):   # (older builds of the library are still loadable for A/B runs)
        L.d4b_mesh_halo_mode.restype = C.c_int32
        L.d4b_mesh_halo_mode.argtypes = [C.c_void_p]
    for fn in (L.d4b_apply, L.d4b_apply_stiffness):
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        fn.restype = None
    L.d4b_sync.argtypes = [C.c_void_p]
    L.d4b_interp_to_quad.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.d4b_apply_galerkin.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.d4b_apply_mass.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.d4b_mesh_set_mass_diag.argtypes = [C.c_void_p, C.c_void_p]
    L.d4b_estimator_bi.argtypes = [C.c_void_p, C.POINTER(EstimatorDesc), C.c_void_p, C.c_void_p]
    L.d4b_build_rhs.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
    L.d4b_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    L.d4b_vec_dot.restype = C.c_double
    L.d4b_vec_dot.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
    L.d4b_vec_dot_noalloc.restype = C.c_int32
    L.d4b_vec_dot_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p,
                                      C.c_char_p, C.c_int64]
    L.d4b_vec_axpy.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.c_int64]
    L.d4b_vec_xpby.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_int64]
    L.d4b_vec_scale.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_int64]
    L.d4b_cg_eigs.restype = C.c_double
    L.d4b_cg_eigs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
    L.d4b_cheby_smooth.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                   C.c_double, C.c_double]
    L.d4b_mg_create.restype = C.c_void_p
    L.d4b_mg_create.argtypes = [C.POINTER(MgDesc)]
    L.d4b_mg_destroy.argtypes = [C.c_void_p]
    L.d4b_mg_vcycle.restype = C.c_double
    L.d4b_mg_vcycle.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.d4b_mg_solve.restype = C.c_int32
    L.d4b_mg_solve.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.d4b_mg_get_eigs.argtypes = [C.c_void_p, C.c_void_p]
    if hasattr(L, "d4b_mg_graph_replays"):
        L.d4b_mg_graph_replays.restype = C.c_int64
        L.d4b_mg_graph_replays.argtypes = [C.c_void_p]
    L.d4b_mg_restrict.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
    L.d4b_mg_prolong.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
    L.d4b_cg_solve.restype = C.c_int32
    L.d4b_cg_solve.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_double, C.c_double]
    L.d4b_fcg_solve.restype = C.c_int32
    L.d4b_fcg_solve.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_double, C.c_double,
                                C.c_int32, C.c_void_p]
    L.d4b_mesh_set_nonlinear_term.argtypes = [C.c_void_p, C.POINTER(NonlinearDesc)]
    L.d4b_mesh_linearize.argtypes = [C.c_void_p, C.c_void_p]
    L.d4b_build_residual.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.d4b_mg_set_coarse_mass.argtypes = [C.c_void_p, C.c_int32]
    L.d4b_newton_solve.restype = C.c_int32
    L.d4b_newton_solve.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(NewtonParams),
                                   C.POINTER(NewtonStats)]
    L.d4b_comm_get_unique_id.argtypes = [C.c_void_p]
    L.d4b_comm_init.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.d4b_comm_size.restype = C.c_int
    L.d4b_comm_rank.restype = C.c_int
    _lib = L
    return L


def cheby_params(imax, eigs_cg_imax, ratio=30.0, multiplier=1.0, use_new=0, zero_guess=0, reuse_down=0, reuse_last=0):
    return ChebyParams(imax, eigs_cg_imax, ratio, multiplier, use_new, zero_guess, reuse_down, reuse_last)


class Multigrid:
    """Device-resident V-cycle (d4b_mg_t) over `meshes` (coarsest first) and, per
    coarse level, (child_first, child_count) int32 arrays."""

    def __init__(self, meshes, transfers, smoother, vcycle_imax=1, vcycle_rtol=1e-9, vcycle_atol=0.0,
                 bottom_solver="cg", bottom_iter=100, bottom_rtol=1e-10, bottom_atol=0.0, bottom_cheby=None):
        L = load_library()
        self.L, self.meshes = L, meshes
        n = len(meshes)
        d = MgDesc()
        d.n_levels = n
        self._mesh_arr = (C.c_void_p * n)(*[m.h for m in meshes])
        d.meshes = C.cast(self._mesh_arr, C.POINTER(C.c_void_p))
        self._keep = []
        firsts, counts = (_I32P_T * (n - 1))(), (_I32P_T * (n - 1))()
        for l, (cf, cc) in enumerate(transfers):
            cf = np.ascontiguousarray(cf, dtype=np.int32); cc = np.ascontiguousarray(cc, dtype=np.int32)
            self._keep += [cf, cc]
            firsts[l] = cf.ctypes.data_as(_I32P_T); counts[l] = cc.ctypes.data_as(_I32P_T)
        self._firsts, self._counts = firsts, counts
        d.child_first = C.cast(firsts, C.POINTER(_I32P_T)); d.child_count = C.cast(counts, C.POINTER(_I32P_T))
        d.vcycle_imax, d.vcycle_rtol, d.vcycle_atol = vcycle_imax, vcycle_rtol, vcycle_atol
        d.smoother = smoother
        d.bottom_solver = 0 if bottom_solver == "cg" else 1
        d.bottom_iter, d.bottom_rtol, d.bottom_atol = bottom_iter, bottom_rtol, bottom_atol
        d.bottom_cheby = bottom_cheby if bottom_cheby is not None else smoother
        self.desc = d
        self.vcycle_imax = vcycle_imax
        self.h = L.d4b_mg_create(C.byref(d))

    def vcycle(self, u, rhs):
        return float(self.L.d4b_mg_vcycle(self.h, _ptr(u), _ptr(rhs)))

    def solve(self, u, rhs):
        hist = np.zeros(self.vcycle_imax + 1)
        n = int(self.L.d4b_mg_solve(self.h, _ptr(u), _ptr(rhs), hist.ctypes.data))
        return n, hist[: n + 1]

    def set_coarse_mass(self, mode="reference"):
        """the reference's multigrid matrix operator on the coarse levels of a linearised problem: "reference" =
        what it executes (d4b_mg_set_coarse_mass mode 1), "galerkin" = P^T M P (mode 2), None / False = off"""
        self.L.d4b_mg_set_coarse_mass(self.h, {None: 0, False: 0, "off": 0, "reference": 1, True: 1, "galerkin": 2}[mode])

    def eigs(self):
        e = np.zeros(len(self.meshes))
        self.L.d4b_mg_get_eigs(self.h, e.ctypes.data)
        return e

    def graph_replays(self):
        """V-cycles replayed from a recorded CUDA graph so far"""
        return int(self.L.d4b_mg_graph_replays(self.h))

    def restrict(self, coarse_level, fine, coarse):
        self.L.d4b_mg_restrict(self.h, coarse_level, _ptr(fine), _ptr(coarse))

    def prolong(self, coarse_level, coarse, fine):
        self.L.d4b_mg_prolong(self.h, coarse_level, _ptr(coarse), _ptr(fine))

    def close(self):
        if getattr(self, "h", None):
            self.L.d4b_mg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


COMM_ID_BYTES = 128


def comm_init_from_torch(device):
    """Create the library's NCCL communicator for the ranks of the default
    torch.distributed group: rank 0 draws the id, torch broadcasts its bytes."""
    import torch
    import torch.distributed as dist

    L = load_library()
    rank, world = dist.get_rank(), dist.get_world_size()
    L.d4b_init(device)
    buf = C.create_string_buffer(COMM_ID_BYTES)
    if rank == 0:
        L.d4b_comm_get_unique_id(buf)
    backend = dist.get_backend()
    t = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
    if backend == "nccl":
        t = t.cuda(device)
    dist.broadcast(t, src=0)
    raw = bytes(t.cpu().numpy().tobytes())
    L.d4b_comm_init(C.create_string_buffer(raw, COMM_ID_BYTES), rank, world)


def comm_destroy():
    load_library().d4b_comm_destroy()


def _ptr(x):
    """numpy array (host) or torch tensor (host/device) -> raw address."""
    if isinstance(x, np.ndarray):
        assert x.dtype == np.float64 and x.flags["C_CONTIGUOUS"]
        return x.ctypes.data
    # torch tensor
    assert x.dtype.is_floating_point and x.element_size() == 8 and x.is_contiguous()
    return x.data_ptr()


def vec_dot_noalloc(x, y, n, nvec, out, stream=None):
    """d4b_vec_dot_noalloc: out[v] = x[v*n:(v+1)*n] . y[v*n:(v+1)*n] for device tensors, enqueued on
    `stream` (raw cudaStream_t as int, None = legacy default stream) with no allocation on the
    path; raises ValueError (bad / too large arguments) or RuntimeError (CUDA) with the library's
    message.  x, y, out may be None only where the C entry point accepts NULL."""
    L = load_library()
    err = C.create_string_buffer(512)
    rc = L.d4b_vec_dot_noalloc(None if x is None else _ptr(x), None if y is None else _ptr(y), n, nvec,
                               None if out is None else _ptr(out), stream, err, len(err))
    if rc != 0:
        raise (RuntimeError if rc == 3 else ValueError)(err.value.decode())


def geom_compute(arrays, geom, want=None, device=0, out=None):
    """d4b_geom_compute: geometric factors of the mesh whose topology is `arrays` (the geometric
    arrays in it are ignored) for the map `geom` (dict, see abi.make_geom_desc).  Returns a dict of
    numpy arrays for the names in `want` (default: all), or fills the torch / numpy buffers in `out`."""
    L = load_library()
    if L.d4b_device_count() <= 0:
        raise RuntimeError("no CUDA device visible: the B200 path has no CPU fallback")
    L.d4b_init(device)
    desc, keep = make_desc({k: v for k, v in arrays.items() if k not in GEOM_ARRAY_FIELDS})
    gd, gkeep = make_geom_desc(geom)
    ln, lq, mp = int(desc.local_nodes), int(desc.local_nodes_quad), int(desc.local_mortar_nodes_quad)
    sizes = {"xyz": 3 * ln, "J_quad": lq, "rst_xyz_quad": 9 * lq, "sj_m_mortar_quad": mp, "n_m_mortar_quad": 3 * mp,
             "drst_dxyz_m_mortar_quad": 9 * mp, "drst_dxyz_p_mortar_quad_porder": 9 * mp, "hm_mortar_quad": mp,
             "hp_mortar_quad": mp}
    ga = GeomArrays()
    res = {}
    if out is None:
        for name in (want or GEOM_ARRAY_FIELDS):
            res[name] = np.zeros(sizes[name])
            setattr(ga, name, res[name].ctypes.data)
    else:
        for name, buf in out.items():
            setattr(ga, name, _ptr(buf))
        res = out
    L.d4b_geom_compute(C.byref(desc), C.byref(gd), C.byref(ga))
    return res


class Mesh:
    """Device-resident mesh (d4b_mesh_t).  With `geom` (dict, abi.make_geom_desc) the geometric
    factors are generated on the device (d4b_mesh_create_with_geometry) and `arrays` only needs the
    topology."""

    def __init__(self, arrays, device=0, geom=None):
        L = load_library()
        if L.d4b_device_count() <= 0:
            raise RuntimeError("no CUDA device visible: the B200 path has no CPU fallback")
        L.d4b_init(device)
        self.L = L
        if geom is not None:
            self.desc, self._keep = make_desc({k: v for k, v in arrays.items() if k not in GEOM_ARRAY_FIELDS})
            self.gdesc, self._gkeep = make_geom_desc(geom)
            self.h = L.d4b_mesh_create_with_geometry(C.byref(self.desc), C.byref(self.gdesc))
        else:
            self.desc, self._keep = make_desc(arrays)
            self.h = L.d4b_mesh_create(C.byref(self.desc))
        self.n = int(L.d4b_mesh_local_nodes(self.h))

    def set_stream(self, cuda_stream_handle):
        self.L.d4b_set_stream(self.h, C.c_void_p(cuda_stream_handle))

    def fast_elements(self):
        return int(self.L.d4b_mesh_fast_elements(self.h))

    def flat_elements(self):
        return int(self.L.d4b_mesh_flat_elements(self.h))

    def hp_elements(self):
        return int(self.L.d4b_mesh_hp_elements(self.h))

    def launches_per_apply(self):
        return int(self.L.d4b_mesh_launches_per_apply(self.h))

    def halo_mode(self):
        """"none" (one rank), "nccl", "peer" (neighbours' memory over NVLink) or "undecided" (before the first apply)"""
        return {0: "none", 1: "nccl", 2: "peer", -1: "undecided"}[int(self.L.d4b_mesh_halo_mode(self.h))]

    def device_bytes(self):
        return int(self.L.d4b_mesh_device_bytes(self.h))

    def apply(self, u, Au):
        self.L.d4b_apply(self.h, _ptr(u), _ptr(Au))

    def apply_stiffness(self, u, Au):
        self.L.d4b_apply_stiffness(self.h, _ptr(u), _ptr(Au))

    def sync(self):
        self.L.d4b_sync(self.h)

    # volume quadrature operators (d4est_quadrature_interpolate / apply_galerkin_integral /
    # apply_mass_matrix, src/Quadrature/d4est_quadrature.c:966-1016, :142-213, :385-477)
    def interp_to_quad(self, u, u_quad):
        self.L.d4b_interp_to_quad(self.h, _ptr(u), _ptr(u_quad))

    def apply_galerkin(self, in_quad, jac_quad, out):
        self.L.d4b_apply_galerkin(self.h, _ptr(in_quad), None if jac_quad is None else _ptr(jac_quad), _ptr(out))

    def apply_mass(self, x, jac_quad, out):
        self.L.d4b_apply_mass(self.h, _ptr(x), _ptr(jac_quad), _ptr(out))

    def build_rhs(self, f, jac_quad, rhs, on_quad=False):
        """d4est_laplacian_build_rhs_with_strong_bc: rhs = M f - A(0) with the mesh's Dirichlet data"""
        self.L.d4b_build_rhs(self.h, _ptr(f), 1 if on_quad else 0, _ptr(jac_quad), _ptr(rhs))

    def estimator_bi(self, u, residual, jac_quad, diam_volume, sj, hm, hp, prefactor, family="houston_maxp_minh",
                     dirichlet_g=None, eta2=None, terms=None):
        """d4est_estimator_bi_compute: eta^2 per local element (and, in `terms`, its four parts [4][n_local])"""
        d = EstimatorDesc()
        d.u, d.residual, d.jac_quad, d.diam_volume = _ptr(u), _ptr(residual), _ptr(jac_quad), _ptr(diam_volume)
        d.sj_m_mortar_quad, d.hm_mortar_quad, d.hp_mortar_quad = _ptr(sj), _ptr(hm), _ptr(hp)
        d.dirichlet_g_mortar_quad = None if dirichlet_g is None else _ptr(dirichlet_g)
        d.penalty_family, d.penalty_prefactor = EST_FAMILY[family], float(prefactor)
        if eta2 is None:
            eta2 = np.zeros(int(self.desc.n_local))
        self.L.d4b_estimator_bi(self.h, C.byref(d), _ptr(eta2), None if terms is None else _ptr(terms))
        return eta2

    def set_nonlinear_term(self, kind=None, punctures=None, star=None, xyz_quad=None, jac_quad=None, a_quad=None,
                           b_quad=None, exponent=0):
        """attach f(x, u) = a(x) (b(x) + u)^k (d4b_mesh_set_nonlinear_term); kind None removes it.
        punctures: array [n][10] = M, X, Y, Z, PX, PY, PZ, SX, SY, SZ; star: (rho0, R, cx, cy, cz)"""
        if kind is None:
            self.L.d4b_mesh_set_nonlinear_term(self.h, None)
            return
        d = NonlinearDesc()
        d.kind = NL_KIND[kind]
        if punctures is not None:
            P = np.asarray(punctures, dtype=np.float64).reshape(-1, 10)
            d.n_punctures = P.shape[0]
            for i in range(P.shape[0]):
                for j in range(10):
                    d.punctures[i][j] = float(P[i, j])
        if star is not None:
            d.star_rho0, d.star_R = float(star[0]), float(star[1])
            for j in range(3):
                d.star_c[j] = float(star[2 + j])
        d.exponent = int(exponent)
        for name, val in (("a_quad", a_quad), ("b_quad", b_quad), ("xyz_quad", xyz_quad), ("jac_quad", jac_quad)):
            setattr(d, name, None if val is None else _ptr(val))
        self.L.d4b_mesh_set_nonlinear_term(self.h, C.byref(d))

    def linearize(self, u0):
        """every later apply is the Jacobian about u0 (None drops the mass-type term)"""
        self.L.d4b_mesh_linearize(self.h, None if u0 is None else _ptr(u0))

    def build_residual(self, u, F, rhs=None):
        """F = A(u) + N(u) - rhs (d4b_build_residual)"""
        self.L.d4b_build_residual(self.h, _ptr(u), None if rhs is None else _ptr(rhs), _ptr(F))

    def newton_solve(self, u, pc=None, mesh_jac=None, atol=1e-15, rtol=1e-8, imin=1, imax=20, fcg_atol=1e-15,
                     fcg_rtol=5e-2, fcg_imax=100, fcg_flags=0):
        """d4est_solver_newton_solve + d4est_solver_fcg_solve + the V-cycle preconditioner on the device;
        returns (ierr, steps, fcg iterations per step, residual norms, fine-level Jacobian applies)"""
        prm = NewtonParams(atol, rtol, imin, imax, fcg_atol, fcg_rtol, fcg_imax, fcg_flags)
        st = NewtonStats()
        mj = mesh_jac if mesh_jac is not None else self
        ierr = int(self.L.d4b_newton_solve(self.h, mj.h, pc.h if pc is not None else None, _ptr(u), C.byref(prm),
                                           C.byref(st)))
        n = int(st.steps)
        return ierr, n, list(st.fcg_iterations[:n]), list(st.residual_norm[:n + 1]), int(st.jacobian_applies)

    def set_mass_diag(self, jac_quad):
        """every later apply adds V^T diag(w jac_quad) V u (None removes the term)"""
        self.L.d4b_mesh_set_mass_diag(self.h, None if jac_quad is None else _ptr(jac_quad))

    def dot(self, x, y):
        return float(self.L.d4b_vec_dot(self.h, _ptr(x), _ptr(y), self.n))

    def cg_eigs(self, u, rhs, imax, use_new=0, use_zero_guess=0):
        return float(self.L.d4b_cg_eigs(self.h, _ptr(u), _ptr(rhs), imax, use_new, use_zero_guess))

    def cg_solve(self, u, rhs, imax, rtol, atol=0.0):
        """d4est_solver_cg_solve on the device; returns the iteration count"""
        return int(self.L.d4b_cg_solve(self.h, _ptr(u), _ptr(rhs), imax, rtol, atol))

    def fcg_solve(self, u, rhs, imax, rtol, atol=0.0, pc=None, textbook=False):
        """d4est_solver_fcg_solve on the device (pc: a Multigrid whose finest mesh is this one, or None);
        returns (iterations, |r_k|^2 at the start of every iteration).  textbook=False reproduces the
        reference's executed arithmetic (see include/d4est_b200.h), True the recurrence of its comments."""
        hist = np.zeros(max(imax, 1))
        its = int(self.L.d4b_fcg_solve(self.h, pc.h if pc is not None else None, _ptr(u), _ptr(rhs), imax, rtol, atol,
                                       1 if textbook else 0, hist.ctypes.data))
        return its, hist[:its]

    def cheby_smooth(self, u, rhs, r, iters, lmin, lmax):
        self.L.d4b_cheby_smooth(self.h, _ptr(u), _ptr(rhs), _ptr(r), iters, lmin, lmax)

    def close(self):
        if getattr(self, "h", None):
            self.L.d4b_mesh_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
