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

from .abi import ChebyParams, MeshDesc, make_desc

_HERE = os.path.dirname(os.path.abspath(__file__))
# D4B_LIBRARY_PATH lets profiling scripts A/B differently compiled builds of the same library
LIB_PATH = os.environ.get("D4B_LIBRARY_PATH") or os.path.join(_HERE, "libd4est_b200.so")
_F64P = C.POINTER(C.c_double)

EXPORTS = [
    "d4b_set_abort_hook", "d4b_abi_version", "d4b_device_count", "d4b_init",
    "d4b_mesh_create", "d4b_mesh_destroy", "d4b_mesh_local_nodes", "d4b_mesh_device_bytes", "d4b_mesh_fast_elements", "d4b_mesh_flat_elements", "d4b_mesh_launches_per_apply",
    "d4b_apply", "d4b_apply_stiffness", "d4b_sync", "d4b_set_stream",
    "d4b_vec_dot", "d4b_vec_axpy", "d4b_vec_xpby", "d4b_vec_scale",
    "d4b_cg_eigs", "d4b_cheby_smooth",
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
    L.d4b_mesh_local_nodes.restype = C.c_int64
    L.d4b_mesh_local_nodes.argtypes = [C.c_void_p]
    L.d4b_mesh_device_bytes.restype = C.c_int64
    L.d4b_mesh_device_bytes.argtypes = [C.c_void_p]
    L.d4b_mesh_fast_elements.restype = C.c_int64
    L.d4b_mesh_fast_elements.argtypes = [C.c_void_p]
    L.d4b_mesh_flat_elements.restype = C.c_int64
    L.d4b_mesh_flat_elements.argtypes = [C.c_void_p]
    L.d4b_mesh_launches_per_apply.restype = C.c_int32
    L.d4b_mesh_launches_per_apply.argtypes = [C.c_void_p]
    for fn in (L.d4b_apply, L.d4b_apply_stiffness):
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        fn.restype = None
    L.d4b_sync.argtypes = [C.c_void_p]
    L.d4b_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    L.d4b_vec_dot.restype = C.c_double
    L.d4b_vec_dot.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
    L.d4b_vec_axpy.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.c_int64]
    L.d4b_vec_xpby.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_int64]
    L.d4b_vec_scale.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_int64]
    L.d4b_cg_eigs.restype = C.c_double
    L.d4b_cg_eigs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
    L.d4b_cheby_smooth.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                   C.c_double, C.c_double]
    L.d4b_comm_get_unique_id.argtypes = [C.c_void_p]
    L.d4b_comm_init.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.d4b_comm_size.restype = C.c_int
    L.d4b_comm_rank.restype = C.c_int
    _lib = L
    return L


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


class Mesh:
    """Device-resident mesh (d4b_mesh_t)."""

    def __init__(self, arrays, device=0):
        L = load_library()
        if L.d4b_device_count() <= 0:
            raise RuntimeError("no CUDA device visible: the B200 path has no CPU fallback")
        L.d4b_init(device)
        self.L = L
        self.desc, self._keep = make_desc(arrays)
        self.h = L.d4b_mesh_create(C.byref(self.desc))
        self.n = int(L.d4b_mesh_local_nodes(self.h))

    def set_stream(self, cuda_stream_handle):
        self.L.d4b_set_stream(self.h, C.c_void_p(cuda_stream_handle))

    def fast_elements(self):
        return int(self.L.d4b_mesh_fast_elements(self.h))

    def flat_elements(self):
        return int(self.L.d4b_mesh_flat_elements(self.h))

    def launches_per_apply(self):
        return int(self.L.d4b_mesh_launches_per_apply(self.h))

    def device_bytes(self):
        return int(self.L.d4b_mesh_device_bytes(self.h))

    def apply(self, u, Au):
        self.L.d4b_apply(self.h, _ptr(u), _ptr(Au))

    def apply_stiffness(self, u, Au):
        self.L.d4b_apply_stiffness(self.h, _ptr(u), _ptr(Au))

    def sync(self):
        self.L.d4b_sync(self.h)

    def dot(self, x, y):
        return float(self.L.d4b_vec_dot(self.h, _ptr(x), _ptr(y), self.n))

    def cg_eigs(self, u, rhs, imax, use_new=0, use_zero_guess=0):
        return float(self.L.d4b_cg_eigs(self.h, _ptr(u), _ptr(rhs), imax, use_new, use_zero_guess))

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
