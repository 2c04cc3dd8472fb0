"""ctypes binding of the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE: the oracle is the checker, never the thing measured or
shipped.  Built by __graft_entry__.build() / oracle/Makefile.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from disco4est_b200.abi import MeshDesc, make_desc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "liboracle.so")
_F64P = C.POINTER(C.c_double)


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"])


def lib():
    if not os.path.exists(LIB):
        build()
    L = C.CDLL(LIB)
    L.d4bo_create.restype = C.c_void_p
    L.d4bo_create.argtypes = [C.POINTER(MeshDesc)]
    L.d4bo_destroy.argtypes = [C.c_void_p]
    for fn in (L.d4bo_apply_stiffness,):
        fn.argtypes = [C.c_void_p, _F64P, _F64P]
    L.d4bo_apply.argtypes = [C.c_void_p, _F64P, _F64P, _F64P]
    L.d4bo_apply_flux.argtypes = [C.c_void_p, _F64P, _F64P, _F64P]
    L.d4bo_interp_to_quad.argtypes = [C.c_void_p, _F64P, _F64P]
    L.d4bo_apply_galerkin.argtypes = [C.c_void_p, _F64P, _F64P, _F64P]
    L.d4bo_apply_mass.argtypes = [C.c_void_p, _F64P, _F64P, _F64P]
    L.d4bo_gauss.argtypes = [C.c_int, _F64P, _F64P]
    L.d4bo_lobatto.argtypes = [C.c_int, _F64P, _F64P]
    L.d4bo_dij.argtypes = [C.c_int, _F64P]
    L.d4bo_eval_matrix.argtypes = [C.c_int, C.c_int, C.c_int, _F64P]
    return L


def _p(a):
    return a.ctypes.data_as(_F64P) if a is not None else None


class Oracle:
    def __init__(self, arrays):
        self.ctx = None
        self.L = lib()
        self.desc, self._keep = make_desc(arrays)
        self.ctx = self.L.d4bo_create(C.byref(self.desc))
        self.n = int(self.desc.local_nodes)

    def apply(self, u, ughost=None):
        u = np.ascontiguousarray(u, dtype=np.float64)
        Au = np.zeros(self.n)
        self.L.d4bo_apply(self.ctx, _p(u), _p(ughost), _p(Au))
        return Au

    def apply_stiffness(self, u):
        u = np.ascontiguousarray(u, dtype=np.float64)
        Au = np.zeros(self.n)
        self.L.d4bo_apply_stiffness(self.ctx, _p(u), _p(Au))
        return Au

    def interp_to_quad(self, u):
        u = np.ascontiguousarray(u, dtype=np.float64)
        uq = np.zeros(int(self.desc.local_nodes_quad))
        self.L.d4bo_interp_to_quad(self.ctx, _p(u), _p(uq))
        return uq

    def apply_galerkin(self, in_quad, jac_quad=None):
        in_quad = np.ascontiguousarray(in_quad, dtype=np.float64)
        jac_quad = None if jac_quad is None else np.ascontiguousarray(jac_quad, dtype=np.float64)
        out = np.zeros(self.n)
        self.L.d4bo_apply_galerkin(self.ctx, _p(in_quad), _p(jac_quad), _p(out))
        return out

    def apply_mass(self, x, jac_quad):
        x = np.ascontiguousarray(x, dtype=np.float64)
        jac_quad = np.ascontiguousarray(jac_quad, dtype=np.float64)
        out = np.zeros(self.n)
        self.L.d4bo_apply_mass(self.ctx, _p(x), _p(jac_quad), _p(out))
        return out

    def close(self):
        if self.ctx:
            self.L.d4bo_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        self.close()
