#!/usr/bin/env python
"""Where a V-cycle's time goes: per-level cg_eigs and Chebyshev sweeps, transfers, bottom solve."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from disco4est_b200 import Mesh
from disco4est_b200.brick import brick_hierarchy

level, deg = int(sys.argv[1]) if len(sys.argv) > 1 else 6, 3
descs, transfers = brick_hierarchy(level, deg)
meshes = [Mesh(d) for d in descs]
out = {}
for l, m in enumerate(meshes):
    n = m.n
    rhs = torch.rand(n, dtype=torch.float64, device="cuda")
    u = torch.zeros_like(rhs); r = torch.zeros_like(rhs); au = torch.zeros_like(rhs)
    def timeit(f, reps=3):
        f(); torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(reps): f()
        torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3
    lmax = m.cg_eigs(u.clone(), rhs, 15)
    out["level%d" % l] = {"nodes": n,
                          "apply_ms": timeit(lambda: (m.apply(u, au), m.sync())),
                          "cg_eigs15_ms": timeit(lambda: m.cg_eigs(u, rhs, 15)),
                          "cheby15_ms": timeit(lambda: m.cheby_smooth(u, rhs, r, 15, lmax / 30, lmax)),
                          "dot_ms": timeit(lambda: m.dot(u, rhs))}
print(json.dumps(out, indent=1))
