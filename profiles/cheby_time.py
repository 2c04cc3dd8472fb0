#!/usr/bin/env python
"""Time one Chebyshev smoothing sweep (cheby_imax iterations + final residual) on the
64^3 p=3 brick, fused epilogue vs apply + vector kernel (D4B_NO_FUSED_CHEBY=1)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from disco4est_b200 import Mesh
from disco4est_b200.brick import brick_desc

level, deg, iters = int(sys.argv[1]) if len(sys.argv) > 1 else 6, 3, 15
m = Mesh(brick_desc(level, deg))
n = m.n
stream = torch.cuda.current_stream()
m.set_stream(stream.cuda_stream)
g = torch.Generator(device="cuda").manual_seed(1)
rhs = torch.rand(n, dtype=torch.float64, device="cuda", generator=g)
u = torch.zeros_like(rhs); r = torch.zeros_like(rhs)
lmax = m.cg_eigs(u.clone(), rhs, 15)
out = {}
for name, env in (("fused", "0"), ("unfused", "1")):
    os.environ["D4B_NO_FUSED_CHEBY"] = env
    for _ in range(2):
        u.zero_(); m.cheby_smooth(u, rhs, r, iters, lmax / 30, lmax)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record(stream)
    for _ in range(reps):
        m.cheby_smooth(u, rhs, r, iters, lmax / 30, lmax)
    e1.record(stream); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    u.zero_(); m.cheby_smooth(u, rhs, r, iters, lmax / 30, lmax); torch.cuda.synchronize()
    out[name] = {"ms_per_sweep": ms, "ms_per_iteration": ms / (iters + 1), "gdof_s": n * (iters + 1) / ms / 1e6,
                 "rnorm": float(r.norm()), "unorm": float(u.norm())}
print(json.dumps({"workload": "%d^3 p=%d Chebyshev(%d) + residual" % (1 << level, deg, iters), "lmax": lmax, **out}))
