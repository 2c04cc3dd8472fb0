#!/usr/bin/env python
"""Device-timed apply (CUDA events, 30 applies after 5 warm-ups) on 2^level cubes, per-point metric:
usage apply_time.py level:p [level:p ...] -- for A/B runs of differently compiled libraries (D4B_LIBRARY_PATH)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("D4B_AFFINE", "0")
import torch
from disco4est_b200 import Mesh
from disco4est_b200.brick import brick_desc

tag = os.path.basename(os.environ.get("D4B_LIBRARY_PATH", "base"))
for arg in sys.argv[1:]:
    level, p = (int(x) for x in arg.split(":"))
    A, g = brick_desc(level, p, geometry=False)
    m = Mesh(A, geom=g)
    st = torch.cuda.current_stream()
    m.set_stream(st.cuda_stream)
    u = torch.rand(m.n, dtype=torch.float64, device="cuda")
    Au = torch.zeros_like(u)
    for _ in range(5):
        m.apply(u, Au)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(30):
            m.apply(u, Au)
        e1.record(st)
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 30)
    print("%-22s level %d p %d  %.4f ms  %.2f GDOF/s  checksum %.10e" % (tag, level, p, best, m.n / best / 1e6, float(Au.sum())), flush=True)
    m.close()
    del u, Au
