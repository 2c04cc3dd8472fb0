"""Time the apply on a unit-cube brick: python profiles/flat_bench.py LEVEL DEG [DEG ...]  (env switches select kernel variants)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from disco4est_b200 import Mesh
from disco4est_b200.brick import brick_desc

level = int(sys.argv[1])
dev = torch.device("cuda:0")
for deg in [int(x) for x in sys.argv[2:]]:
    A, g = brick_desc(level, deg, geometry=False)
    m = Mesh(A, geom=g)
    stream = torch.cuda.current_stream()
    m.set_stream(stream.cuda_stream)
    u = torch.rand(m.n, dtype=torch.float64, device=dev, generator=torch.Generator(device=dev).manual_seed(1))
    Au = torch.empty_like(u)
    for _ in range(5):
        m.apply(u, Au)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 20
    e0.record(stream)
    for _ in range(steps):
        m.apply(u, Au)
    e1.record(stream)
    torch.cuda.synchronize()
    t = e0.elapsed_time(e1) * 1e-3 / steps
    print("brick level %d deg %d: %9d DOF  %.4f ms  %.2f GDOF/s  checksum %.12e" % (level, deg, m.n, 1e3 * t, m.n / t / 1e9, float(Au.sum())), flush=True)
    m.close()
