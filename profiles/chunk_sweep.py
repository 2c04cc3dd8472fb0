"""Host-pointer (pipelined) apply at 64^3 p=3 for several transfer-chunk counts
(D4B_PIPELINE_CHUNKS): ms per call and GDOF/s, checksum against the device-pointer apply.
Run on the GPU box: python profiles/chunk_sweep.py [K ...]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from disco4est_b200 import Mesh  # noqa: E402
from disco4est_b200.brick import brick_desc  # noqa: E402

level, p = int(os.environ.get("LEVEL", 6)), 3
A = brick_desc(level, p)
n = int(A["local_nodes"])
hu = torch.empty(n, dtype=torch.float64).pin_memory()
hu.copy_(torch.from_numpy(np.random.default_rng(1).standard_normal(n)))
hAu = torch.empty(n, dtype=torch.float64).pin_memory()
ref = None
for K in [int(a) for a in sys.argv[1:]] or [32, 16, 64, 128, 256]:
    os.environ["D4B_PIPELINE_CHUNKS"] = str(K)
    m = Mesh(A)
    for _ in range(3):
        m.apply(hu.numpy(), hAu.numpy())
    ts = []
    for _ in range(10):
        t0 = time.perf_counter()
        m.apply(hu.numpy(), hAu.numpy())
        ts.append(time.perf_counter() - t0)
    if ref is None:
        du = hu.cuda()
        dAu = torch.zeros_like(du)
        m.apply(du, dAu)
        torch.cuda.synchronize()
        ref = dAu.cpu()
    same = bool(torch.equal(ref, hAu))
    t = float(np.median(ts))
    print("K=%4d  median %.3f ms  min %.3f ms  %.2f GDOF/s  identical=%s" % (K, t * 1e3, min(ts) * 1e3, n / t / 1e9, same), flush=True)
    m.close()
