"""Worker for the multi-rank tests (launched by tests/test_multi_rank.py, one process per rank).

mode "gloo": CPU.  Every rank builds its part of a partitioned brick
(disco4est_b200.brick), exchanges mirror -> ghost nodal values exactly as the
exchange plan of the mesh description says (what d4est_ghost_data_exchange
does, src/Mesh/d4est_ghost_data.c:142-255) with gloo send/recv, applies the
ORACLE with those ghosts, and rank 0 compares the gathered result with the
single-rank oracle on the whole mesh.  This checks the partition, the ghost
numbering and the plan the CUDA library consumes.

mode "nccl": one GPU per rank.  The same mesh through the C ABI: d4b_comm_init
from a torch-broadcast NCCL id, d4b_apply with the halo exchange inside, plus
the all-reduced dot product; rank 0 compares with the oracle.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def n3_of(p):
    return (p + 1) ** 3


def exchange_host(A, ul, dist, torch):
    """mirror -> ghost exchange by the plan in A, nodal values, host buffers."""
    rank, world = A["mpirank"], A["mpisize"]
    n3 = n3_of(int(A["deg"][0]))
    ug = np.zeros(max(int(A["ghost_nodes"]), 1))
    reqs, keep = [], []
    for q in range(world):
        if q == rank:
            continue
        m0, m1 = A["mirror_proc_offsets"][q], A["mirror_proc_offsets"][q + 1]
        if m1 > m0:
            els = A["mirror_elem"][A["mirror_proc_mirrors"][m0:m1]]
            buf = np.concatenate([ul[A["nodal_stride"][e]:A["nodal_stride"][e] + n3] for e in els])
            t = torch.from_numpy(np.ascontiguousarray(buf))
            keep.append(t)
            reqs.append(dist.isend(t, q))
        g0, g1 = A["ghost_proc_offsets"][q], A["ghost_proc_offsets"][q + 1]
        if g1 > g0:
            lo = int(A["ghost_nodal_stride"][g0])
            t = torch.from_numpy(ug[lo:lo + (g1 - g0) * n3])
            reqs.append(dist.irecv(t, q))
    for r in reqs:
        r.wait()
    return ug


def main():
    mode, level, p = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    import torch
    import torch.distributed as dist
    from disco4est_b200.brick import brick_desc, partition_offsets, trees_for_world
    from oracle_lib import Oracle

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    if mode == "nccl":
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
    else:
        dist.init_process_group("gloo")
    trees = trees_for_world(world)
    full = brick_desc(level, p, trees=trees)
    n3 = n3_of(p)
    u = np.random.default_rng(7).standard_normal(full["local_nodes"])
    offs = partition_offsets(full["n_local"], world) * n3
    ul = np.ascontiguousarray(u[offs[rank]:offs[rank + 1]])
    A = brick_desc(level, p, rank=rank, world=world, trees=trees)

    if mode == "gloo":
        ug = exchange_host(A, ul, dist, torch)
        # the exchange delivered exactly the neighbours' values
        want = np.concatenate([u[g * n3:(g + 1) * n3] for g in A["ghost_global_id"]])
        assert np.array_equal(ug[:want.size], want)
        au = Oracle(A).apply(ul, ug)
        local_dot = float(ul @ au)
        tdot = torch.tensor([local_dot], dtype=torch.float64)
        dist.all_reduce(tdot)
        gdot = float(tdot[0])
    else:
        from disco4est_b200 import Mesh
        from disco4est_b200.runtime import comm_init_from_torch, comm_destroy

        comm_init_from_torch(rank)
        m = Mesh(A, device=rank)
        assert m.fast_elements() == A["n_local"], "every brick element must take a fast kernel, ghosts included"
        au = np.zeros_like(ul)
        m.apply(ul, au)
        # device-pointer path + all-reduced dot
        du = torch.from_numpy(ul).cuda(rank)
        dau = torch.empty_like(du)
        m.apply(du, dau)
        m.sync()
        assert np.array_equal(dau.cpu().numpy(), au)
        gdot = m.dot(du, dau)
        m.close()
        comm_destroy()

    gathered = [torch.zeros(int(offs[r + 1] - offs[r]), dtype=torch.float64) for r in range(world)] if rank == 0 else None
    if mode == "nccl":
        # gather through the host with a gloo side group (sizes may differ per rank)
        side = dist.new_group(backend="gloo")
        dist.gather(torch.from_numpy(au), gathered, dst=0, group=side)
    else:
        dist.gather(torch.from_numpy(au), gathered, dst=0)
    if rank == 0:
        got = np.concatenate([g.numpy() for g in gathered])
        ref = Oracle(full).apply(u)
        err = np.linalg.norm(got - ref) / np.linalg.norm(ref)
        derr = abs(gdot - float(u @ ref)) / abs(float(u @ ref))
        print("MP_RESULT mode=%s world=%d rel_err=%.3e dot_err=%.3e" % (mode, world, err, derr), flush=True)
        assert err < 1e-12 and derr < 1e-12
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
