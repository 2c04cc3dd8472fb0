"""Worker for the multi-rank tests (launched by tests/test_multi_rank.py, one process per rank).

Mesh: a synthetic brick ("<level> <p>") or a reference fixture split by
disco4est_b200.partition ("golden:<case> 0": hp meshes with hanging faces, curved trees).

mode "gloo": CPU.  Every rank builds its part of the partitioned mesh, exchanges mirror -> ghost nodal values exactly as the
exchange plan of the mesh description says (what d4est_ghost_data_exchange
does, src/Mesh/d4est_ghost_data.c:142-255) with gloo send/recv, applies the
ORACLE with those ghosts, and rank 0 compares the gathered result with the
single-rank oracle on the whole mesh.  This checks the partition, the ghost
numbering and the plan the CUDA library consumes.

mode "nccl": one GPU per rank.  The same mesh through the C ABI: d4b_comm_init
from a torch-broadcast NCCL id, d4b_apply with the halo inside -- once over NCCL
send/recv and once over peer memory (the neighbours' vectors mapped with CUDA IPC),
compared bit for bit, also over back-to-back applies and the fused smoother -- plus
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


def _i(x):
    """fixture scalars are stored as 1-element arrays"""
    return int(np.asarray(x).reshape(-1)[0])


def exchange_host(A, ul, dist, torch):
    """mirror -> ghost exchange by the plan in A, nodal values, host buffers."""
    rank, world = A["mpirank"], A["mpisize"]
    ug = np.zeros(max(_i(A["ghost_nodes"]), 1))
    gend = np.concatenate([A["ghost_nodal_stride"], [_i(A["ghost_nodes"])]]).astype(np.int64)
    reqs, keep = [], []
    for q in range(world):
        if q == rank:
            continue
        m0, m1 = A["mirror_proc_offsets"][q], A["mirror_proc_offsets"][q + 1]
        if m1 > m0:
            els = A["mirror_elem"][A["mirror_proc_mirrors"][m0:m1]]
            buf = np.concatenate([ul[A["nodal_stride"][e]:A["nodal_stride"][e] + n3_of(int(A["deg"][e]))] for e in els])
            t = torch.from_numpy(np.ascontiguousarray(buf))
            keep.append(t)
            reqs.append(dist.isend(t, q))
        g0, g1 = A["ghost_proc_offsets"][q], A["ghost_proc_offsets"][q + 1]
        if g1 > g0:
            t = torch.from_numpy(ug[int(gend[g0]):int(gend[g1])])
            reqs.append(dist.irecv(t, q))
    for r in reqs:
        r.wait()
    return ug


def multigrid_main(level, p):
    """mode "mg", one GPU per rank: a brick hierarchy split over the ranks; V-cycles enqueued kernel by kernel against
    V-cycles replayed from a CUDA graph recorded on every rank (peer-memory halo and all-reduce: nothing in the
    sequence calls NCCL, the epochs of the flag protocols are device counters) -- bit-equal residual histories and
    iterates, and the replays must actually have happened."""
    import torch
    import torch.distributed as dist
    from disco4est_b200 import Mesh
    from disco4est_b200.brick import brick_hierarchy, trees_for_world
    from disco4est_b200.runtime import Multigrid, cheby_params, comm_init_from_torch, comm_destroy

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
    comm_init_from_torch(rank)
    trees = trees_for_world(world)
    out = {}
    for graph in ("0", "1"):
        os.environ["D4B_MG_GRAPH"] = graph
        descs, transfers = brick_hierarchy(level, p, rank=rank, world=world, trees=trees, geometry=False)
        meshes = [Mesh(d, device=rank, geom=g) for d, g in descs]
        mg = Multigrid(meshes, transfers, cheby_params(4, 5, 30.0, 1.0), vcycle_imax=1, bottom_solver="cheby",
                       bottom_cheby=cheby_params(5, 6, 30.0, 1.0))
        n = meshes[-1].n
        rhs = torch.rand(n, dtype=torch.float64, device="cuda", generator=torch.Generator(device="cuda").manual_seed(5 + rank))
        u = torch.zeros_like(rhs)
        v = torch.empty_like(rhs)
        hist = []
        for k in range(12):
            hist.append(mg.vcycle(u, rhs))
            if k == 7:
                meshes[-1].apply(u, v)   # an apply between two cycles flips the parity of the top level's uhat copies
        meshes[-1].sync()
        assert all(mm.halo_mode() == "peer" for mm in meshes), [mm.halo_mode() for mm in meshes]
        replays = mg.graph_replays()
        assert (replays > 0) == (graph == "1"), "graph=%s but %d replays" % (graph, replays)
        out[graph] = (hist, u.clone())
        mg.close()
        for mm in meshes:
            mm.close()
    assert out["0"][0] == out["1"][0], "residual histories differ: %r %r" % (out["0"][0], out["1"][0])
    assert torch.equal(out["0"][1], out["1"][1])
    assert out["0"][0][-1] < 1e-3 * out["0"][0][0]
    if rank == 0:
        print("MP_RESULT mode=mg world=%d r2 %.3e -> %.3e" % (world, out["0"][0][0], out["0"][0][-1]), flush=True)
    comm_destroy()
    dist.barrier()
    dist.destroy_process_group()


def main():
    mode, mesh_arg, p = sys.argv[1], sys.argv[2], int(sys.argv[3])
    if mode == "mg":
        return multigrid_main(int(mesh_arg), p)
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
    if mesh_arg.startswith("golden:"):
        import d4bio
        from disco4est_b200.partition import partition_desc

        full = d4bio.load(os.path.join(HERE, "golden", mesh_arg[7:] + ".d4b.gz"))
        eoffs = partition_offsets(_i(full["n_local"]), world)
        nend = np.concatenate([full["nodal_stride"], [_i(full["local_nodes"])]]).astype(np.int64)
        offs = nend[eoffs]
        A = partition_desc(full, rank, world)
        gid = A["ghost_global_id"]
        ghost_want = [(int(nend[g]), int(nend[g + 1])) for g in gid]
    else:
        level = int(mesh_arg)
        trees = trees_for_world(world)
        full = brick_desc(level, p, trees=trees)
        n3 = n3_of(p)
        offs = partition_offsets(full["n_local"], world) * n3
        A = brick_desc(level, p, rank=rank, world=world, trees=trees)
        ghost_want = [(int(g) * n3, (int(g) + 1) * n3) for g in A["ghost_global_id"]]
    u = np.random.default_rng(7).standard_normal(_i(full["local_nodes"]))
    ul = np.ascontiguousarray(u[offs[rank]:offs[rank + 1]])

    if mode == "gloo":
        ug = exchange_host(A, ul, dist, torch)
        # the exchange delivered exactly the neighbours' values
        want = np.concatenate([u[a:b] for a, b in ghost_want]) if ghost_want else np.zeros(0)
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
        # two meshes of the same description: halo over NCCL send/recv, and over peer memory (the default on one node)
        os.environ["D4B_PEER_HALO"] = "0"
        m_nccl = Mesh(A, device=rank)
        au_nccl = np.zeros_like(ul)
        m_nccl.apply(ul, au_nccl)
        assert m_nccl.halo_mode() == "nccl", m_nccl.halo_mode()
        os.environ["D4B_PEER_HALO"] = "1"
        m = Mesh(A, device=rank)
        if not mesh_arg.startswith("golden:"):
            assert m.fast_elements() == A["n_local"], "every brick element must take a fast kernel, ghosts included"
        au = np.zeros_like(ul)
        m.apply(ul, au)
        # when no rank needs the general kernel, the peer halo must be what ran (on every rank or on none)
        t = torch.tensor([1 if m.fast_elements() == _i(A["n_local"]) else 0, 1 if m.halo_mode() == "peer" else 0,
                          -(1 if m.halo_mode() == "peer" else 0)], dtype=torch.int64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        assert int(t[1]) == -int(t[2]), "ranks disagree about the halo mode"
        assert int(t[0]) == 0 or int(t[1]) == 1, "peer halo was not activated: %s" % m.halo_mode()
        assert np.array_equal(au, au_nccl), "peer-memory halo and NCCL halo differ"
        # device-pointer path + all-reduced dot
        du = torch.from_numpy(ul).cuda(rank)
        dau = torch.empty_like(du)
        m.apply(du, dau)
        m.sync()
        assert np.array_equal(dau.cpu().numpy(), au)
        gdot = m.dot(du, dau)
        # back-to-back applies on alternating inputs without a host synchronisation in between: the two copies
        # of uhat and the epoch flags must keep them apart
        du2 = torch.from_numpy(np.random.default_rng(11 + rank).standard_normal(ul.size)).cuda(rank)
        ref2 = torch.empty_like(du)
        m_nccl.apply(du2, ref2)
        m_nccl.sync()
        outs = [torch.empty_like(du) for _ in range(7)]
        for k in range(7):
            m.apply(du if k % 2 == 0 else du2, outs[k])
        m.sync()
        for k in range(7):
            assert torch.equal(outs[k], dau if k % 2 == 0 else ref2), "apply %d of the back-to-back sequence differs" % k
        # fused Chebyshev smoother (u updated in place by the apply's epilogue) and the lambda_max estimate: with the
        # solver fusions off both halos run the same kernels on the same numbers -> bit-equal; with them on (bricks:
        # every element in the tangential-free kernel) the peer mesh carries the recurrence in the Gauss basis, one epoch
        # of the flag protocol per sweep, the NCCL mesh cannot (nodal mirrors) -> equal to round-off
        if m.fast_elements() == _i(A["n_local"]):
            for fused_solvers in ("0", "1"):
                os.environ["D4B_FUSED_SOLVERS"] = fused_solvers
                res = []
                for mm in (m_nccl, m):
                    us = torch.zeros_like(du)
                    rs = torch.empty_like(du)
                    lam = mm.cg_eigs(us, du, 5)
                    us.zero_()
                    mm.cheby_smooth(us, du, rs, 4, lam / 30.0, lam)
                    mm.cheby_smooth(us, du, rs, 3, lam / 30.0, lam)   # (odd and even sweep counts: both copies of uhat)
                    mm.sync()
                    res.append((lam, us.clone(), rs.clone()))
                if fused_solvers == "0":
                    assert res[0][0] == res[1][0], "cg_eigs differs between the two halos: %r %r" % (res[0][0], res[1][0])
                    assert torch.equal(res[0][1], res[1][1]) and torch.equal(res[0][2], res[1][2]), "smoother differs"
                else:
                    assert abs(res[0][0] - res[1][0]) <= 1e-12 * abs(res[0][0])
                    for a_, b_ in ((res[0][1], res[1][1]), (res[0][2], res[1][2])):
                        assert float((a_ - b_).norm()) <= 1e-11 * float(a_.norm()), "smoother differs beyond round-off"
            os.environ.pop("D4B_FUSED_SOLVERS", None)
            # the flag protocol is intact after the smoother's epochs
            chk = torch.empty_like(du)
            m.apply(du, chk)
            m.sync()
            assert torch.equal(chk, dau)
        m_nccl.close()
        m.close()
        comm_destroy()

    # gloo's gather wants equal sizes: pad to the largest rank, trim on rank 0
    sizes = [int(offs[r + 1] - offs[r]) for r in range(world)]
    padded = np.zeros(max(sizes))
    padded[:au.size] = au
    gathered = [torch.zeros(max(sizes), dtype=torch.float64) for r in range(world)] if rank == 0 else None
    if mode == "nccl":
        side = dist.new_group(backend="gloo")   # gather through the host
        dist.gather(torch.from_numpy(padded), gathered, dst=0, group=side)
    else:
        dist.gather(torch.from_numpy(padded), gathered, dst=0)
    if rank == 0:
        got = np.concatenate([g.numpy()[:sizes[r]] for r, g in enumerate(gathered)])
        ref = Oracle(full).apply(u)
        err = np.linalg.norm(got - ref) / np.linalg.norm(ref)
        derr = abs(gdot - float(u @ ref)) / abs(float(u @ ref))
        print("MP_RESULT mode=%s mesh=%s world=%d rel_err=%.3e dot_err=%.3e" % (mode, mesh_arg, world, err, derr), flush=True)
        assert err < 1e-12 and derr < 1e-12
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
