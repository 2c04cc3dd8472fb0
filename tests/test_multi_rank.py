"""Multi-rank path: partitioned mesh + halo exchange.

CPU (gloo, world_size 2 and 3): the partition / exchange plan against the
oracle.  GPU (needs >= 2 devices, skipped otherwise): the same through the C
ABI with the NCCL halo exchange inside d4b_apply.
"""
import os
import socket
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def run_ranks(mode, world, level, p, timeout=600):
    """level: brick refinement level, or "golden:<case>" for a partitioned reference fixture"""
    port = free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), LOCAL_RANK=str(r),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "mp_worker.py"), mode, str(level), str(p)],
                                      env=env, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    try:
        for pr in procs:
            out, _ = pr.communicate(timeout=timeout)
            outs.append(out)
    finally:
        for pr in procs:
            if pr.poll() is None:
                pr.kill()
    for r, (pr, out) in enumerate(zip(procs, outs)):
        assert pr.returncode == 0, "rank %d failed:\n%s" % (r, out)
    assert "MP_RESULT" in outs[0], outs[0]
    return outs[0]


@pytest.mark.parametrize("world,level,p", [(2, 2, 2), (3, 1, 3)])
def test_partition_and_exchange_plan_gloo(world, level, p):
    run_ranks("gloo", world, level, p)


@pytest.mark.gpu
@pytest.mark.parametrize("level,p", [(2, 1), (3, 3)])
def test_two_gpu_apply_matches_oracle(level, p):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    run_ranks("nccl", 2, level, p)


@pytest.mark.parametrize("world,case", [(2, "sphere_hp_amr1_zero"), (3, "sphere_hp_amr1_zero"), (2, "sphere_p2"),
                                        (3, "sphere_p_amr1"), (2, "sphere_p_amr1_robin")])
def test_partitioned_hp_mesh_gloo(world, case):
    """BASELINE config 4 in small: curved cubed sphere, hp-adapted, hanging faces across rank boundaries"""
    run_ranks("gloo", world, "golden:" + case, 0)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["sphere_hp_amr1_zero", "sphere_p2", "sphere_p_amr1"])
def test_two_gpu_hp_mesh_matches_oracle(case):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    run_ranks("nccl", 2, "golden:" + case, 0)


@pytest.mark.gpu
def test_two_gpu_vcycle_graph_replay():
    """multi-rank V-cycles replayed from a CUDA graph == enqueued kernel by kernel, bit for bit"""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    run_ranks("mg", 2, 3, 3)
