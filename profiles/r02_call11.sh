cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L | wc -l
nvidia-smi topo -m 2>&1 | head -8
(timeout 900 python -m pytest tests/test_multi_rank.py -m gpu -q -x 2>&1 | tail -30) > gpurun_out/r02_multi_rank_peer.txt
cat gpurun_out/r02_multi_rank_peer.txt
SECONDS=0
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 --no-sphere --no-punctures --no-affine > gpurun_out/r02_bench_n2_peer.json 2> gpurun_out/r02_bench_n2_peer.err
echo "rc=$? wall=${SECONDS}s"; tail -5 gpurun_out/r02_bench_n2_peer.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_n2_peer.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'halo', d['config']['halo'][:40])
print('parity', d.get('parity',{}).get('nccl_rel_l2'), [c.get('halo_rank0') for c in d.get('parity',{}).get('cases',[])])
print('ab', d.get('halo_comparison'))
print('vcycle', d.get('vcycle',{}).get('seconds_per_vcycle'))
PY
