cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_multi_rank.py -m gpu -q -x 2>&1 | tail -30) > gpurun_out/r02_multi_rank_peer2.txt
cat gpurun_out/r02_multi_rank_peer2.txt
for ar in 1 0; do
SECONDS=0
D4B_PEER_ALLREDUCE=$ar timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 --no-sphere --no-punctures --no-affine > gpurun_out/r02_bench_n2_ar$ar.json 2> gpurun_out/r02_bench_n2_ar$ar.err
echo "ar=$ar rc=$? wall=${SECONDS}s"; tail -3 gpurun_out/r02_bench_n2_ar$ar.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_n2_ar$ar.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'halo', d['config']['halo'][:20])
print('parity', d.get('parity',{}).get('nccl_rel_l2'))
print('ab', d.get('halo_comparison'))
print('vcycle', d.get('vcycle',{}).get('seconds_per_vcycle'), d.get('vcycle',{}).get('residual_reduction_per_vcycle'))
PY
done
