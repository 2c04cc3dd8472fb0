cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python profiles/apply_time.py 6:3 6:1 6:2 5:5 2>&1 | grep level
(timeout 900 python -m pytest tests/test_multi_rank.py -m gpu -q -x 2>&1 | tail -30) > gpurun_out/r02_multi_rank_merged.txt
tail -5 gpurun_out/r02_multi_rank_merged.txt
for mg in 1 0; do
D4B_PEER_MERGED=$mg timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 40 --warmup 5 --no-sphere --no-punctures --no-affine > gpurun_out/r02_bench_n2_merged$mg.json 2> gpurun_out/r02_bench_n2_merged$mg.err
echo "merged=$mg rc=$?"; tail -2 gpurun_out/r02_bench_n2_merged$mg.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_n2_merged$mg.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])
print('parity', d.get('parity',{}).get('nccl_rel_l2'))
print('ab', d.get('halo_comparison'))
print('vcycle', d.get('vcycle',{}).get('seconds_per_vcycle'))
PY
done
