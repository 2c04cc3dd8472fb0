cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L | wc -l
(timeout 900 python -m pytest tests/test_multi_rank.py -m gpu -q -v 2>&1 | tail -12) > gpurun_out/r02_multi_rank_2gpu.txt
cat gpurun_out/r02_multi_rank_2gpu.txt
SECONDS=0
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_n2.json 2> gpurun_out/r02_bench_n2.err
echo "rc=$? wall=${SECONDS}s"; tail -3 gpurun_out/r02_bench_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_n2.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'parity', d['parity']['nccl_rel_l2'], 'warmup', d['warmup'])
print('vcycle', d['vcycle']['seconds_per_vcycle']); a=d['affine_fast_path']; print('affine', a['GDOF_per_s'], a.get('vcycle',{}).get('seconds_per_vcycle'))
print('sphere', {k:round(v['GDOF_per_s'],2) for k,v in d['cubed_sphere'].items() if isinstance(v,dict)})
print('punct', d['two_punctures']['jacobian_apply'])
PY
