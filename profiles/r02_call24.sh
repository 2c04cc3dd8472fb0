cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python profiles/apply_time.py 6:3 6:1 6:2 2>&1 | grep level | tee gpurun_out/r02_apply_times_linear.txt
(timeout 2400 python -m pytest tests -m gpu -q -x 2>&1 | tail -5) | tee gpurun_out/r02_gpu_suite_mid2.txt
python bench.py --steps 20 --warmup 5 --no-sweep --no-cpu > gpurun_out/r02_bench_mid3.json 2> gpurun_out/r02_bench_mid3.err
tail -2 gpurun_out/r02_bench_mid3.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_mid3.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'frac', d['roofline']['frac'], 'e2e', d['e2e']['value'])
print('vcycle', d.get('vcycle',{}).get('seconds_per_vcycle'))
a=d.get('affine_fast_path',{}); print('affine', a.get('GDOF_per_s'), a.get('vcycle',{}).get('seconds_per_vcycle'))
print('sphere', {k:round(v['GDOF_per_s'],2) for k,v in d['cubed_sphere'].items() if isinstance(v,dict)})
print('punct', d['two_punctures']['jacobian_apply']['GDOF_per_s'], d['two_punctures']['residual']['GDOF_per_s'], d['two_punctures'].get('newton_solve',{}).get('seconds'), d['two_punctures'].get('newton_solve',{}).get('fcg_iterations_per_step'))
PY
