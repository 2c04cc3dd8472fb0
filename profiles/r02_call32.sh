cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 2400 python -m pytest tests -m gpu -q -x 2>&1 | tail -5) | tee gpurun_out/r02_gpu_suite_final.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | tee gpurun_out/r02_smoke_final.txt
SECONDS=0
python bench.py > gpurun_out/r02_bench_n1_final.json 2> gpurun_out/r02_bench_n1_final.err
echo "bench rc=$? wall=${SECONDS}s"
tail -2 gpurun_out/r02_bench_n1_final.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_n1_final.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'frac', d['roofline']['frac'], d['roofline']['frac_true_bytes'], 'e2e', d['e2e']['value'])
print('cpu', d['cpu_baseline']['value'], d.get('e2e_apply_lhs',{}).get('e2e_GDOF_per_s'), d.get('e2e_apply_lhs',{}).get('parity_rel_l2_vs_reference_cpu_apply'))
print('vcycle', d['vcycle']['seconds_per_vcycle'], d['vcycle'].get('cuda_graph_replays_in_timed_cycles'))
a=d.get('affine_fast_path',{}); print('affine', a.get('GDOF_per_s'), a.get('vcycle',{}).get('seconds_per_vcycle'))
print('sphere', {k:round(v['GDOF_per_s'],2) for k,v in d['cubed_sphere'].items() if isinstance(v,dict)})
print('sweep', [(x['p'], round(x['GDOF_per_s'],1), round(x['frac'],3)) for x in d['degree_sweep']])
print('punct', d['two_punctures']['jacobian_apply']['GDOF_per_s'], d['two_punctures']['newton_solve']['seconds'], d['two_punctures']['newton_solve']['fcg_iterations_per_step'])
print('clocks', d['clocks'])
PY
SECONDS=0
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_ref.json 2> gpurun_out/r02_bench_ref.err
echo "ref rc=$? wall=${SECONDS}s"; tail -1 gpurun_out/r02_bench_ref.json | cut -c1-400
