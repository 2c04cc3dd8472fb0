cd $GRAFT_REPO_ROOT
SECONDS=0; timeout 1500 python bench.py > gpurun_out/r02_bench_n1_default.json 2> gpurun_out/r02_bench_n1_default.err
echo "rc=$? wall=${SECONDS}s"; tail -3 gpurun_out/r02_bench_n1_default.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_n1_default.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], d['roofline']['frac_true_bytes'])
print('setup', d['config'].get('setup_s'))
print('sweep', [(x['p'], x['elements_per_direction'], round(x['GDOF_per_s'],1), round(x['frac'],3), round(x['setup_s'],2)) for x in d['degree_sweep']])
print('vcycle', d['vcycle']['seconds_per_vcycle'])
a=d['affine_fast_path']; print('affine', a['GDOF_per_s'], a['ms_per_apply'], a.get('vcycle',{}).get('seconds_per_vcycle'))
print('sphere', {k:(round(v['GDOF_per_s'],2), v['setup_s']['device_geometry_and_mesh']) for k,v in d['cubed_sphere'].items() if isinstance(v,dict)})
p=d['two_punctures']; print('punct', p['jacobian_apply'], p['residual']); 
for k in ('newton_solve','newton_solve_textbook'): print(k, p[k]['seconds'], p[k]['newton_steps'], p[k]['fcg_iterations_per_step'], p[k]['ierr'])
print('plugin', d.get('e2e_apply_lhs')); print('cpu', d['cpu_baseline'])
PY
