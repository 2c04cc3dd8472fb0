cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
out=gpurun_out/r02_flat_stream_ab.txt; : > $out
for lib in base r02start streamg streamgrec; do
  if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_$lib.so; fi
  python profiles/apply_time.py 6:3 6:1 6:2 5:5 5:7 6:3 2>&1 | grep level | tee -a $out
done
unset D4B_LIBRARY_PATH
(timeout 1200 python -m pytest tests/test_smoother.py tests/test_multigrid.py tests/test_krylov.py tests/test_newton.py -m gpu -q -x 2>&1 | tail -5)
python bench.py --steps 20 --warmup 5 --no-sphere --no-punctures --no-sweep --no-cpu > gpurun_out/r02_bench_mid2.json 2> gpurun_out/r02_bench_mid2.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_mid2.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'])
print('vcycle', d.get('vcycle',{}).get('seconds_per_vcycle'))
a=d.get('affine_fast_path',{}); print('affine', a.get('GDOF_per_s'), a.get('vcycle',{}).get('seconds_per_vcycle'))
PY
