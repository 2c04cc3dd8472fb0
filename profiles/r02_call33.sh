cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu --no-vcycle --no-sphere --no-affine --no-punctures --no-sweep > gpurun_out/b_small.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench_p3_64cube_final.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-vcycle --no-sphere --no-affine --no-punctures --no-sweep > gpurun_out/ncu_b.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r02_launches_bench_p3_64cube_final.csv')) if len(r)>10 and r[0].isdigit()]
for r in rows[-14:]: print(r[4].split('(')[0][-50:], r[8], r[-1])
PY
python profiles/apply_once.py 6 3 > /dev/null && \
ncu --set full --clock-control none --import-source on -k regex:flat_apply_kernel -s 2 -c 1 -o gpurun_out/r02_flat_p3_final -f python profiles/apply_once.py 6 3 > gpurun_out/ncu_flat2.log 2>&1
tail -2 gpurun_out/ncu_flat2.log
python profiles/solver_kernels.py 6 > /dev/null && D4B_AFFINE=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_solver_final.csv python profiles/solver_kernels.py 6 > gpurun_out/ncu_solver2.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r02_launches_solver_final.csv')) if len(r)>10 and r[0].isdigit()]
for r in rows[-40:]: print('%-52s %-16s %s' % (r[4].split('(')[0][-50:], r[8], r[-1]))
PY
