cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for f in 1 0; do
D4B_FUSED_SOLVERS=$f python profiles/vcycle_breakdown.py 6 > gpurun_out/r02_vcycle_breakdown_fused$f.json 2>&1
python - <<PY
import json
d=json.load(open('gpurun_out/r02_vcycle_breakdown_fused$f.json'))
for k,v in d.items(): print('$f', k, {a:round(b,3) for a,b in v.items()})
PY
done
D4B_AFFINE=0 python profiles/solver_kernels.py 6
D4B_AFFINE=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_solver_fused.csv python profiles/solver_kernels.py 6 > gpurun_out/ncu_solver.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r02_launches_solver_fused.csv')) if len(r)>10 and r[0].isdigit()]
# columns: ID, PID, Process, Host, Kernel Name, Context, Stream, Block Size, Grid Size, Device, CC, Section, Metric Name, Unit, Value
seq=[(r[4].split('(')[0][-60:], r[8], float(r[-1].replace(',',''))) for r in rows]
for name,grid,v in seq[-90:]: print('%-62s %-16s %10.1f' % (name, grid, v))
PY
