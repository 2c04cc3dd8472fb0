cd $GRAFT_REPO_ROOT
for k in 2 4 8 16; do
echo -n "32^3 chunks $k: "
D4B_PIPELINE_CHUNKS=$k python bench.py --level 5 --steps 10 --warmup 3 --no-cpu --no-vcycle --no-sphere --no-affine --no-punctures --no-sweep 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('e2e %.3f GDOF/s  %.3f ms' % (d['e2e']['value'], d['config']['dof_per_gpu']/d['e2e']['value']/1e6))"
done
D4B_PIPELINE_MIN_ELEMS=0 python bench.py --level 5 --steps 10 --warmup 3 --no-cpu --no-vcycle --no-sphere --no-affine --no-punctures --no-sweep 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('32^3 unpipelined: e2e %.3f GDOF/s  %.3f ms' % (d['e2e']['value'], d['config']['dof_per_gpu']/d['e2e']['value']/1e6))"
