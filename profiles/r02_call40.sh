cd $GRAFT_REPO_ROOT
for ns in 1 2; do for k in 16 32 48; do
echo -n "streams $ns chunks $k: "
D4B_PIPE_STREAMS=$ns D4B_PIPELINE_CHUNKS=$k python bench.py --steps 10 --warmup 3 --no-cpu --no-vcycle --no-sphere --no-affine --no-punctures --no-sweep 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('e2e %.3f GDOF/s  %.3f ms' % (d['e2e']['value'], d['config']['dof_per_gpu']/d['e2e']['value']/1e6))"
done; done
