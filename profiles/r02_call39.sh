cd $GRAFT_REPO_ROOT
(timeout 600 python -m pytest tests/test_krylov.py -m gpu -q -x -k blas1 2>&1 | tail -3)
for k in 16 24 32; do
echo "chunks $k"
D4B_PIPELINE_CHUNKS=$k D4B_PIPE_TIMING=1 python bench.py --steps 5 --warmup 3 --no-cpu --no-vcycle --no-sphere --no-affine --no-punctures --no-sweep 2>&1 >/dev/null | grep "pipelined host" | tail -3
done
