python -m pytest tests/test_gpu_apply.py tests/test_sphere.py tests/test_smoother.py -m gpu -x -q 2>&1 | tail -2
for cfg in "6 1" "6 2" "5 4" "5 5"; do set -- $cfg; python bench.py --level $1 --deg $2 --steps 20 --warmup 5 --no-cpu --no-vcycle --no-sphere 2>/dev/null | tail -1 | python -c "import json,sys; j=json.loads(sys.stdin.readline()); print('flat p=$2 %.4f ms %.2f GDOF/s frac %.3f' % (j['ms_per_step'], j['value'], j['roofline']['frac']))"; done
python profiles/hp_bench.py 5 2 uniform | tail -1
python profiles/hp_bench.py 4 5 uniform | tail -1
python profiles/hp_bench.py 5 1 uniform | tail -1
