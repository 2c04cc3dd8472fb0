cd $GRAFT_REPO_ROOT
for lib in base t18 t20 t24; do
  if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_$lib.so; fi
  echo "$lib $(python profiles/hp_bench.py 5 3 uniform 2>&1 | tail -1)"
done
unset D4B_LIBRARY_PATH
(timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -8) > gpurun_out/r02_gpu_suite_1gpu.txt
cat gpurun_out/r02_gpu_suite_1gpu.txt
SECONDS=0; timeout 1200 python bench.py > gpurun_out/r02_bench_n1_default.json 2> gpurun_out/r02_bench_n1_default.err
echo "rc=$? wall=${SECONDS}s"; tail -c 1500 gpurun_out/r02_bench_n1_default.json
