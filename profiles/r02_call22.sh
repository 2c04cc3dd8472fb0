cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
out=gpurun_out/r02_elem_ns_ab.txt; : > $out
for lib in base r02start base r02start; do
  if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_$lib.so; fi
  python profiles/apply_time.py 6:3 6:1 6:2 6:4 5:5 5:6 5:7 2>&1 | grep level | tee -a $out
done
unset D4B_LIBRARY_PATH
(timeout 1200 python -m pytest tests/test_gpu_apply.py tests/test_smoother.py tests/test_sphere.py tests/test_amr_brick.py -m gpu -q -x 2>&1 | tail -5)
