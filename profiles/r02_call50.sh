cd $GRAFT_REPO_ROOT
out=gpurun_out/r02_curved_early_ab.txt; : > $out
for lib in base noearly base noearly; do
  if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_$lib.so; fi
  python profiles/apply_sphere_once.py 5 3 2>&1 | tail -1 | tee -a $out
done
for lib in base noearly; do
  if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_$lib.so; fi
  for p in 1 2 4; do python profiles/apply_sphere_once.py 4 $p 2>&1 | tail -1 | tee -a $out; done
done
unset D4B_LIBRARY_PATH
(timeout 1200 python -m pytest tests/test_gpu_apply.py tests/test_sphere.py tests/test_smoother.py tests/test_newton.py -m gpu -q -x 2>&1 | tail -3)
