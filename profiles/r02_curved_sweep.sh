# the shuffle kernel for curved conforming elements at p=3 on the level-5 sphere under different register caps
cd $GRAFT_REPO_ROOT
for lib in base s12 s20 s24; do
  if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_$lib.so; fi
  echo "$lib $(python profiles/hp_bench.py 5 3 uniform 2>&1 | tail -1)"
done
unset D4B_LIBRARY_PATH
python -m pytest tests/test_gpu_apply.py tests/test_sphere.py -q -x 2>&1 | tail -2
