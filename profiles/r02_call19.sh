cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
out=gpurun_out/r02_flat_regression_check.txt; : > $out
for rep in 1 2; do
for lib in base r02start; do
  if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_$lib.so; fi
  python profiles/apply_time.py 6:3 6:1 6:2 6:4 6:5 5:6 5:7 2>&1 | grep level | tee -a $out
done
done
