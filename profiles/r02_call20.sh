cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
out=gpurun_out/r02_flat_regression_bisect.txt; : > $out
for lib in base r02start novec nodot novecnodot; do
  if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_$lib.so; fi
  python profiles/apply_time.py 6:3 6:1 6:3 2>&1 | grep level | tee -a $out
done
