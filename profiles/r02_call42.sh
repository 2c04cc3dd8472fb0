cd $GRAFT_REPO_ROOT
out=gpurun_out/r02_flat_256thread_ab2.txt; : > $out
for lib in base w256d w256e w256f; do
  if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_$lib.so; fi
  python profiles/apply_time.py 6:5 6:6 6:7 2>&1 | grep level | tee -a $out
done
