cd $GRAFT_REPO_ROOT
out=gpurun_out/r02_flat_minb_ab.txt; : > $out
for lib in base m20 m27 t64m12 low base; do
  if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_$lib.so; fi
  python profiles/apply_time.py 6:3 6:1 6:2 2>&1 | grep level | tee -a $out
done
