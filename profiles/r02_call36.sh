cd $GRAFT_REPO_ROOT
out=gpurun_out/r02_curved_hoist_ab2.txt; : > $out
for lib in base prevhead minb12 minb14 base prevhead minb12 minb14; do
  if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_$lib.so; fi
  python profiles/apply_sphere_once.py 5 3 2>&1 | tail -1 | tee -a $out
done
