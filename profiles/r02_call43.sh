cd $GRAFT_REPO_ROOT
out=gpurun_out/r02_curved_256thread_ab.txt; : > $out
for lib in base c256a c256b; do
  if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_$lib.so; fi
  for p in 5 6 7; do python profiles/apply_sphere_once.py 4 $p 2>&1 | tail -1 | tee -a $out; done
done
