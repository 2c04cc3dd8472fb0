out=gpurun_out/variant_sweep2.txt; : > $out
for d in 1 2 4 5 6 7; do
  lvl=6; if [ $d -ge 4 ]; then lvl=5; fi
  for lib in base $(ls profiles/variants/lib_f*.so); do
    if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/$lib; fi
    ms=$(python bench.py --level $lvl --deg $d --steps 20 --warmup 5 --no-cpu --no-vcycle --no-sphere 2>/dev/null | tail -1 | python -c "import json,sys; j=json.loads(sys.stdin.readline()); print('%.4f ms %.2f GDOF/s frac %.3f' % (j['ms_per_step'], j['value'], j['roofline']['frac']))")
    echo "p=$d $(basename $lib) $ms" | tee -a $out
  done
done
