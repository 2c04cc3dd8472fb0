#!/bin/bash
# curved conforming kernel at every degree with every library variant of profiles/variants/
for cfg in "5 1" "5 2" "4 4" "4 5" "4 6" "4 7"; do
  set -- $cfg
  for lib in base $(ls profiles/variants/lib_c*.so 2>/dev/null); do
    if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/$lib; fi
    echo "$(basename $lib) $(python profiles/hp_bench.py $1 $2 uniform 2>&1 | tail -1)"
  done
done
