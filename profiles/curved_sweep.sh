#!/bin/bash
# curved_sweep.sh -- time the curved conforming kernel (uniform-p cubed sphere) with every library variant
# of profiles/variants/ and with / without the L2 prefetch.  usage: profiles/curved_sweep.sh [level] [deg]
lvl=${1:-5}; deg=${2:-3}
for lib in base $(ls profiles/variants/lib_*.so 2>/dev/null); do
  for pf in 1 0; do
    if [ "$lib" = base ]; then unset D4B_LIBRARY_PATH; else export D4B_LIBRARY_PATH=$PWD/$lib; fi
    echo "$(basename $lib) prefetch=$pf $(D4B_PREFETCH=$pf python profiles/hp_bench.py $lvl $deg uniform 2>&1 | tail -1)"
  done
done
