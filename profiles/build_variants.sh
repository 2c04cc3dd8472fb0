#!/bin/bash
# build_variants.sh -- compile d4b_fast.cu with different launch geometries of the flat kernel and
# link each into profiles/variants/lib_<name>.so (picked up through D4B_LIBRARY_PATH by bench.py).
# usage: profiles/build_variants.sh name 'THREADS-list' 'MINB-list' [name ...]
# KERNEL=FAST selects the curved conforming kernel's tables (D4B_FAST_THREADS / D4B_FAST_MINB)
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"; SRC="$HERE/../disco4est_b200/csrc"; OUT="$HERE/variants"
mkdir -p "$OUT"
make -s -j8 -C "$SRC"
build() {
  local name=$1 thr=$2 minb=$3
  printf '#define D4B_%s_THREADS {%s}\n#define D4B_%s_MINB {%s}\n' "${KERNEL:-FLAT}" "$thr" "${KERNEL:-FLAT}" "$minb" > "$OUT/cfg_$name.h"
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC \
    -I"$SRC/../../include" -I"$SRC" -include "$OUT/cfg_$name.h" -c "$SRC/d4b_fast.cu" -o "$OUT/fast_$name.o"
  objs=$(ls "$SRC"/*.o | grep -v d4b_fast.o)
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT/lib_$name.so" "$OUT/fast_$name.o" $objs -lcudart -ldl
  rm -f "$OUT/fast_$name.o"
  echo "built $name"
}
while [ $# -ge 3 ]; do build "$1" "$2" "$3" & shift 3; done
wait
