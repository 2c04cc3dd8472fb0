#!/bin/bash
# build_flag_variants.sh -- compile d4b_fast.cu with extra -D flags and link each into
# profiles/variants/lib_<name>.so (picked up through D4B_LIBRARY_PATH).
# usage: profiles/build_flag_variants.sh name '-DFLAG ...' [name flags ...]
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"; SRC="$HERE/../disco4est_b200/csrc"; OUT="$HERE/variants"
mkdir -p "$OUT"
make -s -j8 -C "$SRC"
build() {
  local name=$1 flags=$2
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fopenmp \
    -I"$SRC/../../include" -I"$SRC" $flags -c "$SRC/d4b_fast.cu" -o "$OUT/fast_$name.o"
  objs=$(ls "$SRC"/*.o | grep -v d4b_fast.o)
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT/lib_$name.so" "$OUT/fast_$name.o" $objs -lcudart -ldl -lgomp
  rm -f "$OUT/fast_$name.o"
  echo "built $name"
}
while [ $# -ge 2 ]; do build "$1" "$2" & shift 2; done
wait
