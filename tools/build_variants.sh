#!/bin/bash
# Experiment builds of libth_b200.so: same sources, different -D flags for th_search.cu (K2).  Usage:
#   tools/build_variants.sh name1 "-DFLAG ..." name2 "..."   ->  tensorhive-fixed_b200/lib/variants/libth_b200_<name>.so
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
PKG="$ROOT/tensorhive-fixed_b200"
OUT="$PKG/lib/variants"
mkdir -p "$OUT"
python "$ROOT/__graft_entry__.py" >/dev/null
OTHERS=$(ls "$PKG"/lib/obj/*.o | grep -v th_search.o)
while [ $# -gt 1 ]; do
  name=$1; flags=$2; shift 2
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -DTH_BUILD $flags \
       -c "$PKG/csrc/th_search.cu" -o "$OUT/th_search_$name.o"
  nvcc -shared -o "$OUT/libth_b200_$name.so" "$OUT/th_search_$name.o" $OTHERS -gencode arch=compute_100a,code=sm_100a \
       -Xcompiler -fPIC -lcudart
  echo "built $name ($flags)"
done
