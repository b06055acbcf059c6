#!/bin/bash
# full rebuild of the library with extra -D flags (for switches that change shared structs): tools/build_full_variant.sh <name> "<flags>"
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
PKG="$ROOT/tensorhive-fixed_b200"
OUT="$PKG/lib/variants/full_$1"
mkdir -p "$OUT"
for f in "$PKG"/csrc/*.cu; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -DTH_BUILD $2 -c "$f" -o "$OUT/$(basename "$f" .cu).o" &
done
wait
nvcc -shared -o "$PKG/lib/variants/libth_b200_$1.so" "$OUT"/*.o -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -lcudart
echo "built $1 ($2)"
