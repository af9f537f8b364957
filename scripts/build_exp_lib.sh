#!/bin/sh
# Development aid: the library with the K3 experiment modes compiled in (BSGL_EXT_EXPMODE selects one at run time),
# written to scripts/build/libbsgl_exp.so.  Not the product library.
set -e
cd "$(dirname "$0")/.."
mkdir -p scripts/build/obj
for f in baby_sglang_b200/csrc/*.cu; do
  o=scripts/build/obj/$(basename "$f" .cu).o
  extra=""
  [ "$(basename "$f")" = extend_attention.cu ] && extra="-DBSGL_EXT_EXPERIMENTS"
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $extra -I include -c "$f" -o "$o" &
done
wait
nvcc -shared -o scripts/build/libbsgl_exp.so scripts/build/obj/*.o -gencode arch=compute_100a,code=sm_100a
echo scripts/build/libbsgl_exp.so
