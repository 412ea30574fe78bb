#!/bin/sh
# Build a variant of libnbgraft.so with extra -D flags for ONE source file (timing experiments):
#   sh tools/build_variant.sh NAME cull.cu -DNBG_CULL_UNROLL=2 -DNBG_CULL_MINBLOCKS=6
# -> build/variants/libnbgraft_NAME.so (use with NBGRAFT_LIB=...); ptxas log next to it.
set -e
cd "$(dirname "$0")/../odh-kubeflow_b200/csrc"
name=$1; src=$2; shift 2
out=../../build/variants
mkdir -p $out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-mpopcnt -Xptxas -v "$@" \
  -c cuda/$src -o $out/${name}_${src%.cu}.o 2> $out/${name}_${src%.cu}.ptxas.log
objs=""
for f in capi cull diff join join_noalloc store text validate; do
  if [ "$f.cu" = "$src" ]; then objs="$objs $out/${name}_${src%.cu}.o"; else objs="$objs cuda/$f.o"; fi
done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/libnbgraft_$name.so $objs -cudart static
grep -E "registers|spill" $out/${name}_${src%.cu}.ptxas.log | sort | uniq -c | head -8
