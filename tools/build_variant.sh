#!/bin/bash
# Builds a variant of libhppb200.so with extra -D flags into hpp-core_b200/csrc/build/variants/<name>/libhppb200.so (for A/B runs via HPPB200_LIB).
# usage: tools/build_variant.sh NAME "-DFOO=1 -DBAR=2" [file.cu ...]   (listed files are recompiled, the rest reuse build/*.o)
set -e
cd "$(dirname "$0")/../hpp-core_b200/csrc"
name=$1; flags=$2; shift 2
out=build/variants/$name; mkdir -p $out
NV="/usr/local/cuda/bin/nvcc $flags -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC -Xcompiler -ffp-contract=off -ccbin /usr/bin/g++"
objs=""
for o in build/*.o; do
  b=$(basename $o .o); use=$o
  for f in "$@"; do if [ "$(basename $f .cu)" = "$b" ]; then $NV -c $f -o $out/$b.o; use=$out/$b.o; fi; done
  objs="$objs $use"
done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/libhppb200.so $objs -lcudart_static -lpthread -ldl -lrt -ccbin /usr/bin/g++
echo built $out/libhppb200.so
