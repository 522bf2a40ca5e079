#!/bin/bash
# Build a variant of libbert.so that differs from the default build in the compile flags of ONE source file
# (A/B and bisect runs; pick it at run time with BERT_B200_LIB).  Replaces the round-1 tools/gpu_ab_r1*.sh scripts.
#   tools/build_variant.sh <name> <source.cu> <extra nvcc flags...>   ->  embeddings.cpp_b200/lib/variants/libbert_<name>.so
set -e
name=$1; src=$2; shift 2
here=$(cd "$(dirname "$0")/.." && pwd)
pk=$here/embeddings.cpp_b200
python "$pk/build.py" > /dev/null                      # default objects up to date
mkdir -p "$pk/build/variants" "$pk/lib/variants"
obj=$pk/build/variants/${name}_${src%.*}.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 \
    -Xcompiler -fPIC,-fvisibility=hidden,-pthread -I "$here/include" "$@" -c "$pk/csrc/$src" -o "$obj"
objs=""
for o in "$pk"/build/*.o; do
    if [ "$(basename "$o")" = "${src%.*}.o" ]; then objs="$objs $obj"; else objs="$objs $o"; fi
done
/usr/local/cuda/bin/nvcc -shared -cudart static -o "$pk/lib/variants/libbert_$name.so" $objs -lpthread -ldl -lrt
echo "$pk/lib/variants/libbert_$name.so"
