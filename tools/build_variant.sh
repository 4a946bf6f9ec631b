#!/bin/bash
# usage: build_variant.sh <name> "<nvcc defines>"   -> tools/_tmp/libpcp_<name>.so (same sources, extra -D switches for knn_tile5.cu / search_pca.cu)
set -e
cd "$(dirname "$0")/../pointcloudprocessing_b200/csrc"
NAME=$1; DEFS=$2
OUT=../../tools/_tmp; mkdir -p $OUT/obj_$NAME
NV="/usr/local/cuda/bin/nvcc $DEFS -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -ccbin /usr/bin/g++ -Xcompiler -fPIC,-fvisibility=hidden -I../../include -I. --expt-relaxed-constexpr"
OBJS=""
for f in *.cu; do
  o=${f%.cu}.o
  if [ "$f" = knn_tile6.cu ] || [ "$f" = search_pca.cu ]; then $NV -c $f -o $OUT/obj_$NAME/$o & OBJS="$OBJS $OUT/obj_$NAME/$o"; else OBJS="$OBJS $o"; fi
done
wait
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -ccbin /usr/bin/g++ -o $OUT/libpcp_$NAME.so $OBJS -lcudart_static -lpthread -ldl -lrt
echo built $OUT/libpcp_$NAME.so
