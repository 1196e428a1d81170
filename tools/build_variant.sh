#!/bin/bash
# build_variant.sh NAME FILE.cu "-DFLAG=.. ..." : compile one source with extra flags, link a tuning copy of the
# engine as variants/libb200stereo_NAME.so (select it with B200S_LIB=...). Cross-compiles without a GPU.
set -e
NAME=$1; SRC=$2; EXTRA=$3
cd "$(dirname "$0")/../depthestimation_b200/csrc"
mkdir -p ../../variants/obj
O=../../variants/obj/${SRC%.cu}_$NAME.o
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC --fmad=false $EXTRA -c $SRC -o $O
OBJS=""
for f in api cost paths wta imgops speckle rows vert; do
  if [ "$f.cu" == "$SRC" ]; then OBJS="$OBJS $O"; else OBJS="$OBJS $f.o"; fi
done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../variants/libb200stereo_$NAME.so $OBJS
echo built variants/libb200stereo_$NAME.so
