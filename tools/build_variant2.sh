#!/bin/bash
# build_variant2.sh NAME "-DFLAG=.. ..." FILE.cu [FILE.cu ...] : compile the listed sources with extra flags (each as ONE
# translation unit: without its *_PART macro a source holds all its instantiations), link a tuning copy of the engine as
# variants/libb200stereo_NAME.so (select it with B200S_LIB=...). Cross-compiles without a GPU.
set -e
NAME=$1; EXTRA=$2; shift 2
cd "$(dirname "$0")/../depthestimation_b200/csrc"
mkdir -p ../../variants/obj
OBJS=""
for f in api paths wta imgops speckle cost rows vert; do
  hit=0
  for s in "$@"; do if [ "$f.cu" == "$s" ]; then hit=1; fi; done
  if [ $hit == 1 ]; then
    O=../../variants/obj/${f}_$NAME.o
    /usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC --fmad=false $EXTRA -c $f.cu -o $O 2>/dev/null &
    OBJS="$OBJS $O"
  else
    case $f in
      cost) OBJS="$OBJS $(ls cost_p*.o | tr '\n' ' ')" ;;
      rows) OBJS="$OBJS $(ls rows_p*.o | tr '\n' ' ')" ;;
      vert) OBJS="$OBJS $(ls vert_p*.o | tr '\n' ' ')" ;;
      *) OBJS="$OBJS $f.o" ;;
    esac
  fi
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../variants/libb200stereo_$NAME.so $OBJS
echo built variants/libb200stereo_$NAME.so
