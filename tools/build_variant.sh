#!/bin/bash
# usage: tools/build_variant.sh <name> <source-stem> <-D flags...>   ->  lz77_sss_b200/_variants/lib_<name>.so
# (the library with ONE translation unit compiled with extra -D flags; for A/B runs on the GPU box)
NAME=$1; STEM=$2; shift 2
OBJ=lz77_sss_b200/csrc/_obj
nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden "$@" \
  -c lz77_sss_b200/csrc/$STEM.cu -o /tmp/${STEM}_$NAME.o || exit 1
OBJS=$(ls $OBJ/*.o | grep -v "/$STEM.o")
nvcc -shared -o lz77_sss_b200/_variants/lib_$NAME.so $OBJS /tmp/${STEM}_$NAME.o -gencode arch=compute_100a,code=sm_100a -lcudart_static -lpthread -ldl -lrt
