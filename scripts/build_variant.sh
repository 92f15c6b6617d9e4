#!/bin/bash
# usage: scripts/build_variant.sh NAME "-DTA_NX_=8 -DTA_NB_=5"  ->  variants/libgnn_b200_NAME.so (tuning experiments)
set -e
cd "$(dirname "$0")/../gnn_b200/csrc"
make -s -j8 >/dev/null
mkdir -p ../../variants/obj_$1
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC $2 -c ${3:-linear_tc.cu} -o ../../variants/obj_$1/var.o
OBJS=$(ls build/*.o | grep -v "build/$(basename ${3:-linear_tc.cu} .cu).o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../variants/libgnn_b200_$1.so $OBJS ../../variants/obj_$1/var.o -lcudart_static
echo built variants/libgnn_b200_$1.so
