#!/bin/bash
# build_variant.sh NAME "EXTRA NVCC FLAGS" -> variants/libppq_NAME.so (an experiment build of the product library)
set -e
cd "$(dirname "$0")/../ppqsflhe_b200/csrc"
mkdir -p ../../variants/obj_$1
O=../../variants/obj_$1
ARCH="-gencode arch=compute_100a,code=sm_100a"
for f in kernels ckks sampler capi; do
  nvcc $ARCH -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fopenmp,-O3 $2 -c $f.cu -o $O/$f.o &
done
g++ -O2 -std=c++17 -fPIC -c params.cpp -o $O/params.o
wait
nvcc $ARCH -shared -o ../../variants/libppq_$1.so $O/kernels.o $O/ckks.o $O/sampler.o $O/capi.o $O/params.o -Xcompiler -fopenmp -lcudart_static -lpthread -ldl -lrt
rm -rf $O
echo built variants/libppq_$1.so
