#!/bin/bash
# Builds libaz_b200.so (product) and tools/_build/libaz_trace.so (-DAZ_FRONT_TRACE, for tools/front_trace.py) side by side.
set -e
cd "$(dirname "$0")/../alpha-zero_b200/csrc"
F="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -I ../../include"
L="-lcudart_static -lpthread -ldl -lrt"
mkdir -p ../../tools/_build
( nvcc $F -DAZ_FRONT_TRACE -c az_net.cu -o ../../tools/_build/az_net_trace.o && nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../tools/_build/libaz_trace.so ../../tools/_build/az_net_trace.o az_forest.o $L ) &
nvcc $F -c az_net.cu -o az_net.o && nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libaz_b200.so az_forest.o az_net.o $L
wait
