#!/bin/bash
# tools/libaz_attn_trace.so = the product library with az_net.cu compiled -DAZ_ATTN_TRACE (phase clocks of block 0 of
# az_attn_mma_kernel, printed by the kernel).  Run after `python __graft_entry__.py build`; use with
#   AZ_B200_LIB=tools/libaz_attn_trace.so python tools/attn_bench.py --child 7 256 2 5 4096
set -e
cd "$(dirname "$0")/../alpha-zero_b200/csrc"
mkdir -p ../../tools/_build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -I ../../include \
     -DAZ_ATTN_TRACE -c az_net.cu -o ../../tools/_build/az_net_attn_trace.o
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../tools/libaz_attn_trace.so ../../tools/_build/az_net_attn_trace.o az_forest.o az_selfplay.o \
     az_noalloc.o az_replay.o -lcudart_static -lpthread -ldl -lrt
