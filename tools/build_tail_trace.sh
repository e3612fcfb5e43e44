#!/bin/bash
# tools/libaz_tail_trace.so = the product library with az_net.cu compiled -DAZ_TAIL_TRACE (phase clocks of block 0 of
# az_tail_kernel, printed by the kernel).  Run after `python __graft_entry__.py build`; use with
#   AZ_B200_LIB=tools/libaz_tail_trace.so python tools/tail_bench.py --child connect4 1792
set -e
cd "$(dirname "$0")/../alpha-zero_b200/csrc"
mkdir -p ../../tools/_build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -I ../../include \
     -DAZ_TAIL_TRACE -c az_net.cu -o ../../tools/_build/az_net_tail_trace.o
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../tools/libaz_tail_trace.so ../../tools/_build/az_net_tail_trace.o az_forest.o az_selfplay.o \
     az_noalloc.o az_replay.o -lcudart_static -lpthread -ldl -lrt
