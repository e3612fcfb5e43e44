#!/bin/bash
# tools/libaz_trace.so = the product library with az_forest.cu compiled -DAZ_ADV_TRACE (per-phase clocks of az_advance_kernel,
# read by tools/adv_trace.py).  Run after `python __graft_entry__.py build` (reuses the other units' objects).
set -e
cd "$(dirname "$0")/../alpha-zero_b200/csrc"
mkdir -p ../../tools/_build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -I ../../include \
     -fmad=false -DAZ_ADV_TRACE -c az_forest.cu -o ../../tools/_build/az_forest_trace.o
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../tools/libaz_trace.so ../../tools/_build/az_forest_trace.o az_net.o az_selfplay.o \
     az_noalloc.o az_replay.o -lcudart_static -lpthread -ldl -lrt
