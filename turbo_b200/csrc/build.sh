#!/bin/bash
# Build libturbo_b200.so in-tree for sm_100a.  Usage: turbo_b200/csrc/build.sh [extra nvcc flags]
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
# TB200_MBAR_TRAP_SECONDS: wall-clock bound of one mbarrier wait (ozaki_i8.cuh); -UTB200_MBAR_TRAP_SECONDS removes it
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -DTB200_MBAR_TRAP_SECONDS=30"
OUT=${TB200_OUT:-../libturbo_b200.so}   # TB200_OUT: A/B builds next to the shipped library (loaded with TB200_LIB)
$NVCC $FLAGS "$@" -shared -o "$OUT" fit.cu acquire.cu
