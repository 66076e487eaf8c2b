#!/bin/bash
# round 2: ncu --set full of the persistent fit GEMM on LARGE launches (8 restarts, n = 4096): the K^-1 = Linv^T Linv
# product (last gemm_i8 launch of an evaluation) and the top level of the recursive inverse (the two before it)
set -u
mkdir -p gpurun_out
python tools/fit_only.py 8 2 > gpurun_out/r2ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_i8_kernel -s 137 -c 3 -o gpurun_out/r2_gemm_i8_large_full python tools/fit_only.py 8 2 > gpurun_out/r2ncu.log 2>&1
echo "ncu exit $?"; tail -n 3 gpurun_out/r2ncu.log
echo done
