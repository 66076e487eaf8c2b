#!/bin/bash
# round 2: fit GEMM epilogue in two phases (drain all columns, release TMEM, then scale/merge/store) vs interleaved (base)
set -u
mkdir -p gpurun_out
R=$PWD
for lib in base new base new; do
  if [ $lib = new ]; then unset TB200_LIB; else export TB200_LIB=$R/turbo_b200/libturbo_b200_$lib.so; fi
  python tools/fit_only.py 64 4 > gpurun_out/r2ep_fit_${lib}_R64.log 2>&1; echo "$lib R64"; tail -n 2 gpurun_out/r2ep_fit_${lib}_R64.log
  python tools/fit_only.py 8 5 > gpurun_out/r2ep_fit_${lib}_R8.log 2>&1; echo "$lib R8"; tail -n 2 gpurun_out/r2ep_fit_${lib}_R8.log
done
unset TB200_LIB
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_scale.py tests/test_gpu_i8.py -m gpu -q -x -p no:cacheprovider > gpurun_out/r2ep_pytest_fit.txt 2>&1
echo "pytest fit exit $?"; tail -n 4 gpurun_out/r2ep_pytest_fit.txt
echo done
