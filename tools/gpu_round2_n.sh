#!/bin/bash
# round 2, GPU call N: recursive-doubling inverse of the factor (trtri_batched_i8): fit tests, then LML+gradient at
# R = 64 / 8 and the n = 16384 factorisation time
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_scale.py tests/test_gpu_i8.py -m gpu -q -x -p no:cacheprovider > gpurun_out/r2n_pytest_fit.txt 2>&1
echo "pytest fit exit $?"; tail -n 15 gpurun_out/r2n_pytest_fit.txt
python tools/fit_only.py 64 4 > gpurun_out/r2n_fit_R64.log 2>&1; tail -n 2 gpurun_out/r2n_fit_R64.log
python tools/fit_only.py 8 4 > gpurun_out/r2n_fit_R8.log 2>&1; tail -n 2 gpurun_out/r2n_fit_R8.log

echo done
