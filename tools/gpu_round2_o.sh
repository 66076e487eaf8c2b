#!/bin/bash
# round 2, GPU call O: diagonal-block kernel with a column-major shared-memory layout (16-byte pivot loads), row stride
# 132 vs 136 vs the row-major kernel (base): LML+gradient at R = 8 / 64, then the fit tests on the default
set -u
mkdir -p gpurun_out
R=$PWD
for lib in base ld136 new; do
  if [ $lib = new ]; then unset TB200_LIB; else export TB200_LIB=$R/turbo_b200/libturbo_b200_$lib.so; fi
  python tools/fit_only.py 8 5 > gpurun_out/r2o_fit_${lib}_R8.log 2>&1; echo "$lib R8"; tail -n 2 gpurun_out/r2o_fit_${lib}_R8.log
  python tools/fit_only.py 64 4 > gpurun_out/r2o_fit_${lib}_R64.log 2>&1; echo "$lib R64"; tail -n 2 gpurun_out/r2o_fit_${lib}_R64.log
done
unset TB200_LIB
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_scale.py -m gpu -q -x -p no:cacheprovider > gpurun_out/r2o_pytest_fit.txt 2>&1
echo "pytest fit exit $?"; tail -n 6 gpurun_out/r2o_pytest_fit.txt
python tools/fit_only.py 8 2 > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r2o_launches_fit_R8.csv python tools/fit_only.py 8 2 > gpurun_out/r2o_ncu.log 2>&1
echo "ncu exit $?"
echo done
