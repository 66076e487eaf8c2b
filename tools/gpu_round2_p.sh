#!/bin/bash
# round 2, GPU call P: potrf's main-stream INT8 products leave the SM after 0 (= fully persistent) / 2 / 4 / 8 tiles, so
# that the high-priority diagonal-block kernel of the next column gets SMs while they run: LML+gradient at R = 8 / 64;
# then the whole GPU suite on the default (4)
set -u
mkdir -p gpurun_out
R=$PWD
for lib in tpc0 tpc2 new tpc8; do
  if [ $lib = new ]; then unset TB200_LIB; else export TB200_LIB=$R/turbo_b200/libturbo_b200_$lib.so; fi
  python tools/fit_only.py 8 5 > gpurun_out/r2p_fit_${lib}_R8.log 2>&1; echo "$lib R8"; tail -n 2 gpurun_out/r2p_fit_${lib}_R8.log
  python tools/fit_only.py 64 4 > gpurun_out/r2p_fit_${lib}_R64.log 2>&1; echo "$lib R64"; tail -n 2 gpurun_out/r2p_fit_${lib}_R64.log
done
unset TB200_LIB
python -m pytest tests -m gpu -x -q -rs -p no:cacheprovider > gpurun_out/r2p_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2p_pytest_gpu.txt; tail -n 8 gpurun_out/r2p_pytest_gpu.txt
echo done
