#!/bin/bash
# round 2: ncu --set full of the staged K* kernel in its final configuration (128-point chunks, 3 CTAs per SM)
set -u
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --m 32768 --fit 0 --cpu-sample 0 --parity-n 0"
$B > gpurun_out/r2k2_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:kstar_mean_i8_staged -s 20 -c 1 -o gpurun_out/r2_kstar_full $B > gpurun_out/r2k2_ncu.log 2>&1
echo "ncu exit $?"; tail -n 2 gpurun_out/r2k2_ncu.log
echo done
