#!/bin/bash
# round 2, GPU call S: GPU suite after the model factor went back to 28 products in its inverse; fit timings; 1-GPU lines
# of cfg2 and cfg4 on the final kernels
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs -p no:cacheprovider > gpurun_out/r2s_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2s_pytest_gpu.txt; tail -n 8 gpurun_out/r2s_pytest_gpu.txt
cp gpurun_out/variance_error.json gpurun_out/r2s_variance_error.json 2>/dev/null
for R in 64 8; do python tools/fit_only.py $R 4 > gpurun_out/r2s_fit_only_R$R.log 2>&1; tail -n 2 gpurun_out/r2s_fit_only_R$R.log; done
python bench.py --config cfg2 > gpurun_out/r2s_bench_cfg2.json 2> gpurun_out/r2s_bench_cfg2.err; echo "cfg2 exit $?"; head -c 250 gpurun_out/r2s_bench_cfg2.json; echo
python bench.py --config cfg4 --steps 3 > gpurun_out/r2s_bench_cfg4.json 2> gpurun_out/r2s_bench_cfg4.err; echo "cfg4 exit $?"; head -c 250 gpurun_out/r2s_bench_cfg4.json; echo
echo done
