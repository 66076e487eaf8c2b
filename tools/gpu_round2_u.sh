#!/bin/bash
# round 2, GPU call U: three staged K* CTAs per SM: GPU suite, default bench line
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs -p no:cacheprovider > gpurun_out/r2u_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2u_pytest_gpu.txt; tail -n 8 gpurun_out/r2u_pytest_gpu.txt
cp gpurun_out/variance_error.json gpurun_out/r2u_variance_error.json 2>/dev/null
python bench.py > gpurun_out/r2u_bench_headline.json 2> gpurun_out/r2u_bench_headline.err; echo "headline exit $?"; head -c 300 gpurun_out/r2u_bench_headline.json; echo
echo done
