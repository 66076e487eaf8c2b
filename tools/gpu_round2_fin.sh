#!/bin/bash
# round 2, final GPU call: GPU suite (-x, as the round's end runs it), smoke, the default bench line and the reference arm
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs -p no:cacheprovider > gpurun_out/r2fin_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2fin_pytest_gpu.txt; tail -n 8 gpurun_out/r2fin_pytest_gpu.txt
cp gpurun_out/variance_error.json gpurun_out/r2fin_variance_error.json 2>/dev/null
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2fin_smoke.txt 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/r2fin_smoke.txt
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2fin_bench_reference.json 2> gpurun_out/r2fin_bench_reference.err; echo "reference exit $?"
python bench.py > gpurun_out/r2fin_bench_headline.json 2> gpurun_out/r2fin_bench_headline.err; echo "headline exit $?"; head -c 300 gpurun_out/r2fin_bench_headline.json; echo
echo done
