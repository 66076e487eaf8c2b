#!/bin/bash
# round 2, GPU call M: the allocation-free entry point's test alone in a fresh process, then the whole GPU suite (-x, as
# the round's end runs it), smoke, and the default bench line of the final state.
set -u
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x -p no:cacheprovider > gpurun_out/r2m_pytest_no_alloc_alone.txt 2>&1
echo "no_alloc alone exit $?"; tail -n 3 gpurun_out/r2m_pytest_no_alloc_alone.txt
python -m pytest tests -m gpu -x -q -rs -p no:cacheprovider > gpurun_out/r2m_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2m_pytest_gpu.txt; tail -n 8 gpurun_out/r2m_pytest_gpu.txt
cp gpurun_out/variance_error.json gpurun_out/r2m_variance_error.json 2>/dev/null
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2m_smoke.txt 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/r2m_smoke.txt
python bench.py > gpurun_out/r2m_bench_headline.json 2> gpurun_out/r2m_bench_headline.err; echo "headline exit $?"; head -c 300 gpurun_out/r2m_bench_headline.json; echo
echo done
