#!/bin/bash
# round 2, GPU call A: whole GPU test suite, the new bench configurations on one GPU, the predict precision study, and
# ncu --set full of the non-GEMM fit kernels.  Run under gpurun from the repo root.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2a_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2a_pytest_gpu.txt
tail -5 gpurun_out/r2a_pytest_gpu.txt
python bench.py > gpurun_out/r2a_bench_headline.json 2> gpurun_out/r2a_bench_headline.err; echo "headline exit $?"
python bench.py --config cfg2 > gpurun_out/r2a_bench_cfg2.json 2> gpurun_out/r2a_bench_cfg2.err; echo "cfg2 exit $?"
python bench.py --config cfg4 --steps 3 > gpurun_out/r2a_bench_cfg4.json 2> gpurun_out/r2a_bench_cfg4.err; echo "cfg4 exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2a_bench_reference.json 2> gpurun_out/r2a_bench_reference.err; echo "ref exit $?"
python tools/predict_precision.py --cases headline,cfg2,illcond,n8200 > gpurun_out/r2a_predict_precision.md 2>&1; echo "precision exit $?"
python tools/fit_only.py 8 2 > gpurun_out/r2a_fit_only.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'gram_train_kernel|diag_potrf_inv_kernel|slice_operand_kernel' -s 3 -c 5 \
    -o gpurun_out/r2a_fit_small_kernels python tools/fit_only.py 8 2 > gpurun_out/r2a_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'lml_grad_kernel' -c 1 \
    -o gpurun_out/r2a_lml_grad python tools/fit_only.py 8 2 > gpurun_out/r2a_ncu2.log 2>&1
echo done
