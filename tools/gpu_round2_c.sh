#!/bin/bash
# round 2, GPU call C: the default build after the A/B (per-model slice cut-off, bank-conflict-free Gram / gradient
# kernels): whole GPU suite, smoke, every bench configuration on one GPU, precision study incl. n=16384, ncu.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -rs -p no:cacheprovider > gpurun_out/r2c_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2c_pytest_gpu.txt; tail -4 gpurun_out/r2c_pytest_gpu.txt
cp gpurun_out/variance_error.json gpurun_out/r2c_variance_error.json 2>/dev/null
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2c_smoke.txt 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/r2c_smoke.txt
python bench.py > gpurun_out/r2c_bench_headline.json 2> gpurun_out/r2c_bench_headline.err; echo "headline exit $?"
python bench.py --engine i8 --fit 0 --cpu-sample 0 > gpurun_out/r2c_bench_headline_i8.json 2> gpurun_out/r2c_bench_headline_i8.err; echo "headline i8 exit $?"
python bench.py --config cfg2 > gpurun_out/r2c_bench_cfg2.json 2> gpurun_out/r2c_bench_cfg2.err; echo "cfg2 exit $?"
python bench.py --config cfg4 --steps 3 > gpurun_out/r2c_bench_cfg4.json 2> gpurun_out/r2c_bench_cfg4.err; echo "cfg4 exit $?"
python bench.py --config cfg5 --steps 1 --warmup 1 --parity-n 2048 --cpu-sample 1024 > gpurun_out/r2c_bench_cfg5.json 2> gpurun_out/r2c_bench_cfg5.err; echo "cfg5 exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2c_bench_reference.json 2> gpurun_out/r2c_bench_reference.err; echo "ref exit $?"
python tools/predict_precision.py --cases headline,cfg2,illcond,n8200,cfg5 > gpurun_out/r2c_predict_precision.md 2>&1; echo "precision exit $?"
for R in 64 8; do python tools/fit_only.py $R 4 > gpurun_out/r2c_fit_only_R$R.log 2>&1; tail -2 gpurun_out/r2c_fit_only_R$R.log; done
B="python bench.py --steps 2 --warmup 1 --m 32768 --fit 0 --cpu-sample 0 --parity-n 0"
$B > gpurun_out/r2c_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2c_launches.csv $B > gpurun_out/r2c_ncu0.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:colsumsq_i8 -s 20 -c 1 -o gpurun_out/r2c_variance_full $B > gpurun_out/r2c_ncu1.log 2>&1
python tools/fit_only.py 8 2 > gpurun_out/r2c_fit_only.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'gram_train_kernel|lml_grad_kernel' -c 2 \
    -o gpurun_out/r2c_gram_grad python tools/fit_only.py 8 2 > gpurun_out/r2c_ncu2.log 2>&1
echo done
