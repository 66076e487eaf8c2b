#!/bin/bash
# round 2, GPU call R: final state (recursive inverse at 21 products): GPU suite, smoke, default bench line, reference
# arm, launch lists (acquire step; fit at 64 restarts) and ncu --set full of the variance kernel and the fit GEMM
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs -p no:cacheprovider > gpurun_out/r2r_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2r_pytest_gpu.txt; tail -n 8 gpurun_out/r2r_pytest_gpu.txt
cp gpurun_out/variance_error.json gpurun_out/r2r_variance_error.json 2>/dev/null
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2r_smoke.txt 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/r2r_smoke.txt
python bench.py > gpurun_out/r2r_bench_headline.json 2> gpurun_out/r2r_bench_headline.err; echo "headline exit $?"; head -c 300 gpurun_out/r2r_bench_headline.json; echo
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2r_bench_reference.json 2> gpurun_out/r2r_bench_reference.err; echo "ref exit $?"; head -c 300 gpurun_out/r2r_bench_reference.json; echo
for R in 64 8; do python tools/fit_only.py $R 4 > gpurun_out/r2r_fit_only_R$R.log 2>&1; tail -n 2 gpurun_out/r2r_fit_only_R$R.log; done
B="python bench.py --steps 2 --warmup 1 --m 32768 --fit 0 --cpu-sample 0 --parity-n 0"
$B > gpurun_out/r2r_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2r_launches.csv $B > gpurun_out/r2r_ncu0.log 2>&1
echo "ncu list exit $?"
ncu --set full --clock-control none --import-source on -k regex:colsumsq_i8 -s 20 -c 1 -o gpurun_out/r2r_variance_full $B > gpurun_out/r2r_ncu1.log 2>&1
echo "ncu full exit $?"
python tools/fit_only.py 64 2 > gpurun_out/r2r_fit_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2r_launches_fit_R64.csv python tools/fit_only.py 64 2 > gpurun_out/r2r_ncu2.log 2>&1
echo "ncu fit list exit $?"
ncu --set full --clock-control none --import-source on -k regex:gemm_i8_kernel -s 60 -c 3 -o gpurun_out/r2r_gemm_i8_full python tools/fit_only.py 8 2 > gpurun_out/r2r_ncu3.log 2>&1
echo "ncu gemm full exit $?"
echo done
