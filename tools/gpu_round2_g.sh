#!/bin/bash
# round 2, GPU call G: MMA rate as a function of N; A/B of the instruction chunking (balanced = round 1 vs 256 + remainder)
# on the variance kernel (21 and 28 products) and on the batched fit; then the whole GPU suite on the new default.
set -u
mkdir -p gpurun_out
R=$PWD
python tools/i8_mma_rate.py > gpurun_out/r2g_i8_mma_rate.md 2>&1; cat gpurun_out/r2g_i8_mma_rate.md
for lib in balanced default; do
  if [ $lib = balanced ]; then export TB200_LIB=$R/turbo_b200/libturbo_b200_balanced.so; else unset TB200_LIB; fi
  for eng in i8a i8; do
    timeout 400 python bench.py --engine $eng --fit 0 --cpu-sample 0 --parity-n 4096 --steps 3 --warmup 3 > gpurun_out/r2g_${lib}_${eng}.json 2> gpurun_out/r2g_${lib}_${eng}.err
    echo "$lib $eng exit $?"; python - <<PY
import json
try:
    l = json.load(open("gpurun_out/r2g_${lib}_${eng}.json")); r = l["roofline"]
    print("$lib $eng", "value %.4g ms/step %.2f  var kernel %.4f ms frac %.3f clocks %s parity %s" % (l["value"], l["ms_per_step"], r["avg_launch_ms"], r["frac"], l["clocks"]["sm_mhz"], l["parity_check"]["argmax_equal"]))
except Exception as e:
    print("$lib $eng", "no line:", e)
PY
  done
  python tools/fit_only.py 64 4 > gpurun_out/r2g_fit_${lib}_R64.log 2>&1; tail -2 gpurun_out/r2g_fit_${lib}_R64.log
  python tools/fit_only.py 8 4 > gpurun_out/r2g_fit_${lib}_R8.log 2>&1; tail -2 gpurun_out/r2g_fit_${lib}_R8.log
done
unset TB200_LIB
python -m pytest tests -m gpu -q -rs -p no:cacheprovider > gpurun_out/r2g_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2g_pytest_gpu.txt; tail -12 gpurun_out/r2g_pytest_gpu.txt
cp gpurun_out/variance_error.json gpurun_out/r2g_variance_error.json 2>/dev/null
echo done
