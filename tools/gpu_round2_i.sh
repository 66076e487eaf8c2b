#!/bin/bash
# round 2, GPU call I: (1) what the variance kernel's MMA list costs on resident operands (tb200_i8_mma_list);
# (2) persistent INT8 fit GEMM: fit tests, LML+gradient at R = 64 / 8 against the round-1 per-tile kernel (base library
# built from HEAD); (3) A/B of the headline step, persistent vs per-tile variance kernel, on ONE box.
set -u
mkdir -p gpurun_out
R=$PWD
python tools/i8_mma_list.py > gpurun_out/r2i_i8_mma_list.md 2> gpurun_out/r2i_i8_mma_list.err; echo "mma_list exit $?"; cat gpurun_out/r2i_i8_mma_list.md; tail -3 gpurun_out/r2i_i8_mma_list.err
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_scale.py tests/test_gpu_i8.py -q -x -p no:cacheprovider > gpurun_out/r2i_pytest_fit.txt 2>&1
echo "pytest fit exit $?"; tail -5 gpurun_out/r2i_pytest_fit.txt
for lib in base new; do
  if [ $lib = base ]; then export TB200_LIB=$R/turbo_b200/libturbo_b200_base.so; else unset TB200_LIB; fi
  python tools/fit_only.py 64 4 > gpurun_out/r2i_fit_${lib}_R64.log 2>&1; tail -2 gpurun_out/r2i_fit_${lib}_R64.log
  python tools/fit_only.py 8 4 > gpurun_out/r2i_fit_${lib}_R8.log 2>&1; tail -2 gpurun_out/r2i_fit_${lib}_R8.log
done
for rep in 1 2; do
for lib in base new; do
  if [ $lib = base ]; then export TB200_LIB=$R/turbo_b200/libturbo_b200_base.so; else unset TB200_LIB; fi
  timeout 400 python bench.py --fit 0 --cpu-sample 0 --parity-n 4096 --steps 3 --warmup 3 > gpurun_out/r2i_headline_${lib}_${rep}.json 2> gpurun_out/r2i_headline_${lib}_${rep}.err
  echo "$lib exit $?"; python - <<PY
import json
try:
    l = json.load(open("gpurun_out/r2i_headline_${lib}_${rep}.json")); r = l["roofline"]
    print("$lib $rep", "value %.4g e2e %.4g ms/step %.2f  var kernel %.4f ms frac %.3f clocks %s parity %s" % (l["value"], l["e2e"]["value"], l["ms_per_step"], r["avg_launch_ms"], r["frac"], l["clocks"]["sm_mhz"], l["parity_check"]["argmax_equal"]))
except Exception as e:
    print("$lib", "no line:", e)
PY
done
done
unset TB200_LIB
echo done
