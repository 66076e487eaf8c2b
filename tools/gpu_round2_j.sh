#!/bin/bash
# round 2, GPU call J: compile-time MMA lists on resident operands (what does the variance kernel's instruction mix
# cost on the tensor pipe?), then the A/B the previous call missed: round-1 per-tile kernels (base library built from
# HEAD) against the persistent kernels on ONE box -- LML+gradient at R = 64 / 8 and the headline step.
set -u
mkdir -p gpurun_out
R=$PWD
python tools/i8_mma_list.py > gpurun_out/r2j_i8_mma_list.md 2> gpurun_out/r2j_i8_mma_list.err; echo "mma_list exit $?"; cat gpurun_out/r2j_i8_mma_list.md; tail -3 gpurun_out/r2j_i8_mma_list.err
for lib in base new; do
  if [ $lib = base ]; then export TB200_LIB=$R/turbo_b200/libturbo_b200_base.so; else unset TB200_LIB; fi
  python tools/fit_only.py 64 4 > gpurun_out/r2j_fit_${lib}_R64.log 2>&1; tail -n 2 gpurun_out/r2j_fit_${lib}_R64.log
  python tools/fit_only.py 8 4 > gpurun_out/r2j_fit_${lib}_R8.log 2>&1; tail -n 2 gpurun_out/r2j_fit_${lib}_R8.log
done
for rep in 1 2; do
for lib in base new; do
  if [ $lib = base ]; then export TB200_LIB=$R/turbo_b200/libturbo_b200_base.so; else unset TB200_LIB; fi
  timeout 400 python bench.py --fit 0 --cpu-sample 0 --parity-n 4096 --steps 3 --warmup 3 > gpurun_out/r2j_headline_${lib}_${rep}.json 2> gpurun_out/r2j_headline_${lib}_${rep}.err
  echo "$lib exit $?"; python - <<PY
import json
try:
    l = json.load(open("gpurun_out/r2j_headline_${lib}_${rep}.json")); r = l["roofline"]
    print("$lib $rep", "value %.4g e2e %.4g ms/step %.2f  var kernel %.4f ms frac %.3f clocks %s parity %s" % (l["value"], l["e2e"]["value"], l["ms_per_step"], r["avg_launch_ms"], r["frac"], l["clocks"]["sm_mhz"], l["parity_check"]["argmax_equal"]))
except Exception as e:
    print("$lib", "no line:", e)
PY
done
done
unset TB200_LIB
echo done
