#!/bin/bash
# round 2, GPU call L: integer-pipe recombination in the epilogues + A-collector hints + K* writes only the planes read.
# Correctness first; then on ONE box the headline step with (base) the round-1 per-tile kernels, (pertile) the new code
# one CTA per tile, (new) persistent; cfg2 and the fit for base / new.
set -u
mkdir -p gpurun_out
R=$PWD
timeout 900 python -m pytest tests/test_gpu_i8.py tests/test_gpu_path.py tests/test_gpu_parity.py tests/test_gpu_scale.py -q -x -p no:cacheprovider > gpurun_out/r2l_pytest.txt 2>&1
echo "pytest exit $?"; tail -n 4 gpurun_out/r2l_pytest.txt
line() { python - "$1" "$2" <<'PY'
import json, sys
try:
    l = json.load(open(sys.argv[1])); r = l["roofline"]
    print(sys.argv[2], "value %.4g e2e %.4g ms/step %.2f  var kernel %.4f ms frac %.3f clocks %s parity %s" % (l["value"], l["e2e"]["value"], l["ms_per_step"], r["avg_launch_ms"], r["frac"], l["clocks"]["sm_mhz"], (l.get("parity_check") or {}).get("argmax_equal")))
except Exception as e:
    print(sys.argv[2], "no line:", e)
PY
}
for rep in 1 2; do
for lib in base pertile new; do
  if [ $lib = new ]; then unset TB200_LIB; else export TB200_LIB=$R/turbo_b200/libturbo_b200_$lib.so; fi
  timeout 400 python bench.py --fit 0 --cpu-sample 0 --parity-n 4096 --steps 3 --warmup 3 > gpurun_out/r2l_headline_${lib}_${rep}.json 2> gpurun_out/r2l_headline_${lib}_${rep}.err
  line gpurun_out/r2l_headline_${lib}_${rep}.json "headline $lib $rep"
done
done
for lib in base pertile new; do
  if [ $lib = new ]; then unset TB200_LIB; else export TB200_LIB=$R/turbo_b200/libturbo_b200_$lib.so; fi
  timeout 400 python bench.py --config cfg2 --cpu-sample 0 --steps 3 --warmup 3 > gpurun_out/r2l_cfg2_${lib}.json 2> gpurun_out/r2l_cfg2_${lib}.err
  line gpurun_out/r2l_cfg2_${lib}.json "cfg2 $lib"
  timeout 400 python bench.py --engine i8 --fit 0 --cpu-sample 0 --parity-n 4096 --steps 3 --warmup 3 > gpurun_out/r2l_headline28_${lib}.json 2> gpurun_out/r2l_headline28_${lib}.err
  line gpurun_out/r2l_headline28_${lib}.json "headline 28 products $lib"
done
for lib in base new; do
  if [ $lib = new ]; then unset TB200_LIB; else export TB200_LIB=$R/turbo_b200/libturbo_b200_$lib.so; fi
  python tools/fit_only.py 64 4 > gpurun_out/r2l_fit_${lib}_R64.log 2>&1; tail -n 2 gpurun_out/r2l_fit_${lib}_R64.log
  python tools/fit_only.py 8 4 > gpurun_out/r2l_fit_${lib}_R8.log 2>&1; tail -n 2 gpurun_out/r2l_fit_${lib}_R8.log
done
unset TB200_LIB
echo done
