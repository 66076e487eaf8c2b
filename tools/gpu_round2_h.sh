#!/bin/bash
# round 2, GPU call H: persistent variance kernel (one CTA per SM walks the tiles, cross-tile prefetch, conversion-free
# epilogue) -- correctness first (INT8 tests incl. n=16384 multi-pass), then the headline step, cfg2, and a launch list.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_i8.py tests/test_gpu_path.py -q -x -p no:cacheprovider > gpurun_out/r2h_pytest_i8.txt 2>&1
echo "pytest i8 exit $?"; tail -5 gpurun_out/r2h_pytest_i8.txt
for eng in auto i8; do
  timeout 400 python bench.py --engine $eng --fit 0 --cpu-sample 0 --parity-n 4096 --steps 3 --warmup 3 > gpurun_out/r2h_headline_${eng}.json 2> gpurun_out/r2h_headline_${eng}.err
  echo "$eng exit $?"; python - <<PY
import json
try:
    l = json.load(open("gpurun_out/r2h_headline_${eng}.json")); r = l["roofline"]
    print("$eng", "value %.4g e2e %.4g ms/step %.2f  var kernel %.4f ms frac %.3f clocks %s parity %s" % (l["value"], l["e2e"]["value"], l["ms_per_step"], r["avg_launch_ms"], r["frac"], l["clocks"]["sm_mhz"], l["parity_check"]["argmax_equal"]))
except Exception as e:
    print("$eng", "no line:", e)
PY
done
timeout 400 python bench.py --config cfg2 --cpu-sample 0 --steps 3 --warmup 3 > gpurun_out/r2h_cfg2.json 2> gpurun_out/r2h_cfg2.err
echo "cfg2 exit $?"; head -c 300 gpurun_out/r2h_cfg2.json; echo
timeout 600 python -m pytest tests -m gpu -q -rs -x -p no:cacheprovider > gpurun_out/r2h_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2h_pytest_gpu.txt; tail -6 gpurun_out/r2h_pytest_gpu.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:colsumsq_i8 -s 20 -c 1 -o gpurun_out/r2h_variance_full python bench.py --steps 2 --warmup 1 --m 32768 --fit 0 --cpu-sample 0 --parity-n 0 > gpurun_out/r2h_ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/r2h_ncu.log
echo done
