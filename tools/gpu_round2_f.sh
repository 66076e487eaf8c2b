#!/bin/bash
# round 2, GPU call F: what bounds the variance kernel?  The same launch with (E1/E2/E4/E5) some digit planes NOT loaded
# (results meaningless, time only: is the kernel bound by the bytes that enter the SM?) and (E3) three pipeline stages.
set -u
mkdir -p gpurun_out
R=$PWD
for tag in E0 E1 E2 E3 E4 E5 E0; do
  export TB200_LIB=$R/turbo_b200/libturbo_b200_$tag.so
  timeout 300 python bench.py --engine i8l --fit 0 --cpu-sample 0 --parity-n 0 --steps 3 --warmup 3 > gpurun_out/r2f_$tag.json 2> gpurun_out/r2f_$tag.err
  echo "$tag exit $?"; python - <<PY
import json
try:
    l = json.load(open("gpurun_out/r2f_$tag.json")); r = l["roofline"]
    print("$tag", "value %.4g ms/step %.2f  var kernel %.4f ms  clocks %s" % (l["value"], l["ms_per_step"], r["avg_launch_ms"], l["clocks"]["sm_mhz"]))
except Exception as e:
    print("$tag", "no line:", e)
PY
done
echo done
