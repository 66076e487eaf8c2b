#!/bin/bash
# round 2, GPU call Z: cfg2 (n = 1024, d = 6: K*-bound) with the staged K* kernel (128-point chunks, 3 / 4 CTAs per SM)
# instead of the direct-load kernel
set -u
mkdir -p gpurun_out
R=$PWD
for rep in 1 2; do
for lib in new st1024x3 st1024x4; do
  if [ $lib = new ]; then unset TB200_LIB; else export TB200_LIB=$R/turbo_b200/libturbo_b200_$lib.so; fi
  python bench.py --config cfg2 --cpu-sample 0 --steps 5 --warmup 3 > gpurun_out/r2z_cfg2_${lib}_$rep.json 2> gpurun_out/r2z_cfg2_${lib}_$rep.err
  python - <<PY
import json
try:
    l = json.load(open("gpurun_out/r2z_cfg2_${lib}_$rep.json"))
    print("$lib $rep value %.4g e2e %.4g ms/step %.2f parity %s" % (l["value"], l["e2e"]["value"], l["ms_per_step"], (l.get("parity_check") or {}).get("argmax_equal")))
except Exception as e:
    print("$lib $rep no line", e)
PY
done
done
echo done
