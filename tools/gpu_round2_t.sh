#!/bin/bash
# round 2, GPU call T: staged K* CTAs per SM beside the (now integer-epilogue, persistent) variance kernel: 2 x 256-point
# chunks (default) vs 2 / 3 / 4 x 128-point chunks; headline step, each checked by the in-bench parity_check
set -u
mkdir -p gpurun_out
R=$PWD
line() { python - "$1" "$2" <<'PY'
import json, sys
try:
    l = json.load(open(sys.argv[1])); r = l["roofline"]
    print(sys.argv[2], "value %.4g e2e %.4g ms/step %.2f  var kernel %.4f ms clocks %s parity %s" % (l["value"], l["e2e"]["value"], l["ms_per_step"], r["avg_launch_ms"], l["clocks"]["sm_mhz"], (l.get("parity_check") or {}).get("argmax_equal")))
except Exception as e:
    print(sys.argv[2], "no line:", e)
PY
}
for rep in 1 2; do
for lib in new ks16x2 ks16x3 ks16x4; do
  if [ $lib = new ]; then unset TB200_LIB; else export TB200_LIB=$R/turbo_b200/libturbo_b200_$lib.so; fi
  timeout 400 python bench.py --fit 0 --cpu-sample 0 --parity-n 4096 --steps 3 --warmup 3 > gpurun_out/r2t_headline_${lib}_${rep}.json 2> gpurun_out/r2t_headline_${lib}_${rep}.err
  line gpurun_out/r2t_headline_${lib}_${rep}.json "headline $lib $rep"
done
done
unset TB200_LIB
echo done
