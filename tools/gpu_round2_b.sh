#!/bin/bash
# round 2, GPU call B: A/B of the variance kernel's pipeline granularity (64- vs 32-byte K tiles: 2 vs 4 stages) and of
# the paired-cluster variant (multicast candidate operand), each checked against the oracle by bench.py's parity_check.
set -u
mkdir -p gpurun_out
R=$PWD
for kb in 64 32; do
  for pair in 0 1; do
    tag=kb${kb}_pair${pair}
    export TB200_LIB=$R/turbo_b200/libturbo_b200_kb${kb}.so TB200_DEBUG=acq_pair=${pair}
    timeout 600 python bench.py --fit 0 --cpu-sample 0 --parity-n 4096 --steps 3 --warmup 3 \
        > gpurun_out/r2b_bench_${tag}.json 2> gpurun_out/r2b_bench_${tag}.err
    echo "$tag bench exit $?"; python - <<PY
import json
try:
    l = json.load(open("gpurun_out/r2b_bench_${tag}.json"))
    print("${tag}", "value %.4g e2e %.4g ms/step %.2f  var kernel %.4f ms frac %.3f  parity %s" % (l["value"], l["e2e"]["value"], l["ms_per_step"], l["roofline"]["avg_launch_ms"], l["roofline"]["frac"], l["parity_check"]["argmax_equal"]))
except Exception as e:
    print("${tag}", "no line:", e)
PY
    if [ "$pair" = 1 ]; then
      timeout 900 python -m pytest tests/test_gpu_i8.py -q -x -p no:cacheprovider > gpurun_out/r2b_pytest_${tag}.txt 2>&1
      echo "$tag pytest exit $?"; tail -3 gpurun_out/r2b_pytest_${tag}.txt
    fi
  done
done
for g in i8l i8x; do
  export TB200_LIB=$R/turbo_b200/libturbo_b200_kb32.so TB200_DEBUG=acq_pair=1
  timeout 600 python bench.py --fit 0 --cpu-sample 0 --parity-n 4096 --steps 3 --warmup 3 --engine $g \
      > gpurun_out/r2b_bench_kb32_pair1_${g}.json 2> gpurun_out/r2b_bench_kb32_pair1_${g}.err
  echo "$g exit $?"; head -c 400 gpurun_out/r2b_bench_kb32_pair1_${g}.json; echo
done
echo done
