#!/bin/bash
# round 2, GPU call V: end-to-end path with two reusable device buffers (Acquirer.run_host): GPU suite, then the
# headline step four times in fresh processes -- is the e2e figure stable now?
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -rs -p no:cacheprovider > gpurun_out/r2v_pytest_gpu.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2v_pytest_gpu.txt; tail -n 4 gpurun_out/r2v_pytest_gpu.txt
for rep in 1 2 3 4; do
  python bench.py --fit 0 --cpu-sample 0 --parity-n 4096 > gpurun_out/r2v_headline_$rep.json 2> gpurun_out/r2v_headline_$rep.err
  python - <<PY
import json
l = json.load(open("gpurun_out/r2v_headline_$rep.json"))
print("rep $rep value %.4g e2e %.4g (%.1f ms/step) resident %.1f ms/step clocks %s" % (l["value"], l["e2e"]["value"], l["e2e"]["ms_per_step"], l["ms_per_step"], l["clocks"]["sm_mhz"]))
PY
done
echo done
