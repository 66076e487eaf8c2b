#!/bin/bash
# round 2, GPU call Y (2 GPUs): spread of the fit samples at 8 restarts per GPU (16 restarts over 2 ranks)
set -u
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521"
for rep in 1 2 3; do
  $T bench.py --gpus 2 --restarts 16 --steps 2 --warmup 3 --cpu-sample 0 --parity-n 0 > gpurun_out/r2y_bench_2gpu_R16_$rep.json 2> gpurun_out/r2y_bench_2gpu_R16_$rep.err; echo "exit $?"
  python - <<PY
import json
l = json.load(open("gpurun_out/r2y_bench_2gpu_R16_$rep.json")); f = l["fit"]
print("rep $rep samples", f["lml_grad_batched_ms_samples"], "full fit", f.get("full_fit_ms"), f.get("full_fit_lml_evaluations"))
PY
done
echo done
