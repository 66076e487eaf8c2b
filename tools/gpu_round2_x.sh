#!/bin/bash
# round 2, GPU call X (2 GPUs): new streaming test, NCCL tests, and the spread of the fit samples at 8 restarts per GPU
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_path.py tests/test_gpu_nccl.py tests/test_gpu_scale.py::test_two_devices_in_one_process -m gpu -x -q -rs -p no:cacheprovider > gpurun_out/r2x_pytest.txt 2>&1
echo "pytest exit $?" >> gpurun_out/r2x_pytest.txt; tail -n 4 gpurun_out/r2x_pytest.txt
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29520"
for rep in 1 2; do
  $T bench.py --gpus 2 --restarts 16 --steps 2 --warmup 3 --m 262144 --cpu-sample 0 --parity-n 0 --full-fit 1 > gpurun_out/r2x_bench_2gpu_R16_$rep.json 2> gpurun_out/r2x_bench_2gpu_R16_$rep.err; echo "exit $?"
  python - <<PY
import json
l = json.load(open("gpurun_out/r2x_bench_2gpu_R16_$rep.json")); f = l["fit"]
print("rep $rep samples", f["lml_grad_batched_ms_samples"], "full fit", f.get("full_fit_ms"), f.get("full_fit_lml_evaluations"))
PY
done
$T bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2x_bench_2gpu_headline.json 2> gpurun_out/r2x_bench_2gpu_headline.err; echo "headline exit $?"; head -c 200 gpurun_out/r2x_bench_2gpu_headline.json; echo
$T bench.py --gpus 2 --config cfg4 --steps 2 --warmup 3 --cpu-sample 0 > gpurun_out/r2x_bench_2gpu_cfg4.json 2> gpurun_out/r2x_bench_2gpu_cfg4.err; echo "cfg4 exit $?"; head -c 200 gpurun_out/r2x_bench_2gpu_cfg4.json; echo
echo done
