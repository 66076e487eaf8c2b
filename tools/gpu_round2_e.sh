#!/bin/bash
# round 2, GPU call E (8 GPUs): BASELINE configs 4 and 5 at their named GPU count, the headline, the NCCL check with a
# remote winner, the NCCL / two-device tests
set -u
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518"
$T bench.py --gpus 8 --config cfg4 --steps 3 --warmup 3 --cpu-sample 0 > gpurun_out/r2e_bench_8gpu_cfg4.json 2> gpurun_out/r2e_bench_8gpu_cfg4.err; echo "cfg4 exit $?"
$T bench.py --gpus 8 --steps 3 --warmup 3 --parity-n 4096 > gpurun_out/r2e_bench_8gpu_headline.json 2> gpurun_out/r2e_bench_8gpu_headline.err; echo "headline exit $?"
$T tools/nccl_argmax_check.py > gpurun_out/r2e_nccl_check.txt 2>&1; echo "nccl check exit $?"; grep NCCL_CHECK gpurun_out/r2e_nccl_check.txt | head -c 400; echo
timeout 600 python -m pytest tests/test_gpu_nccl.py tests/test_gpu_scale.py::test_two_devices_in_one_process -m gpu -q -rs -p no:cacheprovider \
    > gpurun_out/r2e_pytest_8gpu.txt 2>&1; echo "pytest exit $?" >> gpurun_out/r2e_pytest_8gpu.txt; tail -n 4 gpurun_out/r2e_pytest_8gpu.txt
timeout 900 $T bench.py --gpus 8 --config cfg5 --steps 1 --warmup 3 --parity-n 512 --cpu-sample 0 > gpurun_out/r2e_bench_8gpu_cfg5.json 2> gpurun_out/r2e_bench_8gpu_cfg5.err; echo "cfg5 exit $?"
for f in cfg4 headline cfg5; do head -c 250 gpurun_out/r2e_bench_8gpu_$f.json; echo; done
echo done
