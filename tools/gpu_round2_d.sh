#!/bin/bash
# round 2, GPU call D (2 GPUs): NCCL correctness with a remote winner, the two-devices-in-one-process test, cfg4 and the
# headline sharded over 2 ranks
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_nccl.py tests/test_gpu_scale.py::test_two_devices_in_one_process -m gpu -q -rs -p no:cacheprovider \
    > gpurun_out/r2d_pytest_2gpu.txt 2>&1; echo "pytest exit $?" >> gpurun_out/r2d_pytest_2gpu.txt; tail -5 gpurun_out/r2d_pytest_2gpu.txt
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
$T tools/nccl_argmax_check.py > gpurun_out/r2d_nccl_check.txt 2>&1; echo "nccl check exit $?"; grep NCCL_CHECK gpurun_out/r2d_nccl_check.txt | head -c 600; echo
$T bench.py --gpus 2 --config cfg4 --steps 3 --cpu-sample 0 > gpurun_out/r2d_bench_2gpu_cfg4.json 2> gpurun_out/r2d_bench_2gpu_cfg4.err; echo "cfg4 exit $?"
$T bench.py --gpus 2 --steps 3 > gpurun_out/r2d_bench_2gpu_headline.json 2> gpurun_out/r2d_bench_2gpu_headline.err; echo "headline exit $?"
echo done
