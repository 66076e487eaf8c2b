#!/bin/bash
# round 2, GPU call E (8 GPUs): BASELINE configs 4 and 5 at their named GPU count, the headline, the NCCL check
set -u
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518"
$T bench.py --gpus 8 --config cfg4 --steps 3 --cpu-sample 0 > gpurun_out/r2e_bench_8gpu_cfg4.json 2> gpurun_out/r2e_bench_8gpu_cfg4.err; echo "cfg4 exit $?"
$T bench.py --gpus 8 --config cfg5 --steps 1 --warmup 1 --parity-n 512 --cpu-sample 0 > gpurun_out/r2e_bench_8gpu_cfg5.json 2> gpurun_out/r2e_bench_8gpu_cfg5.err; echo "cfg5 exit $?"
$T bench.py --gpus 8 --steps 3 --parity-n 4096 > gpurun_out/r2e_bench_8gpu_headline.json 2> gpurun_out/r2e_bench_8gpu_headline.err; echo "headline exit $?"
$T tools/nccl_argmax_check.py > gpurun_out/r2e_nccl_check.txt 2>&1; echo "nccl check exit $?"; grep NCCL_CHECK gpurun_out/r2e_nccl_check.txt | head -c 400; echo
echo done
