#!/bin/bash
# round 2, GPU call W (8 GPUs): the headline on the final code (recursive inverse, reusable e2e buffers), both arms
set -u
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519"
$T bench.py --gpus 8 --impl reference --steps 2 --warmup 1 > gpurun_out/r2w_bench_8gpu_reference.json 2> gpurun_out/r2w_bench_8gpu_reference.err; echo "reference exit $?"
$T bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2w_bench_8gpu_headline.json 2> gpurun_out/r2w_bench_8gpu_headline.err; echo "headline exit $?"
head -c 300 gpurun_out/r2w_bench_8gpu_headline.json; echo
echo done
