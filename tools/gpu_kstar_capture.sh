#!/bin/bash
# one gpurun call: the default bench line, then an ncu --set full capture of the staged K* kernel
mkdir -p gpurun_out
timeout 240 python bench.py > gpurun_out/bench_1gpu_v2.json 2> gpurun_out/bench_1gpu_v2.err; echo "bench rc=$?"
timeout 240 ncu --set full --clock-control none --import-source on -k regex:kstar_mean_i8_staged -s 30 -c 1 -f \
    -o gpurun_out/r1_kstar_staged python bench.py --steps 1 --warmup 1 --m 32768 --restarts 8 --cpu-sample 0 --full-fit 0 \
    > gpurun_out/ncu_kstar.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out
python -c "
import json; d=json.load(open('gpurun_out/bench_1gpu_v2.json')); print(d['value'], d['e2e']['value'], d['fit']['lml_grad_batched_ms_samples'], d['clocks'])"
