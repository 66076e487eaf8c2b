#!/bin/bash
# round 2, GPU call K: MMA lists on resident operands AND inside the two-stage pipeline with 0..7 digit planes refilled
# per k-block: does the fill traffic into shared memory slow the tensor pipe's operand reads?
set -u
mkdir -p gpurun_out
timeout 600 python tools/i8_mma_list.py > gpurun_out/r2k_i8_mma_list.md 2> gpurun_out/r2k_i8_mma_list.err; echo "mma_list exit $?"; cat gpurun_out/r2k_i8_mma_list.md; tail -3 gpurun_out/r2k_i8_mma_list.err
echo done
