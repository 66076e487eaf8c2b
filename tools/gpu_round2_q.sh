#!/bin/bash
# round 2, GPU call Q: slice products per fit stage with the recursive inverse (tools/fit_precision.py): accuracy against
# the FP64 DMMA path and time, n = 4096, R = 64 incl. four hard members
set -u
mkdir -p gpurun_out
TB200_DEBUG=fit_engine=dmma python tools/fit_precision.py run dmma
TB200_DEBUG=fit_gmax=9 python tools/fit_precision.py run g999
python tools/fit_precision.py run g887
TB200_DEBUG=fit_gmax_trtri=7 python tools/fit_precision.py run g877
TB200_DEBUG=fit_gmax=7 python tools/fit_precision.py run g777
TB200_DEBUG=fit_gmax_trtri=9 python tools/fit_precision.py run g897
python tools/fit_precision.py compare dmma g999 g897 g887 g877 g777 | tee gpurun_out/r2q_fit_precision.txt
python tools/fit_precision.py compare g999 g897 g887 g877 g777 | tee -a gpurun_out/r2q_fit_precision.txt
echo done
