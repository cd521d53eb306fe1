#!/bin/bash
# ncu --set full of the 128x128 DGEMM kernel at 4096^3 (one launch)
set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dgemm_nn -s 2 -c 1 -o gpurun_out/t_dgemm_4096 -f python tools/gemm_timing.py 4096x4096x4096 > gpurun_out/t_ncu.log 2>&1
tail -5 gpurun_out/t_ncu.log; ls -la gpurun_out/*.ncu-rep
