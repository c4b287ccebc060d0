#!/bin/bash
mkdir -p gpurun_out
timeout 200 python bench/tc_probe.py > gpurun_out/tc_probe.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_gemm -c 2 -f -o gpurun_out/tc_gemm_full python bench/tc_probe.py > gpurun_out/ncu_tc_gemm.log 2>&1
tail -n 3 gpurun_out/ncu_tc_gemm.log | cut -c1-200
