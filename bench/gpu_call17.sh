#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench/tcw_probe.py > gpurun_out/tcw_probe.log 2>&1
echo "probe rc=$?" >> gpurun_out/tcw_probe.log
tail -n 30 gpurun_out/tcw_probe.log | cut -c1-400
