#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench/tcw_wgrad_probe.py > gpurun_out/tcw_wgrad_probe.log 2>&1
tail -n 10 gpurun_out/tcw_wgrad_probe.log | cut -c1-300
