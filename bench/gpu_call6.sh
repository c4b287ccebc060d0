#!/bin/bash
mkdir -p gpurun_out
timeout 120 python bench/tc_probe.py > gpurun_out/tc_probe.log 2>&1
cat gpurun_out/tc_probe.log | cut -c1-700
