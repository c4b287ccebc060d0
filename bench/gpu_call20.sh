#!/bin/bash
mkdir -p gpurun_out
TCW_ONLY=wide timeout 300 python bench/tcw_probe.py > gpurun_out/tcw_probe_wide.log 2>&1 && \
TCW_ONLY=wide timeout 600 ncu --set full --clock-control none --import-source on -k regex:win_conv -c 6 -f -o gpurun_out/tcw_wide python bench/tcw_probe.py > gpurun_out/ncu_tcw_wide.log 2>&1
tail -n 3 gpurun_out/tcw_probe_wide.log | cut -c1-400
tail -n 3 gpurun_out/ncu_tcw_wide.log | cut -c1-300
