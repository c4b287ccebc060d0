#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_tc_win_gpu.py -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest_tcw.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_tcw.log
tail -n 6 gpurun_out/pytest_tcw.log | cut -c1-300
timeout 300 python bench/tcw_probe.py > gpurun_out/tcw_probe.log 2>&1
tail -n 8 gpurun_out/tcw_probe.log | cut -c1-400
