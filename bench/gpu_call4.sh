#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_tc_gemm_gpu.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest_tc.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_tc.log
tail -n 40 gpurun_out/pytest_tc.log | cut -c1-300
