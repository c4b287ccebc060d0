#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_tc_conv_gpu.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest_tcconv.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_tcconv.log
grep -v "^E   \|^tests" gpurun_out/pytest_tcconv.log | tail -n 30 | cut -c1-200
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=short -p no:cacheprovider -k "tc_conv_training" > gpurun_out/pytest_tctrain.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_tctrain.log
tail -n 15 gpurun_out/pytest_tctrain.log | cut -c1-250
