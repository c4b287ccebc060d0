#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_tc_win_gpu.py tests/test_tc_gemm_gpu.py tests/test_kernels_gpu.py -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest_k.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_k.log
tail -n 3 gpurun_out/pytest_k.log | cut -c1-300
timeout 300 python bench.py --steps 300 --warmup 30 --config wide_cifar > gpurun_out/bench_wide_n1.log 2>&1
echo "rc=$?"; tail -n 1 gpurun_out/bench_wide_n1.log | cut -c1-330
