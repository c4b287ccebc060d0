#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_tc_gemm_gpu.py tests/test_kernels_gpu.py -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest_gemm.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gemm.log
tail -n 8 gpurun_out/pytest_gemm.log | cut -c1-300
timeout 200 python bench/tc_probe.py 2>&1 | grep "split/retiled\|^M" | cut -c1-200
timeout 300 python bench.py --steps 300 --warmup 30 --config wide_cifar > gpurun_out/bench_wide_n1.log 2>&1
echo "rc=$?"; grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_wide_n1.log | head -1
