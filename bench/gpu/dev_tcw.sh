#!/bin/bash
# development loop for the tensor-core convolution kernels: tests, probe (error + phase stamps), wide bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_tc_win_gpu.py tests/test_kernels_gpu.py tests/test_engine_gpu.py -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest_tcw.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_tcw.log
tail -n 6 gpurun_out/pytest_tcw.log | cut -c1-300
timeout 300 python bench/tcw_probe.py 2>&1 | tail -4 | cut -c1-520
timeout 300 python bench.py --steps 300 --warmup 30 --config wide_cifar > gpurun_out/bench_wide_n1.log 2>&1
echo "rc=$?"; grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_wide_n1.log | head -1
