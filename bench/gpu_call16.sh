#!/bin/bash
# step pump v2 (prefetch branch + mapped loss): tests, e2e bench; regression on the kernels that changed signature
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pump_gpu.py tests/test_kernels_gpu.py -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest_pump2.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_pump2.log
tail -n 25 gpurun_out/pytest_pump2.log | cut -c1-300
timeout 300 python bench.py --steps 2000 --warmup 200 > gpurun_out/bench_pump2_n1.log 2>&1
echo "rc=$?" >> gpurun_out/bench_pump2_n1.log
tail -n 3 gpurun_out/bench_pump2_n1.log | cut -c1-1800
