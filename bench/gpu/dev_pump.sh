#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pump_gpu.py tests/test_engine_gpu.py -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest_p.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_p.log
tail -n 3 gpurun_out/pytest_p.log | cut -c1-300
timeout 300 python bench.py --steps 500 --warmup 50 --config wide_cifar > gpurun_out/bench_wide_n1.log 2>&1
echo "rc=$?"; grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_wide_n1.log | head -2
timeout 300 python bench.py --steps 2000 --warmup 200 > gpurun_out/bench_own_n1.log 2>&1
echo "rc=$?"; grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_own_n1.log | head -2
