#!/bin/bash
# native step pump: tests + e2e bench with and without it
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_pump_gpu.py -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest_pump.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_pump.log
tail -n 25 gpurun_out/pytest_pump.log | cut -c1-300
timeout 300 python bench.py --steps 2000 --warmup 200 > gpurun_out/bench_pump_n1.log 2>&1
echo "rc=$?" >> gpurun_out/bench_pump_n1.log
timeout 300 python bench.py --steps 2000 --warmup 200 --no-pump > gpurun_out/bench_nopump_n1.log 2>&1
timeout 300 python bench.py --steps 300 --warmup 30 --config wide_cifar > gpurun_out/bench_pump_wide_n1.log 2>&1
for f in bench_pump_n1 bench_nopump_n1 bench_pump_wide_n1; do echo "== $f"; tail -n 3 gpurun_out/$f.log | cut -c1-1800; done
