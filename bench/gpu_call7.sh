#!/bin/bash
mkdir -p gpurun_out
timeout 120 python bench/tc_probe.py > gpurun_out/tc_probe.log 2>&1
cat gpurun_out/tc_probe.log | cut -c1-400
timeout 900 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 300 --warmup 30 > gpurun_out/bench_tf32.log 2>&1
for f in pytest_gpu bench_tf32; do echo "== $f"; tail -n 4 gpurun_out/$f.log | cut -c1-300; done
