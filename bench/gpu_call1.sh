#!/bin/bash
# first GPU call: tests, smoke, bench (own + reference), launch list
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 300 python bench.py --steps 300 --warmup 30 > gpurun_out/bench_own.log 2>&1
echo "bench rc=$?" >> gpurun_out/bench_own.log
timeout 300 python bench.py --steps 300 --warmup 30 --no-graph > gpurun_out/bench_own_nograph.log 2>&1
timeout 300 python bench.py --impl reference --steps 300 --warmup 30 > gpurun_out/bench_ref.log 2>&1
echo "ref rc=$?" >> gpurun_out/bench_ref.log
tail -5 gpurun_out/pytest_gpu.log gpurun_out/smoke.log gpurun_out/bench_own.log gpurun_out/bench_ref.log
