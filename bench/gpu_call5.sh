#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 300 --warmup 30 > gpurun_out/bench_tf32.log 2>&1
timeout 300 python bench.py --steps 300 --warmup 30 --dtype fp32 > gpurun_out/bench_fp32.log 2>&1
timeout 300 python bench.py --steps 20 --warmup 5 --no-graph --no-overlap > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 110 -c 44 --csv --log-file gpurun_out/launches_tf32.csv \
    python bench.py --steps 20 --warmup 5 --no-graph --no-overlap > gpurun_out/ncu_launches.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu_launches.log
for f in pytest_gpu bench_tf32 bench_fp32; do echo "== $f"; tail -n 4 gpurun_out/$f.log | cut -c1-400; done
