#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 300 --warmup 30 > gpurun_out/bench_auto.log 2>&1
timeout 300 python bench.py --steps 20 --warmup 5 --no-graph --no-overlap > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 110 -c 33 --csv --log-file gpurun_out/launches_fp32.csv \
    python bench.py --steps 20 --warmup 5 --no-graph --no-overlap > gpurun_out/ncu_launches2.log 2>&1
for f in pytest_gpu bench_auto; do echo "== $f"; tail -n 3 gpurun_out/$f.log | cut -c1-250; done
