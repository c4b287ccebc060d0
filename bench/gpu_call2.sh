#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/pytest_noalloc_alone.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_noalloc_alone.log
timeout 300 python bench.py --steps 20 --warmup 5 --no-graph --no-overlap > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 110 -c 44 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 20 --warmup 5 --no-graph --no-overlap > gpurun_out/ncu_launches.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu_launches.log
for f in pytest_gpu pytest_noalloc_alone ncu_launches; do echo "== $f"; tail -n 6 gpurun_out/$f.log; done
