#!/bin/bash
# 1-GPU: full GPU test tier, both bench arms, wide config, then launch lists and ncu --set full captures
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python bench.py --impl reference --steps 300 --warmup 30 > gpurun_out/bench_ref_n1.log 2>&1
timeout 300 python bench.py --steps 2000 --warmup 200 > gpurun_out/bench_own_n1.log 2>&1
echo "own rc=$?" >> gpurun_out/bench_own_n1.log
timeout 300 python bench.py --steps 300 --warmup 30 --dtype tf32 > gpurun_out/bench_own_tf32_n1.log 2>&1
timeout 300 python bench.py --steps 300 --warmup 30 --config wide_cifar > gpurun_out/bench_wide_n1.log 2>&1
echo "wide rc=$?" >> gpurun_out/bench_wide_n1.log
for f in pytest_gpu bench_ref_n1 bench_own_n1 bench_own_tf32_n1 bench_wide_n1; do echo "== $f"; tail -n 4 gpurun_out/$f.log | cut -c1-600; done
# launch lists (only after the un-profiled runs above exited)
KF='regex:conv|gemm|head|reduce_update'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KF" -c 200 --csv --log-file gpurun_out/launches_lenet.csv \
    python bench.py --steps 3 --warmup 3 --no-graph > gpurun_out/ncu_launch_lenet.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KF" -c 200 --csv --log-file gpurun_out/launches_wide.csv \
    python bench.py --steps 3 --warmup 3 --no-graph --config wide_cifar > gpurun_out/ncu_launch_wide.log 2>&1
# full captures: skip the warm-up steps' launches, capture one step's worth of every kernel
timeout 600 ncu --set full --clock-control none --import-source on -k "$KF" --launch-skip 33 -c 11 -f -o gpurun_out/lenet_full \
    python bench.py --steps 3 --warmup 3 --no-graph > gpurun_out/ncu_full_lenet.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k "$KF" --launch-skip 39 -c 13 -f -o gpurun_out/wide_full \
    python bench.py --steps 3 --warmup 3 --no-graph --config wide_cifar > gpurun_out/ncu_full_wide.log 2>&1
ls -la gpurun_out | cut -c1-120
tail -n 3 gpurun_out/ncu_full_lenet.log gpurun_out/ncu_full_wide.log | cut -c1-300
