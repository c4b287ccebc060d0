#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --steps 300 --warmup 30 --dtype tf32 --conv tc > gpurun_out/bench_lenet_tc.log 2>&1
tail -n 1 gpurun_out/bench_lenet_tc.log | cut -c1-300
KF='regex:conv|gemm|head|reduce_update|pack'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KF" -c 200 --csv --log-file gpurun_out/launches_lenet_tc.csv \
    python bench.py --steps 3 --warmup 3 --no-graph --dtype tf32 --conv tc > gpurun_out/ncu_launch_lenet_tc.log 2>&1
tail -n 2 gpurun_out/ncu_launch_lenet_tc.log | cut -c1-200
