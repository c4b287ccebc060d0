#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_tc_win_gpu.py tests/test_tc_gemm_gpu.py tests/test_kernels_gpu.py -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest_tcw.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_tcw.log
tail -n 6 gpurun_out/pytest_tcw.log | cut -c1-300
timeout 300 python bench/tcw_probe.py 2>&1 | tail -2 | cut -c1-500
timeout 300 python bench.py --steps 300 --warmup 30 --config wide_cifar > gpurun_out/bench_wide_n1.log 2>&1
echo "wide rc=$?" >> gpurun_out/bench_wide_n1.log
tail -n 2 gpurun_out/bench_wide_n1.log | cut -c1-330
KF='regex:conv|gemm|head|reduce_update|pack'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KF" -c 200 --csv --log-file gpurun_out/launches_wide4.csv \
    python bench.py --steps 3 --warmup 3 --no-graph --config wide_cifar > gpurun_out/ncu_launch_wide4.log 2>&1
tail -n 1 gpurun_out/ncu_launch_wide4.log | cut -c1-100
