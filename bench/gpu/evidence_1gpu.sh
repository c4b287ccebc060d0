#!/bin/bash
# final evidence pass (1 GPU): full GPU tier, both bench arms, wide config, launch lists, ncu --set full of one step of each model
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 3 gpurun_out/pytest_gpu.log | cut -c1-200
timeout 300 python bench.py --impl reference --steps 300 --warmup 30 > gpurun_out/bench_ref_n1.log 2>&1
timeout 300 python bench.py --steps 2000 --warmup 200 > gpurun_out/bench_own_n1.log 2>&1
echo "own rc=$?" >> gpurun_out/bench_own_n1.log
timeout 300 python bench.py --steps 500 --warmup 50 --config wide_cifar > gpurun_out/bench_wide_n1.log 2>&1
echo "wide rc=$?" >> gpurun_out/bench_wide_n1.log
timeout 300 python bench/step_breakdown.py --config wide_cifar > gpurun_out/breakdown_wide.log 2>&1
timeout 300 python bench/step_breakdown.py --config lenet_b256 > gpurun_out/breakdown_lenet.log 2>&1
for f in bench_ref_n1 bench_own_n1 bench_wide_n1; do echo "== $f"; tail -n 2 gpurun_out/$f.log | cut -c1-420; done
KF='regex:conv|gemm|head|reduce_update|pack'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KF" -c 100 --csv --log-file gpurun_out/launches_lenet_final.csv \
    python bench.py --steps 3 --warmup 3 --no-graph > gpurun_out/ncu_launch_lenet.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KF" -c 100 --csv --log-file gpurun_out/launches_wide_final.csv \
    python bench.py --steps 3 --warmup 3 --no-graph --config wide_cifar > gpurun_out/ncu_launch_wide.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k "$KF" --launch-skip 33 -c 11 -f -o gpurun_out/lenet_full \
    python bench.py --steps 3 --warmup 3 --no-graph > gpurun_out/ncu_full_lenet.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k "$KF" --launch-skip 42 -c 14 -f -o gpurun_out/wide_full \
    python bench.py --steps 3 --warmup 3 --no-graph --config wide_cifar > gpurun_out/ncu_full_wide.log 2>&1
tail -n 2 gpurun_out/ncu_full_lenet.log gpurun_out/ncu_full_wide.log | cut -c1-200
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -n 1 gpurun_out/smoke.log | cut -c1-200
