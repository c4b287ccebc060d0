#!/bin/bash
# Round 2 evidence pass on one B200: GPU test tier, both bench arms, warm-cache breakdown, wide model at batch
# 256 / 1024 / 4096, ncu launch list + `--set full` capture of one step of the default LeNet path, smoke().
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 300 python bench.py --impl reference --steps 300 --warmup 30 > gpurun_out/bench_ref_n1.log 2>&1; echo "ref rc=$?"
timeout 300 python bench.py --steps 2000 --warmup 200 > gpurun_out/bench_own_n1.log 2>&1; echo "own rc=$?"
timeout 300 python bench.py --steps 2000 --warmup 200 --dtype bf16 > gpurun_out/bench_own_bf16_n1.log 2>&1; echo "bf16 rc=$?"
timeout 300 python bench.py --steps 2000 --warmup 200 --dtype fp32 --conv simt > gpurun_out/bench_own_fp32_n1.log 2>&1; echo "fp32 rc=$?"
for f in bench_ref_n1 bench_own_n1 bench_own_bf16_n1 bench_own_fp32_n1; do echo "== $f"; grep -o '"ms_per_step": [0-9.]*' gpurun_out/$f.log | tr '\n' ' '; echo; done
timeout 300 python bench/step_breakdown.py --config lenet_b256 > gpurun_out/breakdown_lenet.log 2>&1
grep -v Warning gpurun_out/breakdown_lenet.log | tail -12
for b in 256 1024 4096; do
  timeout 300 python bench/step_breakdown.py --config wide_cifar --batch $b > gpurun_out/breakdown_wide_b$b.log 2>&1
  echo "wide batch $b:"; grep -v Warning gpurun_out/breakdown_wide_b$b.log | tail -16
done
KF='regex:conv|gemm|head|reduce_update|pack|push'
timeout 300 python bench.py --steps 3 --warmup 3 --no-graph > gpurun_out/plain_nograph.log 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KF" -c 80 --csv --log-file gpurun_out/launches_lenet_r2.csv \
    python bench.py --steps 3 --warmup 3 --no-graph > gpurun_out/ncu_launch_lenet.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k "$KF" --launch-skip 27 -c 9 -f -o gpurun_out/lenet_full_r2 \
    python bench.py --steps 3 --warmup 3 --no-graph > gpurun_out/ncu_full_lenet.log 2>&1
tail -n 2 gpurun_out/ncu_full_lenet.log | cut -c1-200
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -n 1 gpurun_out/smoke.log | cut -c1-200
