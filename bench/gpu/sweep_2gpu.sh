#!/bin/bash
# 2-GPU: reduce+update with unrolled chunk loops: tests, sweep, wide bench with the new algorithm choice
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/pytest_multi_n$N.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_multi_n$N.log
tail -n 4 gpurun_out/pytest_multi_n$N.log | cut -c1-300
SWEEP_MAX_BYTES=67108864 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29631 bench/allreduce_sweep.py > gpurun_out/sweep_n$N.log 2>&1
echo "sweep rc=$?" >> gpurun_out/sweep_n$N.log
grep "^{" gpurun_out/sweep_n$N.log | cut -c1-300 | tail -12
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29634 bench.py --gpus $N --steps 300 --warmup 30 --config wide_cifar > gpurun_out/bench_wide_n$N.log 2>&1
echo "wide rc=$?" >> gpurun_out/bench_wide_n$N.log
tail -n 2 gpurun_out/bench_wide_n$N.log | cut -c1-400
