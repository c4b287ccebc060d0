#!/bin/bash
# 2-GPU: multi-GPU tests, bench (own arm with the step pump on the e2e path, reference arm), wide config
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/pytest_multi_n$N.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_multi_n$N.log
tail -n 5 gpurun_out/pytest_multi_n$N.log | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29632 bench.py --gpus $N --steps 2000 --warmup 200 > gpurun_out/bench_own_n$N.log 2>&1
echo "own rc=$?" >> gpurun_out/bench_own_n$N.log
tail -n 3 gpurun_out/bench_own_n$N.log | cut -c1-1800
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29633 bench.py --impl reference --gpus $N --steps 300 --warmup 30 > gpurun_out/bench_ref_n$N.log 2>&1
tail -n 1 gpurun_out/bench_ref_n$N.log | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29634 bench.py --gpus $N --steps 300 --warmup 30 --config wide_cifar > gpurun_out/bench_wide_n$N.log 2>&1
echo "wide rc=$?" >> gpurun_out/bench_wide_n$N.log
tail -n 2 gpurun_out/bench_wide_n$N.log | cut -c1-600
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29635 baseline/torch_ddp_bench.py --config lenet_b256 > gpurun_out/ddp_lenet_n$N.log 2>&1
tail -n 1 gpurun_out/ddp_lenet_n$N.log | cut -c1-400
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29636 baseline/torch_ddp_bench.py --config wide_cifar --tf32 > gpurun_out/ddp_wide_n$N.log 2>&1
tail -n 1 gpurun_out/ddp_wide_n$N.log | cut -c1-400
