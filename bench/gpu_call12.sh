#!/bin/bash
# 8-GPU confirmation: P2P tests at world 8, all-reduce sweep, scaling bench (own + reference), strong scaling
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q --tb=short -p no:cacheprovider -k "8" > gpurun_out/pytest_multi_n$N.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_multi_n$N.log
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29631 bench/allreduce_sweep.py > gpurun_out/sweep_n$N.log 2>&1
echo "sweep rc=$?" >> gpurun_out/sweep_n$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29632 bench.py --gpus $N --steps 300 --warmup 30 > gpurun_out/bench_own_n$N.log 2>&1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29633 bench.py --impl reference --gpus $N --steps 300 --warmup 30 > gpurun_out/bench_ref_n$N.log 2>&1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29634 bench.py --gpus $N --steps 300 --warmup 30 --config strong4096 > gpurun_out/bench_strong_n$N.log 2>&1
for f in pytest_multi_n$N bench_own_n$N bench_ref_n$N bench_strong_n$N; do echo "== $f"; tail -n 3 gpurun_out/$f.log | cut -c1-420; done
grep -i "nvls" gpurun_out/sweep_n$N.log | head -5
grep "^{" gpurun_out/sweep_n$N.log | cut -c1-400
