#!/bin/bash
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_kernels_gpu.py -m gpu -x -q --tb=short -p no:cacheprovider -W default > gpurun_out/pytest_multi.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_multi.log
SWEEP_MAX_BYTES=$((64<<20)) timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 bench/allreduce_sweep.py > gpurun_out/sweep_n$N.log 2>&1
echo "sweep rc=$?" >> gpurun_out/sweep_n$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29622 bench.py --gpus $N --steps 300 --warmup 30 > gpurun_out/bench_own_n$N.log 2>&1
for f in pytest_multi sweep_n$N bench_own_n$N; do echo "== $f"; tail -n 25 gpurun_out/$f.log | cut -c1-330; done
