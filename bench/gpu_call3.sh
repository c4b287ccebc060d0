#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/pytest_noalloc_alone.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_noalloc_alone.log
N=$(nvidia-smi -L | wc -l)
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 300 --warmup 30 > gpurun_out/bench_own_n$N.log 2>&1
echo "rc=$?" >> gpurun_out/bench_own_n$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 bench.py --impl reference --gpus $N --steps 300 --warmup 30 > gpurun_out/bench_ref_n$N.log 2>&1
echo "rc=$?" >> gpurun_out/bench_ref_n$N.log
for f in pytest_gpu pytest_noalloc_alone bench_own_n$N bench_ref_n$N; do echo "== $f"; tail -n 5 gpurun_out/$f.log | cut -c1-600; done
