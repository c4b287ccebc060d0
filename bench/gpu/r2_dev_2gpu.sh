#!/bin/bash
# Round 2 development loop on N B200s (gpurun --gpus N): multi-GPU test file + parity tier, bench at N (default and the
# per-bucket handshake protocol for comparison).
N=${N:-2}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_multi_gpu.py tests/test_parity_gpu.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest_multi_n$N.log 2>&1
echo "pytest rc=$?"; tail -n 40 gpurun_out/pytest_multi_n$N.log | cut -c1-600
for algo in auto oneshot; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
      bench.py --gpus $N --steps 1000 --warmup 100 --algo $algo > gpurun_out/bench_n${N}_$algo.log 2>&1
  echo "bench N=$N algo=$algo rc=$?"; grep -o '"ms_per_step": [0-9.]*\|"exposed_comm_us_per_step": [0-9.-]*\|"replica_divergence": [0-9.e-]*' gpurun_out/bench_n${N}_$algo.log
done
timeout 300 python bench.py --steps 1000 --warmup 100 > gpurun_out/bench_n1.log 2>&1
echo "bench N=1 rc=$?"; grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_n1.log
