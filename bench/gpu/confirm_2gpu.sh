#!/bin/bash
# 2-GPU confirmation: all multi-GPU tests, LeNet + wide bench (own arm), reference arm
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/pytest_multi_n$N.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_multi_n$N.log
tail -n 4 gpurun_out/pytest_multi_n$N.log | cut -c1-300
for cfg in lenet_b256 wide_cifar; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29634 bench.py --gpus $N --steps 1000 --warmup 100 --config $cfg > gpurun_out/bench_${cfg}_n$N.log 2>&1
echo "$cfg rc=$?" >> gpurun_out/bench_${cfg}_n$N.log
python - <<PY
import json
for l in open("gpurun_out/bench_${cfg}_n$N.log"):
    if l.startswith("{"):
        d=json.loads(l); print("$cfg", round(d["value"]), round(d["ms_per_step"]*1e3,1), "us/step; exposed", d["exposed_comm_us_per_step"], "e2e", round(d["e2e"]["value"]), "div", d["replica_divergence"])
PY
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29633 bench.py --impl reference --gpus $N --steps 300 --warmup 30 > gpurun_out/bench_ref_n$N.log 2>&1
grep -o '"value": [0-9.]*' gpurun_out/bench_ref_n$N.log | head -1
