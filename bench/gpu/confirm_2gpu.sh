#!/bin/bash
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q --tb=short -p no:cacheprovider -k "pumped" > gpurun_out/pytest_pump_n$N.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_pump_n$N.log
tail -n 12 gpurun_out/pytest_pump_n$N.log | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29634 bench.py --gpus $N --steps 500 --warmup 50 --config wide_cifar > gpurun_out/bench_wide_cifar_n$N.log 2>&1
python - <<PY
import json
for l in open("gpurun_out/bench_wide_cifar_n$N.log"):
    if l.startswith("{"):
        d=json.loads(l); print("wide", round(d["value"]), round(d["ms_per_step"]*1e3,1), "us/step; exposed", d["exposed_comm_us_per_step"], "e2e", round(d["e2e"]["value"]))
PY
