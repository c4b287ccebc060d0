#!/bin/bash
N=2
for cfg in lenet_b256 wide_cifar; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29634 bench.py --gpus $N --steps 1000 --warmup 100 --config $cfg > gpurun_out/bench_${cfg}_n$N.log 2>&1
python - <<PY
import json
for l in open("gpurun_out/bench_${cfg}_n$N.log"):
    if l.startswith("{"):
        d=json.loads(l); print("$cfg", round(d["value"]), round(d["ms_per_step"]*1e3,1), "us/step; exposed", d["exposed_comm_us_per_step"], "e2e", round(d["e2e"]["value"]), "div", d["replica_divergence"])
PY
done
