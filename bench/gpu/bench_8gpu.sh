#!/bin/bash
# minimal 8-GPU check of the default bench after kernel changes: own arm at 8 and 4 GPUs
mkdir -p gpurun_out
for n in 8 4; do
CUDA_VISIBLE_DEVICES=$(seq -s, 0 $((n-1))) timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py --gpus $n --steps 2000 --warmup 200 > gpurun_out/bench_own_n$n.log 2>&1
echo "n=$n rc=$?"
python - <<PY
import json
for l in open("gpurun_out/bench_own_n$n.log"):
    if l.startswith("{"):
        d=json.loads(l); print("N=$n", round(d["value"]), round(d["ms_per_step"]*1e3,1), "us/step; exposed", d["exposed_comm_us_per_step"], "e2e", round(d["e2e"]["value"]), "div", d["replica_divergence"], d["clocks"]["reasons"])
PY
done
