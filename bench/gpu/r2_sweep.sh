#!/bin/bash
# Round 2: all-reduce + update sweep with the push protocol's rows (gpurun --gpus N), plus the engine tests that exercise eval graphs.
N=${N:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_engine_gpu.py tests/test_pump_gpu.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -2
SWEEP_MAX_BYTES=${SWEEP_MAX_BYTES:-16777216} timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port 29631 bench/allreduce_sweep.py > gpurun_out/sweep_n$N.log 2>&1
echo "sweep rc=$?"; grep '^{' gpurun_out/sweep_n$N.log | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l)
    print(r['bytes'], ' '.join(f\"{k[:-3]}={r[k]:.1f}\" for k in ('nccl_us','push_us','oneshot_us','twoshot_us','nvls_us') if k in r))"
