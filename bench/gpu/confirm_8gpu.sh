#!/bin/bash
# 8-GPU confirmation: P2P tests, all-reduce sweep, scaling bench at 8 and 4 (own + reference), strong scaling, wide, torch DDP
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
run() { local n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@"; }
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q --tb=short -p no:cacheprovider -k "8" > gpurun_out/pytest_multi_n$N.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_multi_n$N.log
tail -n 3 gpurun_out/pytest_multi_n$N.log | cut -c1-200
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT run $N bench/allreduce_sweep.py > gpurun_out/sweep_n$N.log 2>&1
echo "sweep rc=$?" >> gpurun_out/sweep_n$N.log
run $N bench.py --gpus $N --steps 2000 --warmup 200 > gpurun_out/bench_own_n$N.log 2>&1
run $N bench.py --impl reference --gpus $N --steps 300 --warmup 30 > gpurun_out/bench_ref_n$N.log 2>&1
run $N bench.py --gpus $N --steps 1000 --warmup 100 --config strong4096 > gpurun_out/bench_strong_n$N.log 2>&1
run $N bench.py --gpus $N --steps 300 --warmup 30 --config wide_cifar > gpurun_out/bench_wide_n$N.log 2>&1
CUDA_VISIBLE_DEVICES=0,1,2,3 run 4 bench.py --gpus 4 --steps 2000 --warmup 200 > gpurun_out/bench_own_n4.log 2>&1
CUDA_VISIBLE_DEVICES=0,1,2,3 run 4 bench.py --impl reference --gpus 4 --steps 300 --warmup 30 > gpurun_out/bench_ref_n4.log 2>&1
CUDA_VISIBLE_DEVICES=0,1,2,3 run 4 bench.py --gpus 4 --steps 300 --warmup 30 --config wide_cifar > gpurun_out/bench_wide_n4.log 2>&1
run $N baseline/torch_ddp_bench.py --config lenet_b256 > gpurun_out/ddp_lenet_n$N.log 2>&1
run $N baseline/torch_ddp_bench.py --config wide_cifar --tf32 > gpurun_out/ddp_wide_n$N.log 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/bench_*_n[48].log")+glob.glob("gpurun_out/ddp_*_n8.log")):
    got=False
    for l in open(f):
        if l.startswith("{"):
            d=json.loads(l); got=True
            print(f.split("/")[-1], round(d.get("value",0)), "img/s", round(d.get("ms_per_step",0)*1e3,1), "us/step exposed", d.get("exposed_comm_us_per_step"), "e2e", round(d.get("e2e",{}).get("value",0)), "div", d.get("replica_divergence"))
    if not got: print(f.split("/")[-1], "NO RESULT:", open(f).read()[-300:].replace("\n"," | "))
PY
grep -ci "nvls" gpurun_out/sweep_n$N.log
grep "^{" gpurun_out/sweep_n$N.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['bytes'], round(d['nccl_us'],1), round(d['oneshot_us'],1) if d.get('oneshot_us') else None, round(d['twoshot_us'],1), round(d['nvls_us'],1) if d.get('nvls_us') else None)"
