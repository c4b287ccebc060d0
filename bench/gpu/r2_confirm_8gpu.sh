#!/bin/bash
# Round 2, 8xB200: the WHOLE multi-GPU test file (world sizes 2, 4 and 8, one spawn each), then the weak-scaling
# bench at 8/4/2/1 GPUs on this one box, the per-bucket handshake protocol at 8 for comparison, strong scaling.
mkdir -p gpurun_out
if [ -z "$SKIP_TESTS" ]; then
timeout 1500 python -m pytest tests/test_multi_gpu.py -m gpu -v --tb=short -p no:cacheprovider > gpurun_out/pytest_multi_n8.log 2>&1
echo "pytest rc=$?"; tail -n 4 gpurun_out/pytest_multi_n8.log | cut -c1-300
fi
run() {  # N tag extra-args
  N=$1; tag=$2; shift 2
  if [ "$N" = 1 ]; then
    timeout 300 python bench.py --gpus 1 --steps 1000 --warmup 100 "$@" > gpurun_out/scale_${tag}.log 2>&1
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 \
        bench.py --gpus $N --steps 1000 --warmup 100 "$@" > gpurun_out/scale_${tag}.log 2>&1
  fi
  echo "bench $tag rc=$?"; grep -o '"ms_per_step": [0-9.]*\|"exposed_comm_us_per_step": [0-9.-]*\|"replica_divergence": [0-9.e-]*' gpurun_out/scale_${tag}.log | tr '\n' ' '; echo
}
run 8 n8; run 4 n4; run 2 n2; run 1 n1
if [ -z "$SHORT" ]; then
run 8 n8_oneshot --algo oneshot
run 8 n8_bf16wire --wire bf16
run 8 n8_strong --config strong4096
fi
if [ -n "$SWEEP" ]; then
SWEEP_MAX_BYTES=4194304 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
    --master-port 29641 bench/allreduce_sweep.py > gpurun_out/sweep_n8.log 2>&1
echo "sweep rc=$?"; grep '^{' gpurun_out/sweep_n8.log | cut -c1-400
fi
