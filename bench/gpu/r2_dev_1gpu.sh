#!/bin/bash
# Round 2 development loop on one B200: GPU test tier, default bench, the pure CUDA-core path for comparison,
# warm-cache per-kernel breakdown of the default step.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -n 15 gpurun_out/pytest_gpu.log | cut -c1-400
for v in "auto auto" "fp32 simt" "$EXTRA"; do
  [ -z "$v" ] && continue
  set -- $v
  timeout 300 python bench.py --steps 1000 --warmup 100 --dtype $1 --conv $2 > gpurun_out/bench_$1_$2.log 2>&1
  echo "bench dtype=$1 conv=$2 rc=$?"; grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_$1_$2.log | head -2
  tail -n 3 gpurun_out/bench_$1_$2.log | grep -v '^{' | cut -c1-300
done
timeout 300 python bench/step_breakdown.py --config lenet_b256 > gpurun_out/breakdown_default.log 2>&1
grep -v Warning gpurun_out/breakdown_default.log | tail -25
