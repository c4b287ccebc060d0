#!/bin/bash
# Round 2, call 0: re-measure the reference LeNet on the CURRENT tensor-core kernels (warm-cache per-kernel times
# and the graph-replayed step) against the CUDA-core path -- the decision behind dtype="auto".
mkdir -p gpurun_out
for mode in "fp32 simt" "tf32 simt" "fp32 tc" "tf32 tc"; do
  set -- $mode
  echo "=== dtype $1 conv $2 ===" >> gpurun_out/r2_breakdown.log
  timeout 300 python bench/step_breakdown.py --config lenet_b256 --dtype $1 --conv $2 >> gpurun_out/r2_breakdown.log 2>&1
  echo "rc=$?" >> gpurun_out/r2_breakdown.log
  timeout 300 python bench.py --steps 1000 --warmup 100 --dtype $1 --conv $2 > gpurun_out/r2_bench_$1_$2.log 2>&1
  echo "bench $1 $2 rc=$?"; grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_bench_$1_$2.log | head -2
done
cat gpurun_out/r2_breakdown.log | grep -v Warning | tail -80
