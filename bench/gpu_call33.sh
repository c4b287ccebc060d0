#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench/sanitize_step.py > gpurun_out/sanitize_plain.log 2>&1 && \
timeout 900 compute-sanitizer --tool memcheck --log-file gpurun_out/sanitizer_memcheck.txt python bench/sanitize_step.py > gpurun_out/sanitize_memcheck_run.log 2>&1
echo "rc=$?"
tail -n 4 gpurun_out/sanitize_plain.log | cut -c1-200
tail -n 4 gpurun_out/sanitize_memcheck_run.log | cut -c1-200
tail -n 12 gpurun_out/sanitizer_memcheck.txt | cut -c1-200
