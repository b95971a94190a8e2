#!/bin/bash
# e2e after the issue-before-wait change: host time profile + loader geometry sweep
mkdir -p gpurun_out
for ga in "8 1" "8 2" "16 1" "4 1"; do
  set -- $ga
  timeout 120 python scripts/e2e_profile.py $1 $2 2>/dev/null | tail -1 | tee -a gpurun_out/e2e_profile.jsonl
done
timeout 200 python -m pytest tests/test_pipe_gpu.py tests/test_service_gpu.py -x -q -m gpu 2>&1 | tail -3
timeout 300 python bench.py --no-cpu 2>/dev/null | tee gpurun_out/bench_next.json | cut -c1-600
