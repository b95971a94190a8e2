#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/e2e_mp.jsonl
DEFAULTS=1 timeout 200 python scripts/e2e_mp.py 2>gpurun_out/e2e_mp.err | tail -1 | tee -a gpurun_out/e2e_mp.jsonl
for ga in "8 2" "16 2" "16 3"; do
  set -- $ga
  GROUP=$1 AHEAD=$2 timeout 200 python scripts/e2e_mp.py 2>gpurun_out/e2e_mp.err | tail -1 | tee -a gpurun_out/e2e_mp.jsonl
done
WIRE=rows DEFAULTS=1 timeout 200 python scripts/e2e_mp.py 2>gpurun_out/e2e_mp.err | tail -1 | tee -a gpurun_out/e2e_mp.jsonl
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
