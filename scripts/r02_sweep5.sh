#!/bin/bash
N=${1:-8}; mkdir -p gpurun_out
OUT=gpurun_out/r02_sweep_config5_modeg_${N}gpu.jsonl; : > $OUT
OUT=$OUT timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 scripts/sweep_config5_modeg.py 2> gpurun_out/r02_sweep5_${N}.err | tail -14
tail -2 gpurun_out/r02_sweep5_${N}.err
