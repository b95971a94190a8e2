#!/bin/bash
# Mode G: NCCL root all-gather vs the roots board (peer stores over NVLink) vs the device loop.
# usage: scripts/modeg_board.sh N_GPUS [CAP]
N=${1:-2}; export CAP=${2:-262144}; export STEPS=${STEPS:-200}
mkdir -p gpurun_out
OUT=gpurun_out/modeG_board_${N}gpu.jsonl; : > $OUT
export CASES="local:4096,local:16384,local:65536"
for B in ${BOARDS:-0 1}; do
  BOARD=$B timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
      scripts/bench_assembly.py 2>gpurun_out/modeG_board_err_$B.log | grep '^{' >> $OUT
done
CASES="loop:4096,loop:16384,loop:65536" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port 29534 scripts/bench_assembly.py 2>gpurun_out/modeG_board_err_loop.log | grep '^{' >> $OUT
cat $OUT | cut -c1-260
