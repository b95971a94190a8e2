#!/bin/bash
# Mode G device loop on N GPUs: repair-kernel shape on the chain (RETH_MODEG_CRITICAL) and ring depth
N=${1:-8}; mkdir -p gpurun_out; : > gpurun_out/r02o_modeg.jsonl
for C in 1 0; do
  RETH_MODEG_CRITICAL=$C DEPTH=3 CAP=262144 ITERS=300 CASES=4096:8,16384:8,4096:1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 scripts/modeg_group.py 2>gpurun_out/r02o_modeg.err | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); d['critical']=$C; d.pop('note',None); print(json.dumps(d))
    print('critical=$C', d['global_batch'], d['prefetch_group'], round(d['us_per_global_batch'],2), round(d['transitions_per_s']/1e6,1), file=sys.stderr)
" >> gpurun_out/r02o_modeg.jsonl
done
tail -2 gpurun_out/r02o_modeg.err
