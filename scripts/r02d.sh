#!/bin/bash
# 2-GPU pass: board / multi-GPU / parity tests, then the Mode G loop with the exchange folded into the chain
# (RETH_MODEG_FUSED) and the share-gather grid slack (RETH_SHARE_GRID) toggled.
mkdir -p gpurun_out
python -m pytest tests/test_board_gpu.py tests/test_multi_gpu.py tests/test_gpu_parity.py -x -q > gpurun_out/r02d_pytest.log 2>&1; tail -4 gpurun_out/r02d_pytest.log
: > gpurun_out/r02d_modeg.jsonl
for F in 1 0; do for SG in 1.25 2.0; do
  RETH_MODEG_FUSED=$F RETH_SHARE_GRID=$SG CAP=262144 ITERS=300 CASES=1024:1,1024:8,4096:1,4096:8 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/modeg_group.py 2>gpurun_out/r02d_modeg.err | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); d['fused']=$F; d['share_grid']=$SG; d.pop('note',None); print(json.dumps(d))
    print('fused=$F grid=$SG', d['global_batch'], d['prefetch_group'], round(d['us_per_global_batch'],2), round(d['transitions_per_s']/1e6,1), file=sys.stderr)
" >> gpurun_out/r02d_modeg.jsonl
done; done
tail -3 gpurun_out/r02d_modeg.err
