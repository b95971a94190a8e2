#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_board_gpu.py tests/test_multi_gpu.py tests/test_nstep_gpu.py tests/test_frames_gpu.py -x -q > gpurun_out/r02e_pytest.log 2>&1; tail -4 gpurun_out/r02e_pytest.log
: > gpurun_out/r02e_modeg.jsonl
for D in 2 3 4; do
  DEPTH=$D CAP=262144 ITERS=300 CASES=1024:8,4096:8,1024:1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/modeg_group.py 2>gpurun_out/r02e_modeg.err | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); d['depth']=$D; d.pop('note',None); print(json.dumps(d))
    print('depth=$D', d['global_batch'], d['prefetch_group'], round(d['us_per_global_batch'],2), round(d['transitions_per_s']/1e6,1), file=sys.stderr)
" >> gpurun_out/r02e_modeg.jsonl
done
tail -3 gpurun_out/r02e_modeg.err
python scripts/sparse_gather.py 2>&1 | tail -5
