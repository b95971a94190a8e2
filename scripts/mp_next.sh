#!/bin/bash
mkdir -p gpurun_out
for m in 0 1; do
RETH_GATHER_WARP=$m timeout 300 python bench.py --no-cpu --steps 6 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('warp mode $m: value %.2fM e2e %.2fM roofline %.3f launch_us %.1f whole %.0f' % (d['value']/1e6, d['e2e']['value']/1e6, d['roofline']['frac'], d['roofline']['avg_launch_us'], d['hbm_gbs_whole_loop']))"
done
RETH_GATHER_WARP=1 timeout 300 python -m pytest tests/test_gpu_storage.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
