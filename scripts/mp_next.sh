#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_pipe_gpu.py tests/test_service_gpu.py tests/test_service_mp_gpu.py tests/test_samplers_gpu.py -x -q -m gpu 2>&1 | tail -6
for m in 0 1; do
RETH_PIPE_ZEROCOPY=$m timeout 300 python bench.py --no-cpu --steps 6 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('zero copy $m: value %.2fM e2e %.2fM strict_e2e %.2fM strict_dev %.2fM' % (d['value']/1e6, d['e2e']['value']/1e6, d['e2e']['strict_alternation']/1e6, d['strict_alternation']['value']/1e6))"
RETH_PIPE_ZEROCOPY=$m timeout 100 python scripts/e2e_profile.py 8 1 2>/dev/null | tail -1 | cut -c1-400
done
