#!/bin/bash
for u in 448 600 1024 2048; do
RETH_GATHER_UNITS=$u timeout 300 python bench.py --no-cpu --steps 6 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('units $u: value %.2fM e2e %.2fM roofline %.3f launch_us %.1f whole %.0f' % (d['value']/1e6, d['e2e']['value']/1e6, d['roofline']['frac'], d['roofline']['avg_launch_us'], d['hbm_gbs_whole_loop']))"
done
