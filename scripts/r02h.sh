#!/bin/bash
# lean gather (32 registers, CTA cap) x repair-kernel block size, headline loop on one GPU
mkdir -p gpurun_out; : > gpurun_out/r02h_footprint.jsonl
run() { echo "## $*" >> gpurun_out/r02h_footprint.jsonl; env "$@" STEPS=4 python scripts/footprint_sweep.py >> gpurun_out/r02h_footprint.jsonl 2>> gpurun_out/r02h.err; }
run RETH_GATHER_LEAN=0 VARIANTS=-1:512:0
run RETH_GATHER_LEAN=1 RETH_GATHER_CTAS=6 VARIANTS=-1:512:0,-1:256:0
run RETH_GATHER_LEAN=1 RETH_GATHER_CTAS=7 VARIANTS=-1:512:0,-1:256:0
run RETH_GATHER_LEAN=1 RETH_GATHER_CTAS=8 VARIANTS=-1:256:0
run RETH_GATHER_LEAN=1 RETH_GATHER_CTAS=5 VARIANTS=-1:512:0
run RETH_GATHER_LEAN=1 RETH_GATHER_CTAS=6 RETH_GATHER_UNITS=2048 VARIANTS=-1:256:0
python - <<'PY'
import json
for l in open('gpurun_out/r02h_footprint.jsonl'):
    if l.startswith('#'): print(l.strip()); continue
    d=json.loads(l); print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ('threads','value_Mps','gather_frac','strict_Mps','tree_only_us_per_group','gather_only_Mps','no_update_Mps')})
PY
tail -3 gpurun_out/r02h.err
