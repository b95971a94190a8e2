#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/r02l.jsonl
for G in 2 4; do for F in 1024 2048 4096; do
  echo "## group $G fast_n $F" >> gpurun_out/r02l.jsonl
  GROUP=$G RETH_UPDATE_FAST_N=$F VARIANTS=-1:512:0 STEPS=3 python scripts/footprint_sweep.py >> gpurun_out/r02l.jsonl 2>> gpurun_out/r02l.err
done; done
echo "## group 8 fast_n 2048" >> gpurun_out/r02l.jsonl
GROUP=8 VARIANTS=-1:512:0 STEPS=3 python scripts/footprint_sweep.py >> gpurun_out/r02l.jsonl 2>> gpurun_out/r02l.err
python - <<'PY'
import json
for l in open('gpurun_out/r02l.jsonl'):
    if l.startswith('#'): print(l.strip()); continue
    d=json.loads(l); print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ('value_Mps','gather_frac','strict_Mps','tree_only_us_per_group','gather_only_Mps')})
PY
tail -3 gpurun_out/r02l.err
