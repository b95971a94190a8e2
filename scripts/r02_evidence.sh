#!/bin/bash
# Round-2 evidence pass on one GPU (every step under its own timeout): ncu launch list of the bench command, ncu --set full
# of the gather at the BENCH configuration (capacity 2^20, prefetch group 8), of the TMA bulk-copy gather, of the tree
# kernels.  (compute-sanitizer is closed on this pool: profiles/r02_sanitizer_unavailable.log.)
mkdir -p gpurun_out
B="bench.py --steps 2 --warmup 3 --iters 128 --no-cpu --no-process"
timeout 200 python $B > gpurun_out/r02_bench_plain.json 2> gpurun_out/r02_bench_plain.err; echo "plain bench rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"per_|row_move|ring_append|sampler_|global_quota" -s 336 -c 96 --csv \
    --log-file gpurun_out/r02_launches_bench.csv python $B > gpurun_out/r02_ncu_launches.log 2>&1; echo "launch list rc=$?"
export CAP=1048576 STEPS=48 GROUP=8
timeout 120 python scripts/prof_step.py > gpurun_out/r02_prof_plain.log 2>&1; echo "prof_step rc=$?"
timeout 500 ncu --set full --clock-control none --import-source on -k regex:"row_move" -s 2 -c 3 \
    -o gpurun_out/r02_gather_full python scripts/prof_step.py > gpurun_out/r02_ncu_gather.log 2>&1; echo "ncu gather rc=$?"
timeout 500 ncu --set full --clock-control none --import-source on -k regex:"per_update|per_sample" -s 130 -c 4 \
    -o gpurun_out/r02_tree_full python scripts/prof_step.py > gpurun_out/r02_ncu_tree.log 2>&1; echo "ncu tree rc=$?"
RETH_GATHER_MODE=2 timeout 120 python scripts/prof_step.py > gpurun_out/r02_prof_bulk_plain.log 2>&1; echo "bulk prof_step rc=$?"
RETH_GATHER_MODE=2 timeout 500 ncu --set full --clock-control none --import-source on -k regex:"row_gather_bulk" -s 2 -c 2 \
    -o gpurun_out/r02_bulk_full python scripts/prof_step.py > gpurun_out/r02_ncu_bulk.log 2>&1; echo "ncu bulk rc=$?"
ls -la gpurun_out | grep r02_ | tail -20
