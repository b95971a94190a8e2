#!/bin/bash
# Evidence pass: parity tests, smoke, bench (+ reference arm), then ncu: launch list of bench.py itself and
# --set full captures of the gather, append, update and sample kernels (scripts/prof_step.py, group 8).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
# launch list of the bench command itself: skip the 192 fill launches (64 appends x (copy + 2 updates)),
# then 3 warm-up steps x 48 launches; capture two timed steps
python bench.py --steps 2 --warmup 3 --iters 128 --no-cpu > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"per_|row_move|ring_append|sampler_" -s 336 -c 96 --csv \
    --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 3 --iters 128 --no-cpu > gpurun_out/ncu_launches.log 2>&1
export CAP=262144 STEPS=48 GROUP=8
python scripts/prof_step.py > gpurun_out/prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"row_move|per_update|per_sample|ring_append" -s 48 -c 9 \
    -o gpurun_out/prof_all python scripts/prof_step.py > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -20
