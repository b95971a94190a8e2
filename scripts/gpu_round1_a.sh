#!/bin/bash
# First GPU call of a session: parity tests, bench, ncu launch list, ncu full capture of the gather.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
export CAP=262144 STEPS=48 GROUP=8
python scripts/prof_step.py > gpurun_out/prof_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"per_|row_move|ring_append" -s 48 -c 18 --csv --log-file gpurun_out/launches.csv python scripts/prof_step.py > gpurun_out/ncu_launches.log 2>&1
python scripts/prof_step.py > gpurun_out/prof_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:row_move -s 2 -c 2 -o gpurun_out/prof_gather python scripts/prof_step.py > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
