#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_frames_gpu.py -x -q -m gpu 2>&1 | tail -25
timeout 300 python scripts/frame_bench.py 2>gpurun_out/frame_bench.err | tail -1 | tee gpurun_out/frame_bench.json
tail -3 gpurun_out/frame_bench.err
