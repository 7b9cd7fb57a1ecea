#!/bin/bash
python -m pytest tests/test_gpu_accel.py tests/test_shim.py -m gpu -x -q 2>&1 | tail -4
timeout 600 python tools/accel_bench.py > gpurun_out/r2h_accel.jsonl 2> gpurun_out/r2h_accel.err; tail -12 gpurun_out/r2h_accel.jsonl; tail -3 gpurun_out/r2h_accel.err
