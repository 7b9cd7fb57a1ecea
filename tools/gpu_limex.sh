#!/bin/bash
# LimEx bring-up on one B200: parity tests, then the c3 bench line.
set -x
timeout 900 python -m pytest tests/test_gpu_limex.py -m gpu -x -q 2>&1 | tail -15
timeout 600 python bench.py --config c3 --steps 10 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; tail -c 2500 gpurun_out/bench_c3.json; tail -5 gpurun_out/bench_c3.err
