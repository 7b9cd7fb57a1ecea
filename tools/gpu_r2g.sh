#!/bin/bash
python -m pytest tests/test_gpu_ext.py -m gpu -x -q 2>&1 | tail -5
timeout 900 python bench.py --config c4 --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2g_c4.json 2> gpurun_out/r2g_c4.err; tail -c 1500 gpurun_out/r2g_c4.json; tail -3 gpurun_out/r2g_c4.err
timeout 600 python bench.py --config c4 --impl reference --steps 2 --warmup 1 > gpurun_out/r2g_ref_c4.json 2> gpurun_out/r2g_ref_c4.err; cat gpurun_out/r2g_ref_c4.json; tail -3 gpurun_out/r2g_ref_c4.err
