#!/bin/bash
python -m pytest tests/test_gpu_rose.py tests/test_gpu_ext.py -m gpu -x -q -k "packed or multi_outfix or c3_shaped or mixed" 2>&1 | tail -4
timeout 900 python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2m_c3_256.json 2> gpurun_out/r2m_c3_256.err; tail -c 700 gpurun_out/r2m_c3_256.json; tail -2 gpurun_out/r2m_c3_256.err
timeout 900 python bench.py --config c3 --bytes $((64<<20)) --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2m_c3_64.json 2> gpurun_out/r2m_c3_64.err; tail -c 700 gpurun_out/r2m_c3_64.json
timeout 900 python bench.py --config c4 --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2m_c4.json 2> gpurun_out/r2m_c4.err; tail -c 700 gpurun_out/r2m_c4.json
