#!/bin/bash
python -m pytest tests/test_gpu_rose.py -m gpu -x -q -k "packed or fused" 2>&1 | tail -3
for pk in 2 1; do
HS_GPU_PACK=$pk timeout 900 python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2r_c3_pk$pk.json 2> gpurun_out/r2r_c3_pk$pk.err; tail -c 230 gpurun_out/r2r_c3_pk$pk.json; echo
HS_GPU_PACK=$pk HS_BENCH_DFA_STATES=256 timeout 900 python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2r_c3_d256_pk$pk.json 2> gpurun_out/r2r_c3_d256_pk$pk.err; tail -c 230 gpurun_out/r2r_c3_d256_pk$pk.json; echo
done
