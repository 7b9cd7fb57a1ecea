#!/bin/bash
python -m pytest tests/test_gpu_rose.py tests/test_gpu_ext.py tests/test_golden_rose.py -m gpu -x -q -k "packed or fused or multi_outfix or c3_shaped or mixed or recorded" 2>&1 | tail -4
for pk in 2 1; do
HS_GPU_PACK=$pk timeout 900 python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2p_c3_pk$pk.json 2> gpurun_out/r2p_c3_pk$pk.err; tail -c 420 gpurun_out/r2p_c3_pk$pk.json; tail -2 gpurun_out/r2p_c3_pk$pk.err
done
timeout 900 python bench.py --config c3 --bytes $((64<<20)) --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2p_c3_64.json 2> gpurun_out/r2p_c3_64.err; tail -c 420 gpurun_out/r2p_c3_64.json
HS_BENCH_DFA_STATES=256 timeout 900 python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2p_c3_d256.json 2> gpurun_out/r2p_c3_d256.err; tail -c 420 gpurun_out/r2p_c3_d256.json
HS_BENCH_DFA_STATES=512 timeout 900 python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2p_c3_d512.json 2> gpurun_out/r2p_c3_d512.err; tail -c 420 gpurun_out/r2p_c3_d512.json
timeout 900 python bench.py --config c4 --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2p_c4.json 2> gpurun_out/r2p_c4.err; tail -c 420 gpurun_out/r2p_c4.json
