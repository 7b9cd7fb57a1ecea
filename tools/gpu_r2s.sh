#!/bin/bash
python -m pytest tests/test_gpu_limex.py tests/test_gpu_mcclellan.py tests/test_gpu_sheng.py tests/test_gpu_eod.py tests/test_gpu_rose.py tests/test_golden_rose.py -m gpu -x -q 2>&1 | tail -3
for ord in 1 0; do
HS_GPU_ORDER=$ord python bench.py --config c3d --steps 10 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2s_c3d_o$ord.json 2> gpurun_out/r2s_c3d_o$ord.err; tail -c 200 gpurun_out/r2s_c3d_o$ord.json; echo
done
for pk in 1 2; do
HS_GPU_PACK=$pk timeout 900 python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2s_c3_pk$pk.json 2> gpurun_out/r2s_c3_pk$pk.err; tail -c 200 gpurun_out/r2s_c3_pk$pk.json; echo
done
HS_GPU_PACK=1 timeout 900 python bench.py --config c3 --bytes $((64<<20)) --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2s_c3_64.json 2> gpurun_out/r2s_c3_64.err; tail -c 200 gpurun_out/r2s_c3_64.json; echo
HS_GPU_PACK=1 timeout 900 python bench.py --config c4 --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2s_c4.json 2> gpurun_out/r2s_c4.err; tail -c 200 gpurun_out/r2s_c4.json; echo
python bench.py --config c3n --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2s_c3n.json 2> gpurun_out/r2s_c3n.err; tail -c 200 gpurun_out/r2s_c3n.json; echo
