#!/bin/bash
# multi-outfix full-Rose databases: parity, then C3 at 2000 regexes (ours + reference arm)
python -m pytest tests/test_gpu_rose.py tests/test_gpu_ext.py -m gpu -x -q 2>&1 | tail -5
for ds in 1024 256; do
HS_BENCH_DFA_STATES=$ds timeout 900 python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2f_c3_$ds.json 2> gpurun_out/r2f_c3_$ds.err; tail -c 1800 gpurun_out/r2f_c3_$ds.json; tail -3 gpurun_out/r2f_c3_$ds.err
done
timeout 600 python bench.py --config c3 --impl reference --steps 2 --warmup 1 > gpurun_out/r2f_ref_c3.json 2> gpurun_out/r2f_ref_c3.err; cat gpurun_out/r2f_ref_c3.json; tail -3 gpurun_out/r2f_ref_c3.err
