#!/bin/bash
# McClellan DFA evidence on one B200: tests, bench (6-pattern table in shared memory, 8-pattern table in L2), reference arm, ncu.
set -x
timeout 600 python -m pytest tests/test_gpu_mcclellan.py -m gpu -x -q 2>&1 | tail -3
python bench.py --config c3d --steps 20 --warmup 3 > gpurun_out/bench_c3d.json 2> gpurun_out/bench_c3d.err; tail -c 2200 gpurun_out/bench_c3d.json; tail -3 gpurun_out/bench_c3d.err
python bench.py --config c3d --patterns 8 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3d_p8.json 2> gpurun_out/bench_c3d_p8.err; tail -c 1500 gpurun_out/bench_c3d_p8.json; tail -3 gpurun_out/bench_c3d_p8.err
python bench.py --config c3d --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_c3d.json 2> gpurun_out/bench_ref_c3d.err; cat gpurun_out/bench_ref_c3d.json
python bench.py --config c3d --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c3d.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_mcclellan -s 4 -c 1 -o gpurun_out/prof_c3d_r01 -f python bench.py --config c3d --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c3d.log 2>&1
python bench.py --config c3d --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_c3d.csv python bench.py --config c3d --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l_c3d.log 2>&1
ls -la gpurun_out | tail -12
