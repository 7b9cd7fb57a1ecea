#!/bin/bash
python -m pytest tests/test_gpu_accel.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python tools/accel_bench.py > gpurun_out/r2l_accel.jsonl 2> gpurun_out/r2l_accel.err; grep 268435456 gpurun_out/r2l_accel.jsonl | grep "no match" | cut -c1-150; tail -3 gpurun_out/r2l_accel.err
python tools/accel_bench.py 268435456 > gpurun_out/plain_cs.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_class_scan -s 4 -c 1 -o gpurun_out/prof_class_scan -f python tools/accel_bench.py 268435456 > gpurun_out/ncu_cs.log 2>&1; tail -1 gpurun_out/ncu_cs.log
timeout 900 python tools/sweep_c5.py --regex --patterns 100 300 1000 2000 --bufsizes 256 1500 16384 262144 --bytes $((64<<20)) --steps 3 > gpurun_out/r2l_c5_regex.jsonl 2> gpurun_out/r2l_c5_regex.err; cut -c1-330 gpurun_out/r2l_c5_regex.jsonl | tail -16; tail -3 gpurun_out/r2l_c5_regex.err
