#!/bin/bash
python -m pytest tests/test_gpu_accel.py tests/test_hsbench_cli.py -m gpu -x -q 2>&1 | tail -4
timeout 600 python tools/accel_bench.py > gpurun_out/r2i_accel.jsonl 2> gpurun_out/r2i_accel.err; grep 268435456 gpurun_out/r2i_accel.jsonl | cut -c1-260; tail -3 gpurun_out/r2i_accel.err
ncu --set full --clock-control none --import-source on -k regex:k_class_scan -s 30 -c 1 -o gpurun_out/prof_class_scan -f python tools/accel_bench.py > gpurun_out/ncu_cs.log 2>&1; tail -2 gpurun_out/ncu_cs.log
