#!/bin/bash
python -m pytest tests/test_hsbench_cli.py tests/test_gpu_parity.py -m gpu -x -q -k "cli or singlematch" 2>&1 | tail -3
python tools/accel_bench.py 268435456 > gpurun_out/plain_cs.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_class_scan -s 4 -c 1 -o gpurun_out/prof_class_scan -f python tools/accel_bench.py 268435456 > gpurun_out/ncu_cs.log 2>&1; tail -2 gpurun_out/ncu_cs.log
