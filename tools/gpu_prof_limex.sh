#!/bin/bash
# ncu --set full capture of k_limex on the c3 workload (16 MiB batches keep the ~40 replays short)
set -x
CMD="python bench.py --config c3 --bytes $((16<<20)) --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_c3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_limex -s 4 -c 1 -o gpurun_out/prof_c3 -f $CMD > gpurun_out/ncu_c3.log 2>&1
tail -3 gpurun_out/ncu_c3.log
