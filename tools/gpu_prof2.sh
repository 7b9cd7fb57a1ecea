#!/bin/bash
# ncu full capture of the literal front end: tools/gpu_prof2.sh <cfg> <bytes> <frontend> <tag>
cfg=${1:-c2}; bytes=${2:-67108864}; fe=${3:-0}; tag=${4:-x}
python bench.py --config $cfg --bytes $bytes --frontend $fe --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_$cfg.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_turbo -s 4 -c 1 -o gpurun_out/prof_${cfg}_$tag -f python bench.py --config $cfg --bytes $bytes --frontend $fe --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_$cfg.log 2>&1
tail -1 gpurun_out/ncu_$cfg.log
