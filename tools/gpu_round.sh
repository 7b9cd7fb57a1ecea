#!/bin/bash
# Round evidence on one B200: tests, bench (ours + reference arm), ncu launch list, ncu full capture of k_turbo.
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 40 --warmup 5 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; tail -c 3000 gpurun_out/bench_c2.json
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_c2.json 2> gpurun_out/bench_ref_c2.err; cat gpurun_out/bench_ref_c2.json
python bench.py --config c1 --bytes $((256<<20)) --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; tail -c 1500 gpurun_out/bench_c1.json
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_turbo -s 4 -c 1 -o gpurun_out/prof_c2_r01 -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c2.log 2>&1
python bench.py --config c1 --bytes $((256<<20)) --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c1.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_turbo -s 4 -c 1 -o gpurun_out/prof_c1_r01 -f python bench.py --config c1 --bytes $((256<<20)) --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c1.log 2>&1
nproc; lscpu | grep -i "model name\|^CPU(s)"
