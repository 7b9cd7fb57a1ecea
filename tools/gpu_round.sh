#!/bin/bash
# Round evidence on one B200: tests, smoke, bench (ours + reference arm) for c2 / c3 / c3d, ncu launch lists and full captures.
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
python bench.py --steps 40 --warmup 5 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; tail -c 1500 gpurun_out/bench_c2.json
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_c2.json 2> gpurun_out/bench_ref_c2.err; cat gpurun_out/bench_ref_c2.json
python bench.py --config c3 --steps 10 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; tail -c 1200 gpurun_out/bench_c3.json
python bench.py --config c3d --steps 20 --warmup 3 > gpurun_out/bench_c3d.json 2> gpurun_out/bench_c3d.err; tail -c 1200 gpurun_out/bench_c3d.json
python bench.py --config c3d --bytes $((256<<20)) --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3d_256MiB.json 2> gpurun_out/bench_c3d_256MiB.err; tail -c 700 gpurun_out/bench_c3d_256MiB.json
python bench.py --config c3d --patterns 8 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3d_p8.json 2> gpurun_out/bench_c3d_p8.err; tail -c 600 gpurun_out/bench_c3d_p8.json
python bench.py --config c3d --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_c3d.json 2> gpurun_out/bench_ref_c3d.err; cat gpurun_out/bench_ref_c3d.json
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
python bench.py --config c3d --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_c3d.csv python bench.py --config c3d --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l_c3d.log 2>&1
python bench.py --config c3d --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_c3d.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_mcclellan -s 4 -c 1 -o gpurun_out/prof_c3d_r01 -f python bench.py --config c3d --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c3d.log 2>&1
nproc; lscpu | grep -i "model name\|^CPU(s)"
