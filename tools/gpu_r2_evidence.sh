#!/bin/bash
# Round-2 evidence on one B200: full GPU suite, smoke, bench lines of every config (ours + reference arm), byte-class
# microbench, C5 sweeps, ncu launch lists + full captures of the literal kernels.
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
python bench.py --steps 20 --warmup 5 > gpurun_out/ev_c2.json 2> gpurun_out/ev_c2.err; tail -c 600 gpurun_out/ev_c2.json
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/ev_ref_c2.json 2> gpurun_out/ev_ref_c2.err; tail -c 400 gpurun_out/ev_ref_c2.json
for cfg in c1 c4l c3d; do python bench.py --config $cfg --steps 10 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/ev_$cfg.json 2> gpurun_out/ev_$cfg.err; tail -c 500 gpurun_out/ev_$cfg.json; done
python bench.py --config c1 --bytes $((256<<20)) --steps 10 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/ev_c1_256MiB.json 2> gpurun_out/ev_c1_256MiB.err; tail -c 500 gpurun_out/ev_c1_256MiB.json
for cfg in c3 c4; do timeout 900 python bench.py --config $cfg --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ev_$cfg.json 2> gpurun_out/ev_$cfg.err; tail -c 500 gpurun_out/ev_$cfg.json; done
timeout 600 python tools/accel_bench.py > gpurun_out/ev_accel.jsonl 2> gpurun_out/ev_accel.err; grep 268435456 gpurun_out/ev_accel.jsonl | grep "no match" | cut -c1-140
timeout 900 python tools/sweep_c5.py --bytes $((128<<20)) --steps 5 > gpurun_out/ev_c5_literal.jsonl 2> gpurun_out/ev_c5_literal.err; grep -c '"parity": "ok"' gpurun_out/ev_c5_literal.jsonl; grep -c MISMATCH gpurun_out/ev_c5_literal.jsonl
timeout 900 python tools/sweep_c5.py --regex --patterns 100 300 1000 2000 --bufsizes 256 1500 16384 --bytes $((64<<20)) --steps 3 > gpurun_out/ev_c5_regex.jsonl 2> gpurun_out/ev_c5_regex.err; grep -c '"parity": "ok"' gpurun_out/ev_c5_regex.jsonl; grep -c MISMATCH gpurun_out/ev_c5_regex.jsonl
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ev_launches_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_turbo -s 40 -c 1 -o gpurun_out/ev_prof_c2 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/ncu_c2.log 2>&1; tail -1 gpurun_out/ncu_c2.log
python bench.py --config c1 --steps 2 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/plain1.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_turbo2 -s 40 -c 1 -o gpurun_out/ev_prof_c1 -f python bench.py --config c1 --steps 2 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/ncu_c1.log 2>&1; tail -1 gpurun_out/ncu_c1.log
nproc; lscpu | grep -i "model name"
