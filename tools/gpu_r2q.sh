#!/bin/bash
python -m pytest tests/test_gpu_rose.py -m gpu -x -q -k "packed or fused" 2>&1 | tail -4
python bench.py --config c3 --bytes $((64<<20)) --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/plain_q.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size,smsp__inst_executed.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,launch__registers_per_thread,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"k_dfa_pack|k_mcclellan" -s 170 -c 60 --csv --log-file gpurun_out/r2q_launches.csv python bench.py --config c3 --bytes $((64<<20)) --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/ncu_q.log 2>&1; tail -2 gpurun_out/ncu_q.log
