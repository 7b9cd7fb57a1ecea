#!/bin/bash
# Session baseline: GPU tests, smoke, default bench (C2), reference arm, C1 and C4 lines.
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
python bench.py --steps 20 --warmup 5 > gpurun_out/r2e_c2.json 2> gpurun_out/r2e_c2.err; tail -c 2500 gpurun_out/r2e_c2.json
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2e_ref_c2.json 2> gpurun_out/r2e_ref_c2.err; cat gpurun_out/r2e_ref_c2.json
python bench.py --config c1 --steps 10 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2e_c1.json 2> gpurun_out/r2e_c1.err; tail -c 1500 gpurun_out/r2e_c1.json
python bench.py --config c4 --steps 10 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2e_c4.json 2> gpurun_out/r2e_c4.err; tail -c 1500 gpurun_out/r2e_c4.json
nproc; lscpu | grep -i "model name"
