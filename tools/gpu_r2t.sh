#!/bin/bash
python -m pytest tests/test_gpu_rose.py tests/test_gpu_mcclellan.py -m gpu -x -q 2>&1 | tail -3
python bench.py --config c3d --steps 10 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2t_c3d.json 2> gpurun_out/r2t_c3d.err; tail -c 200 gpurun_out/r2t_c3d.json; echo
timeout 900 python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2t_c3.json 2> gpurun_out/r2t_c3.err; tail -c 200 gpurun_out/r2t_c3.json; echo
timeout 900 python bench.py --config c4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2t_c4.json 2> gpurun_out/r2t_c4.err; tail -c 200 gpurun_out/r2t_c4.json; echo
python bench.py --config c3n --steps 3 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2t_c3n.json 2> gpurun_out/r2t_c3n.err; tail -c 200 gpurun_out/r2t_c3n.json; echo
timeout 900 python tools/sweep_c5.py --regex --patterns 100 300 1000 2000 --bufsizes 256 1500 16384 --bytes $((64<<20)) --steps 3 > gpurun_out/r2t_c5_regex.jsonl 2> gpurun_out/r2t_c5_regex.err; grep -c '"parity": "ok"' gpurun_out/r2t_c5_regex.jsonl
