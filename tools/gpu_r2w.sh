#!/bin/bash
# round-2 closing evidence, second pass (after stream-mode DFA engines): full GPU suite, smoke, c3d / c3 / c4 bench lines
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2w_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2w_smoke.log 2>&1
python bench.py --config c3d > gpurun_out/r2w_bench_c3d.json 2> gpurun_out/r2w_c3d.err
python bench.py --config c3 --steps 3 > gpurun_out/r2w_bench_c3.json 2> gpurun_out/r2w_c3.err
python bench.py --config c4 --steps 3 > gpurun_out/r2w_bench_c4.json 2> gpurun_out/r2w_c4.err
cat gpurun_out/r2w_pytest.log; tail -n 2 gpurun_out/r2w_smoke.log gpurun_out/r2w_c3d.err gpurun_out/r2w_c3.err gpurun_out/r2w_c4.err
