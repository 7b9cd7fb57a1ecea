#!/bin/bash
# round-2 closing evidence: full GPU suite, smoke, default bench + reference arm, C1 at its stated 1 GiB
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2v_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2v_smoke.log 2>&1
python bench.py --impl reference > gpurun_out/r2v_bench_ref_c2.json 2> gpurun_out/r2v_bench_ref.err
python bench.py > gpurun_out/r2v_bench_c2.json 2> gpurun_out/r2v_bench.err
python bench.py --config c1 --bytes $((1<<30)) --steps 10 > gpurun_out/r2v_bench_c1_1GiB.json 2> gpurun_out/r2v_bench_c1.err
tail -3 gpurun_out/r2v_pytest.log gpurun_out/r2v_smoke.log gpurun_out/r2v_bench.err gpurun_out/r2v_bench_c1.err
