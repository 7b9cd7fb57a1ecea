#!/bin/bash
# round-2 last pass: Sherman engines stepped in the reference's compressed form vs expanded at upload (c3d set, 8 patterns so
# that Sherman states exist), then the full GPU suite
set -x
HS_BENCH_C3D_ENGINE=2 HS_GPU_SHERMAN_EXACT=1 python bench.py --config c3d --patterns 8 --steps 10 > gpurun_out/r2x_c3d_sherman_exact.json 2> gpurun_out/r2x_a.err
HS_BENCH_C3D_ENGINE=2 python bench.py --config c3d --patterns 8 --steps 10 > gpurun_out/r2x_c3d_sherman_expanded.json 2> gpurun_out/r2x_b.err
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2x_pytest.log
cat gpurun_out/r2x_pytest.log; tail -n 2 gpurun_out/r2x_a.err gpurun_out/r2x_b.err
