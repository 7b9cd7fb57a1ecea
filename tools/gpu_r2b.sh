#!/bin/bash
python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -5
run() { # cfg bytes fe tag
python bench.py --config $1 --frontend $3 --bytes $2 --steps 20 --warmup 3 --no-cpu-baseline 2>gpurun_out/bench_$4.err > gpurun_out/bench_$4.json
python - <<PY
import json
try:
    d=json.load(open('gpurun_out/bench_$4.json')); r=d['roofline']
    print('$4', 'value',d['value'],'kernel_ms',r['kernel_ms'],'frac',r['frac'],'e2e',d['e2e']['value'],d['diag'],'matches',d['matches_per_step'])
except Exception as e:
    print('$4 FAILED', e); print(open('gpurun_out/bench_$4.err').read()[-1500:])
PY
}
run c2 $((64<<20)) 0 c2_fe0
run c2 $((256<<20)) 0 c2_256_fe0
HS_GPU_TURBO_WIDE=0 run c2 $((256<<20)) 0 c2_256_fe0_narrow
run c2 $((256<<20)) 2 c2_256_fe2
run c4 $((256<<20)) 0 c4_fe0
run c1 $((256<<20)) 0 c1_fe0
