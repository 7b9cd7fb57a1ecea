#!/bin/bash
# A/B of build variants of the literal front end (temporary libraries vectorscan_b200/libhs_gpu_*.so)
run() { # cfg bytes tag
python bench.py --config $1 --bytes $2 --steps 20 --warmup 3 --no-cpu-baseline 2>gpurun_out/bench_$3.err > gpurun_out/bench_$3.json
python - <<PY
import json
try:
    d=json.load(open('gpurun_out/bench_$3.json')); r=d['roofline']
    print('$3', 'kernel_ms',r['kernel_ms'],'frac',r['frac'],'matches',d['matches_per_step'])
except Exception as e:
    print('$3 FAILED', e); print(open('gpurun_out/bench_$3.err').read()[-800:])
PY
}
for lib in v000 v100 v010 v001 v011 n000 n011 n111; do
export HS_GPU_LIB=$PWD/vectorscan_b200/libhs_gpu_$lib.so
run c4 $((256<<20)) c4_$lib
run c1 $((256<<20)) c1_$lib
run c2 $((256<<20)) c2_$lib
done
