#!/bin/bash
# round 2, first check: parity of all front ends, then C2 / C1 / C4 with k_turbo2 (0) against k_turbo (2)
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for cfg in c2 c1 c4; do for fe in 0 2; do
BYTES=$((64<<20)); [ $cfg = c1 ] && BYTES=$((256<<20)); [ $cfg = c4 ] && BYTES=$((256<<20))
python bench.py --config $cfg --frontend $fe --bytes $BYTES --steps 20 --warmup 3 --no-cpu-baseline 2>gpurun_out/bench_${cfg}_fe$fe.err > gpurun_out/bench_${cfg}_fe$fe.json
python - <<PY
import json
try:
    d=json.load(open('gpurun_out/bench_${cfg}_fe$fe.json'))
    r=d['roofline']
    print('$cfg fe$fe', 'value',d['value'],'ms/step',d['ms_per_step'],'kernel_ms',r['kernel_ms'],'achieved',r['achieved'],'frac',r['frac'],'e2e',d['e2e']['value'],d['diag'],'matches',d['matches_per_step'])
except Exception as e:
    print('$cfg fe$fe FAILED', e); print(open('gpurun_out/bench_${cfg}_fe$fe.err').read()[-1500:])
PY
done; done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_c2.json 2> gpurun_out/bench_ref_c2.err; cat gpurun_out/bench_ref_c2.json
nproc; lscpu | grep -i "model name\|^CPU(s)\|numa"
