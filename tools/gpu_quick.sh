#!/bin/bash
# quick GPU check: parity tests, then C2 and C1 bench lines (compact)
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for cfg in c2 c1; do
python bench.py --config $cfg --steps 20 --warmup 3 --no-cpu-baseline 2>gpurun_out/bench_$cfg.err > gpurun_out/bench_$cfg.json
python - <<PY
import json
d=json.load(open('gpurun_out/bench_$cfg.json'))
r=d['roofline']
print('$cfg', 'value',d['value'],'ms/step',d['ms_per_step'],'kernel_ms',r['kernel_ms'],'achieved',r['achieved'],'frac',r['frac'],'e2e',d['e2e']['value'],'cand/kb',d['config'].get('candidates_per_kb'),'probes/kb',d['config'].get('confirm_probes_per_kb'),'matches',d['matches_per_step'])
PY
done
