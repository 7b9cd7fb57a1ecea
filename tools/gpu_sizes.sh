#!/bin/bash
# kernel time vs batch size (intercept = fixed cost, slope = per-byte cost)
for cfg in ${CFGS:-c2 c1}; do
for mb in ${SIZES:-1 4 16 64 256}; do
python bench.py --config $cfg --bytes $((mb<<20)) --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null > /tmp/b.json
python - <<PY
import json
d=json.load(open('/tmp/b.json'))
r=d['roofline']
print('$cfg', $mb, 'MiB kernel_ms',r['kernel_ms'],'ms/step',d['ms_per_step'],'achieved',r['achieved'])
PY
done; done
