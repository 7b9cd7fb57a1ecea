#!/bin/bash
# kernel time vs batch size (intercept = fixed cost, slope = per-byte cost)
for cfg in ${CFGS:-c2 c1}; do
for mb in ${SIZES:-1024 4096 16384 65536 262144}; do
python bench.py --config $cfg --bytes $((mb<<10)) --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null > /tmp/b.json
python - <<PY
import json
d=json.load(open('/tmp/b.json'))
r=d['roofline']
print("$cfg", $mb, "KiB kernel_ms",r['kernel_ms'],'ms/step',d['ms_per_step'],'achieved',r['achieved'])
PY
done; done
