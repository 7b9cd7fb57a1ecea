#!/bin/bash
python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
for p in 0 2; do
  echo "== pipeline $p"
  HS_GPU_TRACE=1 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-verify --pipeline $p > gpurun_out/p$p.json 2> gpurun_out/p$p.err
  python - <<EOF
import json
l=json.loads(open("gpurun_out/p$p.json").read().strip().splitlines()[-1])
print(json.dumps(l["e2e"])); print(l["value"], l["roofline"]["kernel_ms"])
EOF
  grep "hs_gpu trace" gpurun_out/p$p.err | sed -n 20,30p
done
