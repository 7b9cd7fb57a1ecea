#!/bin/bash
# k_mcclellan: CTA size sweep at 64 MiB, then 256 MiB (C3's corpus size)
for b in 128 256 512 1024; do
  echo "block $b"; HS_GPU_MCC_BLOCK=$b python bench.py --config c3d --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | grep -o '"value": [0-9.]*, "unit": "GB/s", "n_gpus"\|"kernel_ms": [0-9.]*'
done
echo "auto 64MiB"; python bench.py --config c3d --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | grep -o '"value": [0-9.]*, "unit": "GB/s", "n_gpus"\|"kernel_ms": [0-9.]*'
for b in 512 1024; do
  echo "256MiB block $b"; HS_GPU_MCC_BLOCK=$b python bench.py --config c3d --bytes $((256<<20)) --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | grep -o '"value": [0-9.]*, "unit": "GB/s", "n_gpus"\|"kernel_ms": [0-9.]*'
done
