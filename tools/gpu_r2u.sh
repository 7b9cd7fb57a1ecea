#!/bin/bash
# four GPUs of one box under torchrun: C2 (default), C3, C4 -- weak scaling, sharded-equals-single verify leg on hardware
for cfg in c2 c3 c4; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 4 --config $cfg --steps $([ $cfg = c2 ] && echo 20 || echo 3) --warmup 3 --no-cpu-baseline > gpurun_out/r2u_${cfg}_n4.json 2> gpurun_out/r2u_${cfg}_n4.err; tail -c 300 gpurun_out/r2u_${cfg}_n4.json; echo; tail -2 gpurun_out/r2u_${cfg}_n4.err
done
