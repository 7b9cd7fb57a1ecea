#!/bin/bash
# two GPUs of one box: in-process two-device tests, then the default bench under torchrun (verify leg: sharded == single)
python -m pytest tests/test_gpu_multidevice.py -m gpu -x -q 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2n_c2_n2.json 2> gpurun_out/r2n_c2_n2.err; tail -c 2200 gpurun_out/r2n_c2_n2.json; tail -3 gpurun_out/r2n_c2_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --impl reference --steps 3 --warmup 1 > gpurun_out/r2n_ref_n2.json 2> gpurun_out/r2n_ref_n2.err; cat gpurun_out/r2n_ref_n2.json | cut -c1-300
