#!/bin/bash
# two GPUs: multi-GPU tests, then the bench under torchrun
timeout 900 python -m pytest tests/test_multigpu.py -x -q -m gpu 2>&1 | tail -12 > gpurun_out/g2_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 10 > gpurun_out/g2_bench.log 2> gpurun_out/g2_bench.err
