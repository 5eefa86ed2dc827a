#!/bin/bash
# N GPUs (N = $1): multi-GPU tests, then the bench under torchrun
N=${1:-2}
timeout 900 python -m pytest tests/test_multigpu.py -x -q -m gpu 2>&1 | grep -v "UserWarning\|Consider using\|e_l = abs" | tail -40 > gpurun_out/g${N}_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 50 --warmup 10 > gpurun_out/g${N}_bench.log 2> gpurun_out/g${N}_bench.err
