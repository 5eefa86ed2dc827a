#!/bin/bash
for f in test_supcon_gpu test_supcon_parity_gpu test_syncbn_gpu test_peer_gpu; do timeout 600 python -m pytest tests/$f.py -m gpu -q 2>&1 | tail -12 > gpurun_out/r2d_$f.log; done
python -c "
from simclr_pytorch_distributed_b200 import _lib
import torch; torch.cuda.init()
lib=_lib.load()
print('cluster occupancy', {c: lib.scb_bn_cluster_occupancy(c) for c in range(1,17)})
" > gpurun_out/r2d_occ.log 2>&1
for d in 0 1 8 9; do SCB_DBG=$d timeout 120 python tools/time_loss.py 2>&1 | grep SCB_DBG; done > gpurun_out/r2d_tl.log 2>&1
timeout 120 python tools/time_bn.py > gpurun_out/r2d_tb.log 2>&1
timeout 120 python tools/time_bn.py 512 1024 8 8 >> gpurun_out/r2d_tb.log 2>&1
timeout 120 python tools/time_bn.py 512 64 32 32 >> gpurun_out/r2d_tb.log 2>&1
timeout 400 python tools/bn_census.py --iters 10 > gpurun_out/r2d_census.log 2>&1
