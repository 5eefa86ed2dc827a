#!/bin/bash
# GPU session script (scratch): tests, loss timing toggles, BN timings
for f in test_supcon_gpu test_supcon_parity_gpu test_syncbn_gpu; do timeout 600 python -m pytest tests/$f.py -m gpu -q 2>&1 | tail -15 > gpurun_out/r2b_$f.log; done
for d in 0 1 3 7 8 16 23 31; do SCB_DBG=$d timeout 120 python tools/time_loss.py; done > gpurun_out/r2b_tl.log 2>&1
SCB_BWD_DIRECT=0 timeout 120 python tools/time_loss.py >> gpurun_out/r2b_tl.log 2>&1
SCB_PDL=0 timeout 120 python tools/time_loss.py >> gpurun_out/r2b_tl.log 2>&1
for c in 16 8 4 2 1; do SCB_BN_MAX_CLUSTER=$c timeout 120 python tools/time_bn.py; done > gpurun_out/r2b_tb.log 2>&1
SCB_BN_FUSED=0 timeout 120 python tools/time_bn.py >> gpurun_out/r2b_tb.log 2>&1
timeout 120 python tools/time_bn.py 512 512 4 4 >> gpurun_out/r2b_tb.log 2>&1
SCB_BN_FUSED=0 timeout 120 python tools/time_bn.py 512 1024 8 8 >> gpurun_out/r2b_tb.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:bn_ -s 8 -c 4 -o gpurun_out/r2b_bn_big python tools/time_bn.py 512 1024 8 8 > gpurun_out/r2b_ncu_bn.log 2>&1
