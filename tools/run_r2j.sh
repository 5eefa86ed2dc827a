#!/bin/bash
timeout 600 python -m pytest tests/test_syncbn_gpu.py tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -8 > gpurun_out/r2j_t.log
for m in reg stream; do SCB_BN_FUSED_MODE=$m timeout 120 python tools/time_bn.py; SCB_BN_FUSED_MODE=$m timeout 120 python tools/time_bn.py 512 512 4 4; SCB_BN_FUSED_MODE=$m timeout 120 python tools/time_bn.py 512 256 8 8;  SCB_BN_FUSED_MODE=$m timeout 120 python tools/time_bn.py 512 128 16 16; done > gpurun_out/r2j_tb.log 2>&1
