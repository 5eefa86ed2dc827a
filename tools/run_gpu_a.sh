#!/bin/bash
timeout 900 python -m pytest tests/test_syncbn_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3 > gpurun_out/r3p_t.log
for s in "256 256 8 8" "512 256 8 8" "512 512 8 8" "512 128 16 16" "512 2048 4 4" "512 256 16 16" "512 1024 8 8" "512 256 32 32"; do
  timeout 200 python tools/time_bn.py $s 2>&1 | grep shape
done > gpurun_out/r3p_tb.log 2>&1
