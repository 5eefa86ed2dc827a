#!/bin/bash
timeout 900 python -m pytest tests/test_syncbn_gpu.py tests/test_zz_no_alloc.py tests/test_model_gpu.py tests/test_peer_gpu.py -x -q -m gpu 2>&1 | tail -6 > gpurun_out/r2z_t.log
timeout 200 python tools/time_bn.py 2>&1 | grep shape > gpurun_out/r2z_tb.log
timeout 200 python tools/time_bn.py 512 256 8 8 2>&1 | grep shape >> gpurun_out/r2z_tb.log
timeout 600 python bench.py --steps 50 --warmup 10 > gpurun_out/r2z_bench.log 2> gpurun_out/r2z_bench.err
