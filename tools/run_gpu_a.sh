#!/bin/bash
timeout 900 python -m pytest tests/test_syncbn_gpu.py tests/test_zz_no_alloc.py tests/test_model_gpu.py -x -q -m gpu 2>&1 | tail -4 > gpurun_out/r2y_t.log
SCB_BN_FUSED_MODE=coop timeout 900 python -m pytest tests/test_syncbn_gpu.py -x -q -m gpu 2>&1 | tail -4 >> gpurun_out/r2y_t.log
timeout 900 python bench.py --workload syncbn-census --steps 20 --warmup 3 > gpurun_out/r2y_census.log 2> gpurun_out/r2y_census.err
