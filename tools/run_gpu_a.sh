#!/bin/bash
timeout 900 python bench.py --workload syncbn-census --steps 20 --warmup 3 > gpurun_out/r2r_census.log 2> gpurun_out/r2r_census.err
timeout 900 python -m pytest tests/test_syncbn_gpu.py tests/test_model_gpu.py tests/test_peer_gpu.py -x -q -m gpu 2>&1 | tail -3 > gpurun_out/r2r_t.log
