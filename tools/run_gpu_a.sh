#!/bin/bash
timeout 900 python -m pytest tests/test_model_gpu.py tests/test_head_gpu.py -x -q -m gpu 2>&1 | tail -8 > gpurun_out/r2u_t.log
timeout 900 python bench.py --workload model --steps 20 --warmup 5 > gpurun_out/r2u_model.log 2> gpurun_out/r2u_model.err
timeout 300 python __graft_entry__.py smoke > gpurun_out/r2u_smoke.log 2>&1
