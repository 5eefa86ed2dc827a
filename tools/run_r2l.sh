#!/bin/bash
timeout 600 python -m pytest tests/test_supcon_parity_gpu.py -m gpu -q -x 2>&1 | tail -12 > gpurun_out/r2l_t.log
timeout 600 python bench.py --steps 50 --warmup 10 --workload ntxent --no-cpu-baseline --no-gpu-baseline > gpurun_out/r2l_bench.log 2> gpurun_out/r2l_bench.err
