#!/bin/bash
for d in 0 1; do SCB_DBG=$d timeout 120 python tools/time_loss.py 2>&1 | grep SCB_DBG; done > gpurun_out/r2e_tl.log 2>&1
timeout 300 python -m pytest tests/test_supcon_parity_gpu.py tests/test_supcon_gpu.py -m gpu -q 2>&1 | tail -3 > gpurun_out/r2e_t.log
