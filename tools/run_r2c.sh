#!/bin/bash
# GPU session script (scratch): ncu of the fused cluster BN kernels + launch list
timeout 120 python tools/time_bn.py > gpurun_out/r2c_plain.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on -k regex:bn_.*cluster -s 6 -c 2 -o gpurun_out/r2c_bn_fused python tools/time_bn.py > gpurun_out/r2c_ncu.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 60 --csv --log-file gpurun_out/r2c_launches.csv python tools/time_bn.py > gpurun_out/r2c_ncu2.log 2>&1
