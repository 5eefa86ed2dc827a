#!/bin/bash
# one ncu use per call: $1 selects which
case "$1" in
  launches)
    timeout 600 python bench.py --steps 2 --warmup 1 --no-gpu-baseline > gpurun_out/prof_bench_plain.log 2>&1 || exit 1
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_bench.csv \
        python bench.py --steps 2 --warmup 1 --no-gpu-baseline > gpurun_out/prof_bench_ncu.log 2>&1 ;;
  loss)
    timeout 300 python tools/time_loss.py > gpurun_out/prof_loss_plain.log 2>&1 || exit 1
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:supcon_tc -s 6 -c 2 -f -o gpurun_out/r2_loss \
        python tools/time_loss.py > gpurun_out/prof_loss_ncu.log 2>&1 ;;
  bn_small)
    timeout 300 python tools/time_bn.py > gpurun_out/prof_bn_small_plain.log 2>&1 || exit 1
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:bn_ -s 6 -c 2 -f -o gpurun_out/r2_bn_cluster \
        python tools/time_bn.py > gpurun_out/prof_bn_small_ncu.log 2>&1 ;;
  bn_ring)
    timeout 300 python tools/time_bn.py 512 256 16 16 > gpurun_out/prof_bn_ring_plain.log 2>&1 || exit 1
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:bn_ring -s 8 -c 4 -f -o gpurun_out/r2_bn_ring \
        python tools/time_bn.py 512 256 16 16 > gpurun_out/prof_bn_ring_ncu.log 2>&1 ;;
  bn_ring_large)
    timeout 300 python tools/time_bn.py 512 256 32 32 > gpurun_out/prof_bn_ring_large_plain.log 2>&1 || exit 1
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:bn_ring -s 8 -c 4 -f -o gpurun_out/r2_bn_ring_large \
        python tools/time_bn.py 512 256 32 32 > gpurun_out/prof_bn_ring_large_ncu.log 2>&1 ;;
  bn_small_bwd)
    timeout 300 python tools/time_bn.py > gpurun_out/prof_bn_small_plain.log 2>&1 || exit 1
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:bn_bwd_cluster -s 4 -c 2 -f -o gpurun_out/r2_bn_cluster_bwd \
        python tools/time_bn.py > gpurun_out/prof_bn_small_bwd_ncu.log 2>&1 ;;
  head)
    timeout 300 python -m pytest tests/test_head_gpu.py -q -m gpu -k "1024-2048-128-2" > gpurun_out/prof_head_plain.log 2>&1 || exit 1
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:head_tc -c 1 -f -o gpurun_out/r2_head \
        python -m pytest tests/test_head_gpu.py -q -m gpu -k "dtype0-1024-2048-128-2" > gpurun_out/prof_head_ncu.log 2>&1 ;;
esac
