#!/bin/bash
# one GPU: the whole -m gpu suite (one pytest process, as the driver runs it; REPEAT times), smoke, default bench
for i in $(seq 1 ${REPEAT:-1}); do
  timeout 1500 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -6 > gpurun_out/all_tests_$i.log
done
timeout 300 python __graft_entry__.py smoke > gpurun_out/all_smoke.log 2>&1
timeout 900 python bench.py --steps 50 --warmup 10 $BENCH_FLAGS > gpurun_out/all_bench.log 2> gpurun_out/all_bench.err
