#!/bin/sh
# Builds tools/bin/libtrace.so: the library with the kernel timeline instrumentation compiled in
# (-DSCB_TRACE=1).  tools/trace_fwd.py and tools/trace_bwd.py load it through SCB_LIB.
set -e
cd "$(dirname "$0")/.."
python -c "import __graft_entry__ as g; g.build()" > /dev/null
mkdir -p tools/bin
cd simclr_pytorch_distributed_b200/csrc
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr \
     -DSCB_TRACE=1 -c supcon_tc.cu -o ../../tools/bin/tc_trace.o
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr \
     -DSCB_TRACE=1 -c bn_cluster.cu -o ../../tools/bin/bn_cluster_trace.o
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr \
     -DSCB_TRACE=1 -c bn_ring.cu -o ../../tools/bin/bn_ring_trace.o
nvcc -shared -o ../../tools/bin/libtrace.so api.o bn.o supcon_generic.o head_tc.o ../../tools/bin/tc_trace.o \
     ../../tools/bin/bn_cluster_trace.o ../../tools/bin/bn_ring_trace.o \
     -gencode arch=compute_100a,code=sm_100a
echo "built tools/bin/libtrace.so"
