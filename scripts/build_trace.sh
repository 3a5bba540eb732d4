#!/bin/bash
# developer aid: libdualhyp_trace.so = the library with -DDH_ATTN_TRACE (per-event clock64 dumps from one CTA)
set -e
cd "$(dirname "$0")/../dualhyp_b200/csrc"
make -s > /dev/null
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr -DDH_ATTN_TRACE -c attention.cu -o build/attention_trace.o
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libdualhyp_trace.so build/runtime.o build/leaf.o build/gemm.o build/attention_trace.o build/relpred.o -lcudart_static -ldl -lrt -lpthread
