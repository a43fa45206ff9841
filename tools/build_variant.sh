#!/bin/bash
# build_variant.sh NAME [-D...]: compile a tuning variant of libgpb200 into gaussian-processes_b200/libgpb200_NAME.so
set -e
cd "$(dirname "$0")/../gaussian-processes_b200/csrc"
name=$1; shift
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@" \
  -shared -o ../libgpb200_${name}.so gpb_build.cu gpb_chol.cu gpb_epilogue.cu gpb_api.cu gpb_mh.cu gpb_trsm.cu gpb_comm.cu gpb_next.cu gpb_grad.cu gpb_noalloc.cu -lcudart -ldl -Xptxas -v 2> /tmp/variant_${name}.log
grep -A2 "k_chol_tasks" /tmp/variant_${name}.log | grep -E "spill|Used" | head -3
