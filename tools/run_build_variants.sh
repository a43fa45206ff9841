#!/bin/bash
# time kernel build variants (libgpb200_<name>.so next to the product library); scratch
for v in "" "$@"; do
  echo "== variant '$v'"
  GPB200_LIB=$PWD/gaussian-processes_b200/libgpb200$v.so python tools/c3_breakdown.py 2>&1 | grep -E "n=1024 B=256|n=1024 B=32|n=4096 B=1:|n=4096 B=8|n=8192 B=8"
done
