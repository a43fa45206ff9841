#!/bin/bash
rm -f gpurun_out/c19.log
for V in "" nv; do
for D in 0 3; do
  if [ -n "$V" ]; then export GPB200_LIB=$PWD/gaussian-processes_b200/libgpb200_$V.so; fi
  echo "=== variant '$V' GPB_DBG_NOTRAFFIC=$D" >> gpurun_out/c19.log
  GPB_DBG_NOTRAFFIC=$D timeout 300 python tools/c3_breakdown.py 2>&1 | grep -E "n=4096 B=8|n=8192|n=1024 B=256|n=4096 B=1:" >> gpurun_out/c19.log
done; done
