#!/bin/bash
rm -f gpurun_out/c5_stress.log
for v in 1x4 2x2 2x4k8; do
  export GPB200_LIB=$PWD/gaussian-processes_b200/libgpb200_$v.so
  echo "=== variant '$v'" >> gpurun_out/c5_stress.log
  timeout 300 python tools/stress_small.py >> gpurun_out/c5_stress.log 2>&1
  timeout 300 python tools/c3_breakdown.py >> gpurun_out/c5_stress.log 2>&1
done
