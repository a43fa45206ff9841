#!/bin/bash
mkdir -p gpurun_out
for v in "" 1x4; do
  if [ -n "$v" ]; then export GPB200_LIB=$PWD/gaussian-processes_b200/libgpb200_$v.so; fi
  echo "=== variant '$v'" >> gpurun_out/c2_stress.log
  timeout 600 python tools/stress_repro.py >> gpurun_out/c2_stress.log 2>&1
  timeout 600 python -m pytest tests -m gpu -q -k "full_size" >> gpurun_out/c2_stress.log 2>&1
done
