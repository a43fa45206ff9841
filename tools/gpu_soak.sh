#!/bin/bash
# soak: production scheduler vs per-phase scheduler, bit for bit, many seeds x launch modes
rm -f gpurun_out/soak.log
for seed in 101 102 103 104 105 106; do
  for c in 0 1 2; do
    GPB_CTAS_PER_SM_RT=$c SEED=$seed timeout 300 python tools/stress_shapes.py 2>&1 | tail -1 >> gpurun_out/soak.log
  done
done
for i in 1 2 3; do timeout 300 python tools/stress_repro.py 2>&1 | tail -1 >> gpurun_out/soak.log; done
cat gpurun_out/soak.log
