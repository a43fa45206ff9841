#!/bin/bash
# soak: production scheduler vs per-phase scheduler, bit for bit, many seeds x launch modes x chain-pause modes
rm -f gpurun_out/soak.log
for seed in 101 102 103 104; do
  for c in 0 1 2; do
    for p in 0 1; do
      GPB_CHAIN_PAUSE=$p GPB_CTAS_PER_SM_RT=$c SEED=$seed timeout 300 python tools/stress_shapes.py 2>&1 | tail -1 >> gpurun_out/soak.log
    done
  done
done
for p in auto 0 1; do
  for i in 1 2; do
    if [ $p = auto ]; then timeout 300 python tools/stress_repro.py 2>&1 | tail -1 >> gpurun_out/soak.log
    else GPB_CHAIN_PAUSE=$p GPB_CTAS_PER_SM_RT=2 timeout 300 python tools/stress_repro.py 2>&1 | tail -1 >> gpurun_out/soak.log; fi
  done
done
cat gpurun_out/soak.log
