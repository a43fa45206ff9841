#!/bin/bash
rm -f gpurun_out/c21.log
for C in 1 2 0; do
  echo "=== GPB_CTAS_PER_SM_RT=$C" >> gpurun_out/c21.log
  GPB_CTAS_PER_SM_RT=$C timeout 300 python tools/c3_breakdown.py >> gpurun_out/c21.log 2>&1
done
