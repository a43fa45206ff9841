#!/bin/bash
rm -f gpurun_out/c18.log
for D in 0 1 2; do
  echo "=== GPB_DBG_NOTRAFFIC=$D" >> gpurun_out/c18.log
  GPB_DBG_NOTRAFFIC=$D timeout 300 python tools/c3_breakdown.py >> gpurun_out/c18.log 2>&1
done
