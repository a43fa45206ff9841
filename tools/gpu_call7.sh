#!/bin/bash
rm -f gpurun_out/c7.log
for G in 0 1 2 4; do
  export GPB_TICKET_GROUP=$G
  echo "=== ticket group $G" >> gpurun_out/c7.log
  timeout 300 python tools/stress_small.py >> gpurun_out/c7.log 2>&1
  timeout 300 python tools/c3_breakdown.py >> gpurun_out/c7.log 2>&1
done
export GPB_TICKET_GROUP=0
timeout 300 python tools/trace_analyze.py 4096 8 >> gpurun_out/c7.log 2>&1
timeout 300 python tools/trace_analyze.py 1024 256 >> gpurun_out/c7.log 2>&1
export GPB_TICKET_GROUP=2
timeout 300 python tools/trace_analyze.py 4096 8 >> gpurun_out/c7.log 2>&1
