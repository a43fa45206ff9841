#!/bin/bash
rm -f gpurun_out/c10.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c10_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c10_pytest.log
timeout 300 python tools/stress_repro.py >> gpurun_out/c10.log 2>&1
timeout 300 python tools/c3_breakdown.py >> gpurun_out/c10.log 2>&1
for G in 0 8 16 32; do GPB_TICKET_GROUP=$G timeout 300 python tools/group_sweep.py 8192 64 >> gpurun_out/c10.log 2>&1; done
for G in 0 64 128; do GPB_TICKET_GROUP=$G timeout 300 python tools/group_sweep.py 1024 256 >> gpurun_out/c10.log 2>&1; done
for G in 0 8 16; do GPB_TICKET_GROUP=$G timeout 300 python tools/group_sweep.py 4096 32 >> gpurun_out/c10.log 2>&1; done
timeout 300 python tools/trace_analyze.py 1024 256 >> gpurun_out/c10.log 2>&1
timeout 300 python tools/trace_analyze.py 4096 8 >> gpurun_out/c10.log 2>&1
