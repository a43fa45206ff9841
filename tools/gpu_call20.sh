#!/bin/bash
rm -f gpurun_out/c20.log
timeout 300 python tools/stress_small.py >> gpurun_out/c20.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c20_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c20_pytest.log
timeout 300 python tools/stress_repro.py >> gpurun_out/c20.log 2>&1
timeout 300 python tools/c3_breakdown.py >> gpurun_out/c20.log 2>&1
echo "=== cw4" >> gpurun_out/c20.log
GPB200_LIB=$PWD/gaussian-processes_b200/libgpb200_cw4.so timeout 300 python tools/c3_breakdown.py >> gpurun_out/c20.log 2>&1
timeout 300 python tools/trace_analyze.py 4096 8 >> gpurun_out/c20.log 2>&1
timeout 300 python tools/trace_analyze.py 1024 256 >> gpurun_out/c20.log 2>&1
