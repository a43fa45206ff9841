#!/bin/bash
rm -f gpurun_out/c27.log
timeout 300 python tools/stress_small.py >> gpurun_out/c27.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c27_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c27_pytest.log
timeout 300 python tools/stress_repro.py >> gpurun_out/c27.log 2>&1
timeout 300 python tools/c3_breakdown.py >> gpurun_out/c27.log 2>&1
timeout 300 python tools/trace_analyze.py 4096 1 >> gpurun_out/c27.log 2>&1
