#!/bin/bash
rm -f gpurun_out/c36.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c36_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c36_pytest.log
tail -3 gpurun_out/c36_pytest.log
timeout 300 python tools/stress_repro.py >> gpurun_out/c36.log 2>&1
timeout 300 python tools/c3_breakdown.py >> gpurun_out/c36.log 2>&1
cat gpurun_out/c36.log
