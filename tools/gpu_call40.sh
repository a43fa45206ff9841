#!/bin/bash
rm -f gpurun_out/c40.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c40_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c40_pytest.log
tail -3 gpurun_out/c40_pytest.log
timeout 300 python tools/stress_repro.py >> gpurun_out/c40.log 2>&1
timeout 300 python tools/c3_breakdown.py >> gpurun_out/c40.log 2>&1
GPB_CTAS_PER_SM_RT=2 timeout 300 python tools/stress_shapes.py 2>&1 | tail -1 >> gpurun_out/c40.log
timeout 300 python tools/stress_shapes.py 2>&1 | tail -1 >> gpurun_out/c40.log
cat gpurun_out/c40.log
