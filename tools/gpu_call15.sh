#!/bin/bash
rm -f gpurun_out/c15.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c15_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c15_pytest.log
timeout 300 python tools/stress_repro.py >> gpurun_out/c15.log 2>&1
timeout 900 python bench.py > gpurun_out/c15_bench.json 2> gpurun_out/c15_bench.err; echo "bench rc=$?" >> gpurun_out/c15_bench.err
