#!/bin/bash
timeout 600 python bench.py --steps 5 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/c25_short.json 2> gpurun_out/c25_short.err
timeout 900 python bench.py > gpurun_out/c25_bench.json 2> gpurun_out/c25_bench.err; echo "rc=$?" >> gpurun_out/c25_bench.err
timeout 300 python tools/c3_breakdown.py > gpurun_out/c25_breakdown.log 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/c25_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c25_pytest.log
