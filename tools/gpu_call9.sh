#!/bin/bash
timeout 900 python -m pytest tests -m gpu -x -q -k "c4" > gpurun_out/c9_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c9_pytest.log
timeout 1200 python bench.py > gpurun_out/c9_bench.json 2> gpurun_out/c9_bench.err; echo "bench rc=$?" >> gpurun_out/c9_bench.err
