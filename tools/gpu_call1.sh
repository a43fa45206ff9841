#!/bin/bash
# first GPU call of the session: tests, bench, variant comparison, traces
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/c1_smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c1_pytest.log
timeout 600 python bench.py > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err
for v in "" 1x4; do
  if [ -n "$v" ]; then export GPB200_LIB=$PWD/gaussian-processes_b200/libgpb200_$v.so; fi
  echo "=== variant '$v'" >> gpurun_out/c1_variants.log
  timeout 300 python tools/c3_breakdown.py >> gpurun_out/c1_variants.log 2>&1
  timeout 300 python tools/trace_analyze.py 4096 8 >> gpurun_out/c1_variants.log 2>&1
  timeout 300 python tools/trace_analyze.py 4096 1 >> gpurun_out/c1_variants.log 2>&1
  timeout 300 python tools/trace_analyze.py 1024 256 >> gpurun_out/c1_variants.log 2>&1
done
