#!/bin/bash
# round 2 baseline probe: tests, schedule traces and ncu captures of the shapes that limit scaling
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/p_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/p_pytest.log
timeout 300 python tools/c3_breakdown.py > gpurun_out/p_breakdown.log 2>&1
for s in "1024 32" "1024 256" "4096 1"; do
  timeout 300 python tools/trace_analyze.py $s > "gpurun_out/p_trace_${s// /x}.log" 2>&1
done
timeout 120 python tools/chol_shape.py 1024 32 > gpurun_out/p_shape_1024x32.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_chol_tasks -s 2 -c 1 -o gpurun_out/p_chol_1024x32 -f python tools/chol_shape.py 1024 32 > gpurun_out/p_ncu_a.log 2>&1
timeout 120 python tools/chol_shape.py 4096 1 > gpurun_out/p_shape_4096x1.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_chol_tasks -s 2 -c 1 -o gpurun_out/p_chol_4096x1 -f python tools/chol_shape.py 4096 1 > gpurun_out/p_ncu_b.log 2>&1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/p_smi.log 2>&1
nproc >> gpurun_out/p_smi.log
