#!/bin/bash
# tests + bench + ncu evidence for the current default build; outputs under gpurun_out/v_*
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/v_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/v_pytest.log
timeout 900 python bench.py > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; echo "bench rc=$?" >> gpurun_out/v_bench.err
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
timeout 300 $B > gpurun_out/v_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/v_launches.csv $B > gpurun_out/v_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_chol_tasks -s 3 -c 1 -o gpurun_out/v_chol -f $B > gpurun_out/v_ncu2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_build -s 3 -c 1 -o gpurun_out/v_build -f $B > gpurun_out/v_ncu3.log 2>&1
timeout 300 python tools/grad_timing.py > gpurun_out/v_grad_timing.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_grad_contract -s 2 -c 1 -o gpurun_out/v_grad -f python tools/grad_timing.py 4096 8 > gpurun_out/v_ncu4.log 2>&1
