#!/bin/bash
# ncu evidence for round 2 (each capture only after the same command has exited 0 without ncu)
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
timeout 300 $B > gpurun_out/e_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/e_launches.csv $B > gpurun_out/e_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_chol_tasks -s 3 -c 1 -o gpurun_out/e_chol_c2 -f $B > gpurun_out/e_ncu2.log 2>&1
timeout 120 python tools/chol_shape.py 1024 256 > gpurun_out/e_shape_c3.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_chol_tasks -s 2 -c 1 -o gpurun_out/e_chol_c3 -f python tools/chol_shape.py 1024 256 > gpurun_out/e_ncu3.log 2>&1
timeout 120 python tools/chol_shape.py 1024 32 > gpurun_out/e_shape_c3_32.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_chol_tasks -s 2 -c 1 -o gpurun_out/e_chol_c3_32 -f python tools/chol_shape.py 1024 32 > gpurun_out/e_ncu4.log 2>&1
timeout 120 python tools/chol_shape.py 4096 2 > gpurun_out/e_shape_4096x2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_chol_tasks -s 2 -c 1 -o gpurun_out/e_chol_4096x2 -f python tools/chol_shape.py 4096 2 > gpurun_out/e_ncu5.log 2>&1
timeout 120 python tools/mh_shape.py 1024 256 12 truth > gpurun_out/e_mh.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/e_mh_launches.csv python tools/mh_shape.py 1024 256 12 truth > gpurun_out/e_ncu6.log 2>&1
