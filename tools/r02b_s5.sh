#!/bin/bash
# session 5: bitwise soak of the final kernels (reduced: 2 seeds) + ncu evidence of the final build
mkdir -p gpurun_out
rm -f gpurun_out/soak.log
for seed in 201 202; do
  for c in 0 1 2; do
    for p in 0 1; do
      GPB_CHAIN_PAUSE=$p GPB_CTAS_PER_SM_RT=$c SEED=$seed timeout 300 python tools/stress_shapes.py 2>&1 | tail -1 >> gpurun_out/soak.log
    done
  done
done
for p in auto 0 1; do
  if [ $p = auto ]; then timeout 300 python tools/stress_repro.py 2>&1 | tail -1 >> gpurun_out/soak.log
  else GPB_CHAIN_PAUSE=$p GPB_CTAS_PER_SM_RT=2 timeout 300 python tools/stress_repro.py 2>&1 | tail -1 >> gpurun_out/soak.log; fi
done
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
timeout 300 $B > gpurun_out/f_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/f_launches.csv $B > gpurun_out/f_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_chol_tasks -s 3 -c 1 -o gpurun_out/f_chol_c2 -f $B > gpurun_out/f_ncu2.log 2>&1
timeout 120 python tools/chol_shape.py 4096 1 > gpurun_out/f_shape_4096x1.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_chol_tasks -s 2 -c 1 -o gpurun_out/f_chol_4096x1 -f python tools/chol_shape.py 4096 1 > gpurun_out/f_ncu3.log 2>&1
cat gpurun_out/soak.log
