#!/bin/bash
# session 7: ncu --set full of the final kernels at the C3 shapes (each capture after the same command exited 0 without ncu)
mkdir -p gpurun_out
for shape in "1024 256" "1024 32"; do
  tag=$(echo $shape | tr ' ' 'x')
  timeout 120 python tools/chol_shape.py $shape > gpurun_out/l_shape_$tag.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_chol_tasks -s 2 -c 1 -o gpurun_out/l_chol_$tag -f python tools/chol_shape.py $shape > gpurun_out/l_ncu_$tag.log 2>&1
done
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
timeout 300 $B > gpurun_out/l_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/l_launches.csv $B > gpurun_out/l_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_chol_tasks -s 3 -c 1 -o gpurun_out/l_chol_c2 -f $B > gpurun_out/l_ncu2.log 2>&1
timeout 120 python tools/mh_shape.py 1024 256 12 truth > gpurun_out/l_mh.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/l_mh_launches.csv python tools/mh_shape.py 1024 256 12 truth > gpurun_out/l_ncu6.log 2>&1
cat gpurun_out/l_shape_*.log
