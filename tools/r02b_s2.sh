#!/bin/bash
# session 2: asynchronous chain groups (two_group.py) and the launch list of the current MH loop
mkdir -p gpurun_out
bash tools/r02_two_group.sh > gpurun_out/b_two_group.log 2>&1
for shape in "1024 256 12 prior" "1024 32 12 prior"; do
  tag=$(echo $shape | tr ' ' '_')
  timeout 120 python tools/mh_shape.py $shape > gpurun_out/b_mh_$tag.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/b_mh_launches_$tag.csv python tools/mh_shape.py $shape > gpurun_out/b_ncu_$tag.log 2>&1
done
timeout 120 python tools/mh_shape.py 1024 256 200 prior >> gpurun_out/b_mh_long.log 2>&1
timeout 120 python tools/mh_shape.py 1024 256 200 truth >> gpurun_out/b_mh_long.log 2>&1
timeout 120 python tools/mh_shape.py 1024 32 400 prior >> gpurun_out/b_mh_long.log 2>&1
timeout 120 python tools/mh_shape.py 1024 32 400 truth >> gpurun_out/b_mh_long.log 2>&1
cat gpurun_out/b_two_group.log gpurun_out/b_mh_long.log
