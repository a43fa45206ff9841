#!/bin/bash
rm -f gpurun_out/c29.log
for L in 1 0; do
GPB_LIGHT_ROWS=$L timeout 300 python tools/light_ab.py 4096 1 >> gpurun_out/c29.log 2>&1
GPB_LIGHT_ROWS=$L timeout 300 python tools/light_ab.py 1024 1 >> gpurun_out/c29.log 2>&1
done
cat gpurun_out/c29.log
