#!/bin/bash
# session 6: is the long MH loop slower per iteration than a short one because of clocks?  (nvidia-smi samples during the runs)
mkdir -p gpurun_out
nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown,temperature.gpu --format=csv,noheader -lms 100 > gpurun_out/g_smi.csv &
SMI=$!
sleep 1
for it in 12 50 200 800; do
  echo "MARK start iters=$it $(date +%s.%N)" >> gpurun_out/g_mh.log
  python tools/mh_shape.py 1024 256 $it truth >> gpurun_out/g_mh.log 2>&1
  echo "MARK end $(date +%s.%N)" >> gpurun_out/g_mh.log
done
kill $SMI
cat gpurun_out/g_mh.log; sort gpurun_out/g_smi.csv | uniq -c | sort -rn | head -20
