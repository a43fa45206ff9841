#!/bin/bash
# session 1 (round 2b): no-alloc test alone + whole GPU suite + bench + ticket-group sweep at the C3 shapes
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/a_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/a_noalloc_alone.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/a_pytest.log
timeout 900 python bench.py > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err; echo "bench rc=$?" >> gpurun_out/a_bench.err
for tg in 0 4 8 16 32 64 128; do
  for shape in "1024 256" "1024 32" "1024 64"; do
    echo "== GPB_TICKET_GROUP=$tg shape $shape" >> gpurun_out/a_tgroup.log
    GPB_TICKET_GROUP=$tg timeout 120 python tools/chol_shape.py $shape 6 >> gpurun_out/a_tgroup.log 2>&1
  done
done
for tg in 0 8 16 32; do
  GPB_TICKET_GROUP=$tg timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active --clock-control none -k regex:k_chol_tasks -s 2 -c 1 --csv --log-file gpurun_out/a_tg${tg}_c3.csv python tools/chol_shape.py 1024 256 > gpurun_out/a_tg_ncu.log 2>&1
done
tail -3 gpurun_out/a_noalloc_alone.log gpurun_out/a_pytest.log; cat gpurun_out/a_tgroup.log
