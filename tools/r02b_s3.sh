#!/bin/bash
# session 3 (2 GPUs): the 2-GPU gather test and the bench at N = 2 through torchrun
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_mh.py -m gpu -x -q -k "two_gpu or gather" > gpurun_out/c_pytest_2gpu.log 2>&1; echo "rc=$?" >> gpurun_out/c_pytest_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > gpurun_out/c_bench_n2.json 2> gpurun_out/c_bench_n2.err; echo "rc=$?" >> gpurun_out/c_bench_n2.err
tail -4 gpurun_out/c_pytest_2gpu.log; tail -3 gpurun_out/c_bench_n2.err; head -c 400 gpurun_out/c_bench_n2.json
