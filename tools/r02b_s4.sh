#!/bin/bash
# session 4 (8 GPUs): the bench through torchrun at N = 8 (library NCCL communicator, gather of ring segments)
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 > gpurun_out/e_bench_n8.json 2> gpurun_out/e_bench_n8.err; echo "rc=$?" >> gpurun_out/e_bench_n8.err
tail -3 gpurun_out/e_bench_n8.err; head -c 300 gpurun_out/e_bench_n8.json
