#!/bin/bash
# N-GPU bench through the driver's launch line
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 3 > gpurun_out/multi_${N}.json 2> gpurun_out/multi_${N}.err
echo "rc=$?" >> gpurun_out/multi_${N}.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 1 --warmup 0 > gpurun_out/multi_ref_${N}.json 2>> gpurun_out/multi_${N}.err
