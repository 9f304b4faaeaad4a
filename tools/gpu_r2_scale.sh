#!/bin/bash
# K3 strong scaling through the product API at N GPUs (and the K2 weak-scaling headline beside it): bench.py under torchrun.
N=${1:-8}
O=gpurun_out
mkdir -p $O
nvidia-smi -L | wc -l > $O/scale_${N}_ngpus.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus $N --steps 4 --warmup 3 --no-encoder --no-torch-eager --no-streaming > $O/scale_${N}.log 2> $O/scale_${N}.err
echo "scale $N rc=$?"
tail -c 600 $O/scale_${N}.err
