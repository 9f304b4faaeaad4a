#!/bin/bash
# Round-2 GPU call E (2 GPUs): product multi-GPU API over NCCL, ffn_stream constant-bank change, K3 strong scaling.
set +e
O=gpurun_out
mkdir -p $O
step() { local name=$1 lim=$2; shift 2; local t0=$(date +%s); timeout $lim "$@" > $O/$name.log 2> $O/$name.err; local rc=$?
         echo "$name rc=$rc $(( $(date +%s) - t0 ))s" | tee -a $O/steps.txt; }
: > $O/steps.txt
nvidia-smi -L > $O/e_gpus.txt
step e_multi 400 python -m pytest tests/test_gpu_multi.py -m gpu -q -x
step e_ffn 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_decode.py tests/test_gpu_encode.py -m gpu -q -x -k "ffn_half or offline or encode_stress"
python tools/kbench.py --kind ffn --stage 4 --frames 5760 --iters 10 > $O/e_kbench.txt 2>&1; cat $O/e_kbench.txt
step e_bench2 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 4 --warmup 3 --no-encoder --no-torch-eager --no-streaming
cat $O/steps.txt
