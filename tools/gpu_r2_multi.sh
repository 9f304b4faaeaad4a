#!/bin/bash
# Round-2 multi-GPU refresh (final edition): `gpurun --gpus N -- bash tools/gpu_r2_multi.sh N`.  2-rank NCCL tests of the
# product API (N >= 2) and the bench under torch.distributed.run (K2 weak scaling + K3 through decode_sharded).
set +e
N=${1:-2}
O=gpurun_out
mkdir -p $O
step() { local name=$1 lim=$2; shift 2; local t0=$(date +%s); timeout $lim "$@" > $O/$name.log 2> $O/$name.err; local rc=$?
         echo "$name rc=$rc $(( $(date +%s) - t0 ))s" | tee -a $O/steps.txt; }
: > $O/steps.txt
nvidia-smi -L > $O/m_gpus_$N.txt
if [ "$N" = "2" ]; then step m_multi_tests 400 python -m pytest tests/test_gpu_multi.py -m gpu -q -x; fi
step m_bench_$N 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 4 --warmup 3 --no-encoder --no-torch-eager --no-streaming
cat $O/steps.txt
