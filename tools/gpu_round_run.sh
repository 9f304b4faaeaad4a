#!/bin/bash
# One gpurun call that produces the round's evidence (run from the repo root on the GPU box):
#   tests, smoke, the bench line, the ncu launch list of the same command, then the VV_PDL=1 A/B and two ncu captures.
# Every step has its own timeout and writes under gpurun_out/; a failing step does not stop the rest.
set +e
O=gpurun_out
mkdir -p $O
step() { local name=$1 lim=$2; shift 2; local t0=$(date +%s); timeout $lim "$@" > $O/$name.log 2> $O/$name.err; local rc=$?
         echo "$name rc=$rc $(( $(date +%s) - t0 ))s" | tee -a $O/steps.txt; }
: > $O/steps.txt
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $O/gpu.txt 2>&1
step tests_gpu 420 python -m pytest tests -m gpu -q
step smoke 240 python -c "import __graft_entry__ as g; g.smoke()"
step bench 540 python bench.py --kernels-out $O/kernels.json --encoder-kernels-out $O/kernels_encoder.json
BARGS="--steps 1 --warmup 3 --no-cpu-baseline --no-streaming --no-encoder --no-longform --no-torch-eager"
step bench_short 240 python bench.py $BARGS
step ncu_launches 420 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches.csv python bench.py $BARGS
step bench_pdl 300 env VV_PDL=1 python bench.py --steps 3 --no-cpu-baseline --no-encoder --no-longform --no-torch-eager --kernels-out $O/kernels_pdl.json
step ncu_up6 240 ncu --set full --clock-control none --import-source on -k regex:gemm_tc2 -c 1 -f -o $O/gemm_tc2_up6 python tools/kbench.py --kind upsample --stage 6 --frames 5760 --iters 1
step ncu_pcm16 240 ncu --set full --clock-control none --import-source on -k regex:pcm16 -c 3 -f -o $O/pcm16 python -c "
import sys; sys.path.insert(0, 'vibevoice-rocm_b200'); import torch, vvdec
x = (torch.randn(64, 1, 1440000, device='cuda') * 3).to(torch.bfloat16)
vvdec.pcm16(x); vvdec.pcm16(x[:, :, :3200].contiguous()); torch.cuda.synchronize()"
step tests_gpu_pdl 420 env VV_PDL=1 python -m pytest tests -m gpu -q
cat $O/steps.txt
