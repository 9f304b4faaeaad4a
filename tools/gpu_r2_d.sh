#!/bin/bash
# Round-2 GPU call D: block_tma second build (constant-bank biases, accumulator preloaded with y, shared-space addressing in
# every kernel): parity first, then kbench of the kernel classes, short bench, whole suite.
set +e
O=gpurun_out
mkdir -p $O
step() { local name=$1 lim=$2; shift 2; local t0=$(date +%s); timeout $lim "$@" > $O/$name.log 2> $O/$name.err; local rc=$?
         echo "$name rc=$rc $(( $(date +%s) - t0 ))s" | tee -a $O/steps.txt; }
: > $O/steps.txt
step d_tests_block 400 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_decode.py -m gpu -q -x -k "block or stream or offline_stress or state"
: > $O/d_kbench.txt
for K in 1 0; do for S in 5 6; do
  echo "== block_kernel=$K stage $S" >> $O/d_kbench.txt
  VV_BLOCK_KERNEL=$K timeout 120 python tools/kbench.py --kind block --stage $S --frames 5760 --iters 10 >> $O/d_kbench.txt 2>&1
done; done
for KS in "ffn 4" "mixer 4" "ffn 3" "ffn 2" "ffn 0" "upsample 6" "upsample 4" "mixer 3"; do
  set -- $KS; timeout 120 python tools/kbench.py --kind $1 --stage $2 --frames 5760 --iters 10 >> $O/d_kbench.txt 2>&1
done
cat $O/d_kbench.txt
BARGS="--steps 3 --warmup 3 --no-cpu-baseline --no-encoder --no-longform --no-torch-eager"
step d_bench 240 python bench.py $BARGS --kernels-out $O/d_kernels.json
step d_tests_gpu 700 python -m pytest tests -m gpu -q -x
cat $O/steps.txt
