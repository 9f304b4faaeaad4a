#!/bin/bash
# Round-2 GPU call C: the allocation-free PCM16 entry point's test alone in a fresh process, then the whole GPU suite (which
# ends with it), then the block-kernel A/B (kbench, short bench) whose logs were lost with the previous machine.
set +e
O=gpurun_out
mkdir -p $O
step() { local name=$1 lim=$2; shift 2; local t0=$(date +%s); timeout $lim "$@" > $O/$name.log 2> $O/$name.err; local rc=$?
         echo "$name rc=$rc $(( $(date +%s) - t0 ))s" | tee -a $O/steps.txt; }
: > $O/steps.txt
step c_noalloc_alone 240 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x
step c_tests_gpu 700 python -m pytest tests -m gpu -q -x
: > $O/c_kbench.txt
for K in 1 0; do for S in 5 6; do
  echo "== block_kernel=$K stage $S" >> $O/c_kbench.txt
  VV_BLOCK_KERNEL=$K timeout 120 python tools/kbench.py --kind block --stage $S --frames 5760 --iters 10 >> $O/c_kbench.txt 2>&1
done; done
cat $O/c_kbench.txt
BARGS="--steps 3 --warmup 3 --no-cpu-baseline --no-encoder --no-longform --no-torch-eager"
step c_bench_new 240 python bench.py $BARGS --kernels-out $O/c_kernels_new.json
step c_bench_old 240 env VV_BLOCK_KERNEL=0 python bench.py $BARGS --kernels-out $O/c_kernels_old.json
cat $O/steps.txt
