#!/bin/bash
# Round-2 GPU call B: first run of block_tma_kernel -- parity of both whole-block kernels, kbench A/B of library variants
# (chunk width, ring depth), a short bench, then the whole GPU suite.
set +e
O=gpurun_out
mkdir -p $O
step() { local name=$1 lim=$2; shift 2; local t0=$(date +%s); timeout $lim "$@" > $O/$name.log 2> $O/$name.err; local rc=$?
         echo "$name rc=$rc $(( $(date +%s) - t0 ))s" | tee -a $O/steps.txt; }
: > $O/steps.txt
step b_tests_block 400 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_decode.py -m gpu -q -x -k "block or stream or offline_stress or state"
LIBDIR=$PWD/vibevoice-rocm_b200/vvdec
: > $O/b_kbench.txt
for L in libvvdec_b200 $(cd $LIBDIR; ls libvvdec_b200_*.so 2>/dev/null | sed 's/\.so$//'); do
  for K in 1 0; do
    [ "$K" = 0 ] && [ "$L" != libvvdec_b200 ] && continue
    for S in 5 6; do
      echo "== $L block_kernel=$K stage $S" >> $O/b_kbench.txt
      VVDEC_LIB=$LIBDIR/$L.so VV_BLOCK_KERNEL=$K timeout 120 python tools/kbench.py --kind block --stage $S --frames 5760 --iters 10 >> $O/b_kbench.txt 2>&1
    done
  done
done
cat $O/b_kbench.txt
BARGS="--steps 3 --warmup 3 --no-cpu-baseline --no-encoder --no-longform --no-torch-eager"
step b_bench_new 240 python bench.py $BARGS --kernels-out $O/b_kernels_new.json
step b_bench_old 240 env VV_BLOCK_KERNEL=0 python bench.py $BARGS --kernels-out $O/b_kernels_old.json
step b_tests_gpu 600 python -m pytest tests -m gpu -q
cat $O/steps.txt
