#!/bin/bash
# A/B of the block kernel (HEAD~ build vs this build), then the whole GPU suite and a short bench with this build.
set +e
O=gpurun_out; mkdir -p $O
for L in libvvdec_b200_head libvvdec_b200; do for S in 5 6; do echo "== $L stage $S"; VVDEC_LIB=$PWD/vibevoice-rocm_b200/vvdec/$L.so python tools/kbench.py --kind block --stage $S --frames 5760 --iters 5 2>&1 | grep -v Warn; done; done | tee $O/g_kbench.txt
timeout 300 python -m pytest tests -m gpu -q > $O/g_tests_gpu.log 2>&1; echo "tests rc=$?"; tail -2 $O/g_tests_gpu.log
timeout 100 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-streaming --no-encoder --no-longform --no-torch-eager --kernels-out $O/g_kernels.json > $O/g_bench_short.log 2>&1; echo "bench rc=$?"
