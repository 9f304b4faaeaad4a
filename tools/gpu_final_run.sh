#!/bin/bash
# Final verification of HEAD on the GPU box: every GPU test, smoke(), the bench line, the ncu launch list of the
# same command (after it ran alone) and the reference arm.  Outputs under gpurun_out/f_*.
set +e
O=gpurun_out; mkdir -p $O
step() { local name=$1 lim=$2; shift 2; local t0=$(date +%s); timeout $lim "$@" > $O/$name.log 2> $O/$name.err; echo "$name rc=$? $(( $(date +%s) - t0 ))s" | tee -a $O/steps2.txt; }
: > $O/steps2.txt
step f_tests_gpu 420 python -m pytest tests -m gpu -q
step f_smoke 200 python -c "import __graft_entry__ as g; g.smoke()"
step f_bench 400 python bench.py --kernels-out $O/f_kernels.json --encoder-kernels-out $O/f_kernels_encoder.json
BARGS="--steps 1 --warmup 3 --no-cpu-baseline --no-streaming --no-encoder --no-longform --no-torch-eager"
step f_bench_short 200 python bench.py $BARGS
step f_ncu_launches 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/f_launches.csv python bench.py $BARGS
step f_ref_arm 200 python bench.py --impl reference --steps 2 --warmup 1
