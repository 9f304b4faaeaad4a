#!/bin/bash
# Round-2 evidence run, second edition (after block_ws / sampler changes; one GPU): whole GPU suite, smoke, the full bench line, the reference arm, the ncu launch list of the
# short bench command, ncu --set full of the final whole-block kernel.  Every step has its own timeout; outputs in gpurun_out/.
set +e
O=gpurun_out
mkdir -p $O
step() { local name=$1 lim=$2; shift 2; local t0=$(date +%s); timeout $lim "$@" > $O/$name.log 2> $O/$name.err; local rc=$?
         echo "$name rc=$rc $(( $(date +%s) - t0 ))s" | tee -a $O/steps.txt; }
: > $O/steps.txt
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $O/gpu.txt 2>&1; nproc >> $O/gpu.txt
step z_tests_gpu 800 python -m pytest tests -m gpu -q
step z_smoke 240 python -c "import __graft_entry__ as g; g.smoke()"
step z_bench 700 python bench.py --kernels-out $O/z_kernels.json --encoder-kernels-out $O/z_kernels_encoder.json
step z_ref_arm 300 python bench.py --impl reference --steps 5 --warmup 1
BARGS="--steps 1 --warmup 3 --no-cpu-baseline --no-streaming --no-encoder --no-longform --no-torch-eager"
step z_bench_short 240 python bench.py $BARGS
step z_ncu_launches 500 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/z_launches.csv python bench.py $BARGS
for S in 4 5 6; do
  step z_ncu_bt_s$S 300 ncu --set full --clock-control none --import-source on -k regex:block_ -c 1 -f -o $O/block_r02c_s$S python tools/kbench.py --kind block --stage $S --frames 5760 --iters 1
done
for KS in "block 4" "block 5" "block 6" "ffn 3" "mixer 3"; do set -- $KS; python tools/kbench.py --kind $1 --stage $2 --frames 5760 --iters 20; done > $O/z_kbench.txt 2>&1
python tools/kbench.py --kind ffn --stage 3 --frames 5760 --iters 20 --inplace >> $O/z_kbench.txt 2>&1
python tools/sampler_step.py >> $O/z_kbench.txt 2>&1
python tools/stream_step.py --streams 256 --steps 20 --profile >> $O/z_kbench.txt 2>&1
cat $O/z_kbench.txt $O/steps.txt
