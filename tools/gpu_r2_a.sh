#!/bin/bash
# Round-2 GPU call A: the new parity tests, smoke, the full bench line (with the K2/K3/K4/K5 parity samples), the reference
# arm, and compute-sanitizer memcheck / racecheck of the smoke-sized paths.  Every step has its own timeout.
set +e
O=gpurun_out
mkdir -p $O
step() { local name=$1 lim=$2; shift 2; local t0=$(date +%s); timeout $lim "$@" > $O/$name.log 2> $O/$name.err; local rc=$?
         echo "$name rc=$rc $(( $(date +%s) - t0 ))s" | tee -a $O/steps.txt; }
: > $O/steps.txt
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $O/gpu.txt 2>&1
nproc >> $O/gpu.txt
step a_tests_gpu 600 python -m pytest tests -m gpu -q -x
step a_smoke 240 python -c "import __graft_entry__ as g; g.smoke()"
step a_bench 600 python bench.py --kernels-out $O/a_kernels.json --encoder-kernels-out $O/a_kernels_encoder.json
step a_ref_arm 300 python bench.py --impl reference --steps 5 --warmup 1
step a_memcheck 500 compute-sanitizer --tool memcheck --print-limit 20 python tools/sanitize_paths.py
step a_racecheck 500 compute-sanitizer --tool racecheck --print-limit 20 python tools/sanitize_paths.py --small
cat $O/steps.txt
