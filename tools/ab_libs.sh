#!/bin/bash
# A/B of library builds on one box, back to back: K2 headline (3 steps), K4 streaming and the per-kernel table for each
# lib listed below.  Variant libs are built next to the product one with extra -D flags, e.g.
#   nvcc <flags of __graft_entry__.NVCC_FLAGS> -DVV_TRYWAIT_HINT_NS=0 -o vibevoice-rocm_b200/vvdec/libvvdec_b200_hint0.so vibevoice-rocm_b200/csrc/vvdec.cu
# and selected through VVDEC_LIB (vvdec/_capi.py).  Usage: gpurun -- 'bash tools/ab_libs.sh'  (edit the list).
mkdir -p gpurun_out
BARGS="--steps 3 --warmup 3 --no-cpu-baseline --no-encoder --no-longform --no-torch-eager"
for L in libvvdec_b200 libvvdec_b200_hint2000 libvvdec_b200_hint10000000; do
  export VVDEC_LIB=$PWD/vibevoice-rocm_b200/vvdec/$L.so
  echo "== $L"
  python bench.py $BARGS --kernels-out gpurun_out/kernels_$L.json > gpurun_out/ab_$L.log 2> gpurun_out/ab_$L.err
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/ab_$L.log') if l.startswith('{')][-1])
s=d['streaming']
print('K2 ms/step',round(d['ms_per_step'],2),'clk',d['clocks']['sm_mhz'],'stream eager',round(s['eager_p50_ms'],3),'graph',round(s.get('graph_p50_ms',0),3),'sem',round(s['semantic_encode']['eager_p50_ms'],3))
k=json.load(open('gpurun_out/kernels_$L.json'))['kernels']
print('  '+' '.join(f"{x['kind']}@{x['stage']}={x['ms_per_launch']:.3f}" for x in sorted(k,key=lambda x:-x['ms_total'])[:12]))
PY
done
