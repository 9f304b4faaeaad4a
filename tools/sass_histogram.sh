#!/bin/bash
# SASS opcode histogram of the shipped library: proof that the contraction / tile-movement kernels are tcgen05 / TMEM / TMA
# (mnemonics per /opt/skills/guides/B200_PROFILING.md).  Usage: bash tools/sass_histogram.sh > profiles/sass_histogram_rNN.txt
LIB=${1:-vibevoice-rocm_b200/vvdec/libvvdec_b200.so}
echo "# $(basename $LIB)  $(sha256sum $LIB | cut -c1-16)  $(date -u +%FT%TZ)"
cuobjdump -sass $LIB > /tmp/vv_sass.txt
echo "## whole library"
for op in UTCHMMA UTCBAR UTMALDG UTMASTG UTMACMDFLUSH LDTM STTM UTCATOMSWS SYNCS FFMA2 FMUL2 FADD2 MUFU.TANH MUFU.RSQ HMMA IMMA; do
  printf "%-14s %6d\n" $op $(grep -cE "(^|[^A-Z.])$op" /tmp/vv_sass.txt)
done
echo "## per kernel (tcgen05.mma = UTCHMMA, TMA load/store = UTMALDG/UTMASTG, TMEM ld/st = LDTM/STTM)"
awk '/Function :/ {name=$3} /UTCHMMA/ {m[name]++} /UTMALDG/ {l[name]++} /UTMASTG/ {s[name]++} /LDTM/ {t[name]++} /STTM/ {u[name]++}
     END {for (n in m) printf "%s  UTCHMMA %3d UTMALDG %3d UTMASTG %3d LDTM %3d STTM %3d\n", n, m[n], l[n], s[n], t[n], u[n]}' /tmp/vv_sass.txt | sort | c++filt
