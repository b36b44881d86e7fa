#!/bin/bash
# Per-kernel counts of the Blackwell-native SASS opcodes in the built library (the .so is git-ignored, so the evidence
# is committed as text):  tools/sass_opcodes.sh > profiles/sass_opcodes.txt
set -e
LIB=${1:-$(dirname "$0")/../mocov2_whisper_flamingo_b200/libgxa_b200.so}
echo "# cuobjdump -sass $(basename "$LIB") -- SASS opcode counts per kernel (tcgen05.mma = UTC*MMA, tcgen05.ld/st = LDTM/STTM,"
echo "# TMA = UTMALDG / UTMASTG / UTMAREDG / UBLKCP / UBLKRED, legacy mma.sync = HMMA)"
printf "%-110s %8s %6s %6s %6s %8s %8s %8s %6s\n" kernel UTCHMMA .2CTA LDTM STTM UTMALDG UTMAREDG UBLK HMMA
cuobjdump -sass "$LIB" | awk '
  /Function :/ { if (name != "") emit(); name=$3; a=b=c=d=e=f=g=h=0; next }
  /UTCHMMA/ { a++; if ($0 ~ /2CTA/) b++ }
  /LDTM/ { c++ }
  /STTM/ { d++ }
  /UTMALDG/ { e++ }
  /UTMAREDG/ { f++ }
  /UBLKCP|UBLKRED/ { g++ }
  /HMMA/ && !/UTCHMMA/ { h++ }
  function emit() { if (a+c+d+e+f+g+h > 0) printf "%-110s %8d %6d %6d %6d %8d %8d %8d %6d\n", substr(name,1,110), a, b, c, d, e, f, g, h }
  END { if (name != "") emit() }' | sort | c++filt 2>/dev/null || true
