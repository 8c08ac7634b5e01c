#!/bin/bash
# Builds synthesizer_b200/libsynth_b200.so for sm_100a (cross-compiles without a GPU).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="${SB_OUT:-$HERE/../libsynth_b200.so}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --extended-lambda -Xcompiler -fPIC,-ffp-contract=off,-fvisibility=default -shared"
$NVCC $FLAGS ${SB_EXTRA:-} ${SB_PTXAS_V:+-Xptxas -v} -o "$OUT" \
    "$HERE/sb_core.cu" "$HERE/sb_osc.cu" "$HERE/sb_voicebank.cu" "$HERE/sb_pcm.cu" "$HERE/sb_host.cu" \
    -lcudart
echo "built $OUT"
