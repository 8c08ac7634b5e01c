#!/bin/bash
# Builds synthesizer_b200/libsynth_b200.so for sm_100a (cross-compiles without a GPU).
# The translation units are compiled in parallel, then linked.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="${SB_OUT:-$HERE/../libsynth_b200.so}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
OBJ="${SB_OBJ:-$HERE/../../build/obj}"
mkdir -p "$OBJ"
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --extended-lambda -Xcompiler -fPIC,-ffp-contract=off,-fvisibility=default"
SRCS="sb_core sb_osc sb_voicebank sb_pcm sb_pcm32 sb_host"
pids=""
for s in $SRCS; do
    $NVCC $FLAGS ${SB_EXTRA:-} ${SB_PTXAS_V:+-Xptxas -v} -c -o "$OBJ/$s.o" "$HERE/$s.cu" &
    pids="$pids $!"
done
for p in $pids; do wait $p; done
OBJS=""
for s in $SRCS; do OBJS="$OBJS $OBJ/$s.o"; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT" $OBJS -lcudart
echo "built $OUT"
