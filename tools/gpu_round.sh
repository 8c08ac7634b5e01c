#!/bin/bash
# One gpurun call: GPU tests, bench, launch list, and ncu --set full captures of the kernels
# named in VERDICT round 1.   usage: tools/gpu_round.sh <tag> [what...]
#   what: tests bench launches ncu_bank ncu_pcm extra
TAG="$1"; shift
WHAT="${*:-tests bench launches ncu_bank ncu_pcm}"
OUT=gpurun_out
mkdir -p $OUT
NCU="ncu --set full --import-source on --clock-control none --kernel-name-base demangled"
has() { [[ " $WHAT " == *" $1 "* ]]; }
if has tests; then
  python -m pytest tests -m gpu -x -q > $OUT/${TAG}_gputests.log 2>&1; echo "tests rc=$?"; tail -3 $OUT/${TAG}_gputests.log
fi
if has bench; then
  python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; tail -c 600 $OUT/${TAG}_bench.json
fi
if has launches; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > $OUT/${TAG}_ncu_bench.log 2>&1; echo "launches rc=$?"
fi
if has ncu_bank; then
  # the headline workload itself (config 5 on one GPU): the 4th launch of the dominant kernel
  $NCU -k regex:'bank_fast_kernel<.int.0>' -s 3 -c 1 -o $OUT/${TAG}_bank_fast0 -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras > $OUT/${TAG}_cb00.log 2>&1
  $NCU -k regex:'bank_fast_kernel<.int.1>' -s 2 -c 1 -o $OUT/${TAG}_bank_fast1 -f python tools/class_bank.py 3 0 > $OUT/${TAG}_cb30.log 2>&1
  $NCU -k regex:'bank_fast_kernel<.int.2>|bank_lfo_tile' -s 4 -c 2 -o $OUT/${TAG}_bank_fast2 -f python tools/class_bank.py 3 2 > $OUT/${TAG}_cb32.log 2>&1
  $NCU -k regex:'bank_generic_kernel|bank_carry' -s 4 -c 2 -o $OUT/${TAG}_bank_generic -f python tools/class_bank.py 0 5 1024 > $OUT/${TAG}_cb05.log 2>&1
  echo "ncu_bank done"
fi
if has ncu_pcm; then
  $NCU -k regex:'mix_apply' -s 2 -c 1 -o $OUT/${TAG}_mix_apply -f python tools/pcm_target.py mix_at > $OUT/${TAG}_pcm_mix.log 2>&1
  $NCU -k regex:'ratecv' -s 2 -c 1 -o $OUT/${TAG}_ratecv -f python tools/pcm_target.py ratecv > $OUT/${TAG}_pcm_rcv.log 2>&1
  $NCU -k regex:'fadeout' -s 2 -c 1 -o $OUT/${TAG}_fadeout -f python tools/pcm_target.py fadeout > $OUT/${TAG}_pcm_fade.log 2>&1
  echo "ncu_pcm done"
fi
if has mixed; then
  # BASELINE config 3's "mixed waveforms" bank (5 x 5 kinds): who takes the time
  ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file $OUT/${TAG}_mixed_launches.csv \
    python tools/class_bank.py 9 0 > $OUT/${TAG}_mixed_ncu.log 2>&1; echo "mixed launches rc=$?"
  $NCU -k regex:'bank_fast_kernel<.int.2>' -s 2 -c 1 -o $OUT/${TAG}_mixed_fast2 -f python tools/class_bank.py 9 0 > $OUT/${TAG}_cb90.log 2>&1
  $NCU -k regex:'bank_fast_kernel<.int.1>' -s 2 -c 1 -o $OUT/${TAG}_mixed_fast1 -f python tools/class_bank.py 9 0 > $OUT/${TAG}_cb90b.log 2>&1
fi
if has extra; then
  python bench_extra.py > $OUT/${TAG}_extra.jsonl 2> $OUT/${TAG}_extra.err; echo "extra rc=$?"
fi
if has classes; then
  for c in "0 0" "3 0" "2 -1" "0 1" "3 2" "9 0"; do echo "class $c: $(python tools/class_bank.py $c 2>&1 | tail -1)"; done | tee $OUT/${TAG}_classes.log
  echo "class 0 5 1024: $(python tools/class_bank.py 0 5 1024 2>&1 | tail -1)" | tee -a $OUT/${TAG}_classes.log
  for o in mix_at ratecv fadeout add mul tomono; do python tools/pcm_target.py $o 2>&1 | tail -1; done | tee $OUT/${TAG}_pcm.log
fi
# digest every report on the box (raw page + opcode mix); the reports themselves stay only if small
for rep in $OUT/${TAG}_*.ncu-rep; do
  [ -f "$rep" ] || continue
  base="${rep%.ncu-rep}"
  ncu -i "$rep" --page raw --csv > "${base}_ncu_raw.csv" 2>/dev/null
  units=""; case "$rep" in *bank_fast0*) units=61440000;; *bank_fast1*|*bank_fast2*) units=7680000;; esac
  python tools/ncu_summary.py "$rep" $units > "${base}_summary.txt" 2>&1
  [ -n "$KEEP_REP" ] || rm -f "$rep"
done
ls -la $OUT | tail -40
