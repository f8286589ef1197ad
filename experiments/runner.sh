#!/bin/bash
# reference: experiments/runner.sh -- every sweep N=5 times, one log per repetition (<sweep>_log<i>)
HERE="$(cd "$(dirname "$0")" && pwd)"
N="${N:-5}"
LOGN="${LOGN:-27}"
OUT="${OUT:-$HERE/results}"
mkdir -p "$OUT"
for s in motivation SizeToAbortsAndTimeSorted SizeToAbortsAndTimeShuffled TSizeAndShuffleWindowstoTime AtomicsVsHTMVsNoCC adaptive probe; do
  for i in $(seq 1 $N); do
    "$HERE/sweeps.sh" $s $LOGN > "$OUT/${s}_log$i"
  done
done
