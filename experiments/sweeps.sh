#!/bin/bash
# The reference's experiment scripts (experiments/*.sh) against this library's ./main.
#
# Same sweeps, same flags, one function per reference script; the only changes are the binary
# (htm_hashjoin_b200/csrc/main instead of the reference's per-#define builds noretry / retry / adaptive /
# adaptiveWithProbe) and the algorithms that replace the CPU-only ones:
#   htm (RTM transactions)   -> adaptive (sampled contention picks global table or radix join)
#   mchashjoins --algo=PRO   -> --algo radix
#   ENABLE_PROBE builds      -> --probe 0|1
# Every run regenerates its input on the host exactly like the reference does (serial rand(), seconds at
# 2^27), so a full sweep is dominated by input generation; experiments/w_sweep.py runs the same points
# in one process and is what figs/perf.svg is drawn from.
#
#   experiments/sweeps.sh <name> [log2 rSize, default 27]     name = one of the functions below | all
HERE="$(cd "$(dirname "$0")" && pwd)"
HJ="${HJ_MAIN:-$HERE/../htm_hashjoin_b200/csrc/main}"
LOGN="${2:-27}"
N=$((2**LOGN))

AtomicsVsHTMVsNoCC() {            # reference: AtomicsVsHTMVsNoCC.sh
  for algo in nocc atomic adaptive; do
    for d in sorted shuffle; do
      $HJ --algo $algo --rSize $N --transactionSize 1 --probeLength 4 --dataDistr $d --probe 0
    done
  done
}
SizeToAbortsAndTimeSorted() {     # reference: SizeToAbortsAndTimeSorted.sh (tSize 2^0..2^12)
  for i in $(seq 0 12); do
    $HJ --algo adaptive --rSize $N --transactionSize $((2**i)) --probeLength 4 --dataDistr sorted --probe 0
  done
}
SizeToAbortsAndTimeShuffled() {   # reference: SizeToAbortsAndTimeShuffled.sh (tSize 2^0..2^9)
  for i in $(seq 0 9); do
    $HJ --algo adaptive --rSize $N --transactionSize $((2**i)) --probeLength 4 --dataDistr shuffle --probe 0
  done
}
TSizeAndShuffleWindowstoTime() {  # reference: TSizeAndShuffleWindowstoTime.sh / adaptive.sh (tSize x W grid)
  for t in 1 4 8 16; do
    for i in $(seq 0 $LOGN); do
      $HJ --algo atomic --rSize $N --transactionSize $t --probeLength 4 --dataDistr local_shuffle --shuffleRange $((2**i)) --probe 0
    done
  done
}
adaptive() {                      # reference: adaptive2.sh (the adaptive binary over W)
  for i in $(seq 0 $LOGN); do
    $HJ --algo adaptive --rSize $N --probeLength 4 --transactionSize 16 --dataDistr local_shuffle --shuffleRange $((2**i)) --probe 0
  done
}
motivation() {                    # reference: motivation.sh (build only: PRO, nocc, atomic, htm over W)
  for algo in radix nocc atomic adaptive; do
    for i in $(seq 0 $LOGN); do
      $HJ --algo $algo --transactionSize 16 --rSize $N --probeLength 4 --dataDistr local_shuffle --shuffleRange $((2**i)) --probe 0
    done
  done
}
probe() {                         # reference: probe.sh (build + probe over W)
  for algo in nocc atomic adaptive radix; do
    for i in $(seq 0 $LOGN); do
      $HJ --algo $algo --transactionSize 16 --rSize $N --probeLength 4 --dataDistr local_shuffle --shuffleRange $((2**i)) --probe 1
    done
  done
}
ALL="motivation SizeToAbortsAndTimeSorted SizeToAbortsAndTimeShuffled TSizeAndShuffleWindowstoTime AtomicsVsHTMVsNoCC adaptive probe"
case "$1" in
  all) for s in $ALL; do $s; done ;;
  "") echo "usage: $0 <$(echo $ALL | tr ' ' '|')|all> [log2 rSize]"; exit 1 ;;
  *) if declare -F "$1" >/dev/null; then "$1"; else echo "unknown sweep $1"; exit 1; fi ;;
esac
