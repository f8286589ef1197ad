#!/bin/bash
# compute-sanitizer over the small-size GPU tests: memcheck on the kernels that index by
# data (scatter flushes, region layout, bitmap claims, exchange windows), racecheck on the shared-memory
# protocols (tile sort without a trailing barrier, async-proxy fences, claim bands).
#   tools/sanitize.sh [outdir]      logs: <outdir>/r02_sanitizer_{memcheck,racecheck}.log
OUT=${1:-gpurun_out}
mkdir -p "$OUT"
cd "$(dirname "$0")/.."
MEM_K=${MEM_K:-"claim_edge_cases or probe_lengths or fanout_sweep or histogram_free or region_exchange or xchg_pipeline or wide_tuples or one_pass_beyond or zero_keys or materialize_pairs or misaligned or noalloc"}
RACE_K=${RACE_K:-"claim_edge_cases and 16 or fanout_sweep_fk and 11 or histogram_free_equals or xchg_pipeline_logical_ranks and 2-3 or wide_tuples_join_equals_oracle and duplicates"}
timeout ${SAN_TIMEOUT:-900} compute-sanitizer --tool memcheck --error-exitcode 99 --print-limit 20 \
  python -m pytest tests -m gpu -x -q -k "$MEM_K" > "$OUT/r02_sanitizer_memcheck.log" 2>&1
echo "memcheck rc=$?" >> "$OUT/r02_sanitizer_memcheck.log"
timeout ${SAN_TIMEOUT:-900} compute-sanitizer --tool racecheck --racecheck-report analysis --error-exitcode 99 --print-limit 20 \
  python -m pytest tests -m gpu -x -q -k "$RACE_K" > "$OUT/r02_sanitizer_racecheck.log" 2>&1
echo "racecheck rc=$?" >> "$OUT/r02_sanitizer_racecheck.log"
tail -5 "$OUT/r02_sanitizer_memcheck.log" "$OUT/r02_sanitizer_racecheck.log"
