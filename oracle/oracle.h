/*
 * oracle.h -- CPU restatement of HTM-HashJoin's build/probe hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load liboracle.so, and only as the checker.
 * The product path (htm_hashjoin_b200/csrc) never links or calls this.
 *
 * Parity status: PINNED for unique-key inputs (reference logs
 * experiments/new_backup/ logs: inputSum 9007199321849856, totalMatches
 * 134217728, conflicts 0, PRO Results 549688705024) and against the
 * reference itself compiled into oracle/_ref (see oracle/Makefile).
 * For duplicate-key inputs ("uniform") the reference's own conflicts /
 * totalMatches depend on thread schedule; the oracle pins the serial
 * (1-thread) schedule of the reference and the true equi-join.
 *
 * Every function cites the reference file:line it follows (paths relative
 * to the reference repo root).
 */
#ifndef HJ_ORACLE_H
#define HJ_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- inputs ------------------------------------------------------------ */

/* include/DataGen.hpp:26-122.  dist in {uniform,random,sorted,shuffle,
 * local_shuffle}.  Returns 0, or -1 for an unknown distribution
 * (the reference prints "Unknown distribution" and exit(1)s, :117-120). */
int orc_generate_data(const char *dist, uint64_t n, uint64_t distinct_keys,
                      int local_shuffle_range, uint64_t *out);

/* mc/src/generator.c:117-126 + :84-93 (random_unique_gen + knuth_shuffle),
 * seeded like mc/src/main.c:362 (seed_generator -> srand(seed)).
 * Tuples are written as u64 = key | (payload<<32) with payload = 0. */
void orc_mc_pk(uint64_t *out, uint64_t n, unsigned seed);
/* mc/src/generator.c:409-445 create_relation_fk: floor(m/maxid) independent
 * permutations of 1..maxid followed by a permutation of 1..(m%maxid). */
void orc_mc_fk(uint64_t *out, uint64_t m, uint64_t maxid, unsigned seed);

/* ---- global-table path -------------------------------------------------- */

typedef struct orc_gt_result {
  uint64_t conflicts;      /* AtomicHashBuild.hpp:121-124 */
  uint64_t total_matches;  /* AtomicHashBuild.hpp:126-131 (probe window only) */
  uint64_t input_sum;      /* :90-98 */
  uint64_t table_sum;      /* :100-107 */
  uint64_t overflow_sum;   /* sum of conflicts[] actually written (:62-63) */
  uint64_t ref_output_sum; /* what the reference PRINTS (:151) incl. bug §9-3/4 */
} orc_gt_result;

/* Serial (input order) restatement of AtomicHashBuild.hpp:37-85; with one
 * thread a CAS never fails, so this is also NoCCHashBuild.hpp:37-81.
 * nocc != 0 reproduces NoCC's checksum quirk (NoCCHashBuild.hpp:94).
 * S may be NULL (build only, ENABLE_PROBE=0). */
int orc_gt_build_probe_serial(const uint64_t *R, uint64_t n,
                              const uint64_t *S, uint64_t m,
                              uint32_t scale_output, uint32_t num_partitions,
                              uint32_t probe_length, int nocc,
                              orc_gt_result *res);

/* ---- true equi-join (what a join engine must return) -------------------- */
/* matches = sum_k cntR(k)*cntS(k); checksum = sum over matching pairs of
 * (r.value + s.value) mod 2^64.  Key = value & key_mask (all ones for the
 * reference's main, 0xffffffff for mc's {int32 key,int32 payload} tuples,
 * mc/src/types.h:31-37).  Counting follows
 * mc/src/no_partitioning_join.c:270-310 (one match per equal key pair). */
void orc_true_join(const uint64_t *R, uint64_t n, const uint64_t *S,
                   uint64_t m, uint64_t key_mask, uint64_t *matches,
                   uint64_t *checksum);
/* KEY_8B tuples (mc/src/types.h:22-37): R, S = n, m tuples of two words {key, payload}; key 0 = no tuple */
void orc_true_join16(const uint64_t *R, uint64_t n, const uint64_t *S, uint64_t m,
                     uint64_t key_mask, uint64_t *matches, uint64_t *checksum);

/* ---- radix path ---------------------------------------------------------- */

/* One radix partitioning pass on bits [shift, shift+bits) :
 * mc/src/parallel_radix_join.c:559-627 (histogram, prefix, scatter) run by a
 * single thread, no padding.  hist_out (may be NULL) receives fanout+1
 * exclusive offsets.  Stable within a partition. */
void orc_radix_partition(const uint64_t *in, uint64_t n, uint32_t shift,
                         uint32_t bits, uint64_t *out, uint64_t *offsets_out);

/* PRO "Results": sum over partitions of sum idx, idx from
 * bucket_chaining_join (mc/src/parallel_radix_join.c:231-283) with
 * N = NEXT_POW_2(|R_p|) and MASK = (N-1) << radix_bits. */
uint64_t orc_pro_checksum(const uint64_t *R, uint64_t n, uint32_t radix_bits);

/* Partition R and S on the low radix_bits (1 or 2 passes give the same
 * partitions) and join each partition pair with histogram_join
 * (mc/src/parallel_radix_join.c:306-369).  Returns the match count. */
uint64_t orc_radix_join(const uint64_t *R, uint64_t n, const uint64_t *S,
                        uint64_t m, uint32_t radix_bits, uint32_t passes,
                        uint64_t *checksum);

/* ---- HTM bucket table (semantics source for the adaptive signal) -------- */
/* Serial restatement of HTMHashBuild.hpp:157-215 with no aborts: 3-tuple
 * buckets, slot = (key/3) & mask; returns the number of bucket-full
 * conflicts, writes table_sum + conflict_sum. */
uint64_t orc_htm_build_serial(const uint64_t *R, uint64_t n,
                              uint64_t *table_sum, uint64_t *conflict_sum);

/* ---- counter-based generator (no reference counterpart; see oracle.c) ----- */
void orc_feistel_gen(int kind, uint64_t first, uint64_t count, uint64_t domain,
                     uint64_t seed, uint64_t *out);

/* ---- helpers -------------------------------------------------------------- */
uint64_t orc_sum(const uint64_t *a, uint64_t n);
/* possum = sum (i+1)*a[i] mod 2^64: order-sensitive fingerprint */
uint64_t orc_possum(const uint64_t *a, uint64_t n);

#ifdef __cplusplus
}
#endif
#endif
