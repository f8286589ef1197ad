/*
 * oracle.c -- CPU restatement of HTM-HashJoin's build/probe hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Plain C99, single threaded.
 * Each function cites the reference file:line it restates; paths are relative
 * to the reference repo root.  Nothing here is copied: the loops are
 * re-derived from the semantics and validated against the reference itself
 * (oracle/_ref, built by oracle/Makefile) and its committed logs.
 */
#include "oracle.h"

#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------ */
/* helpers                                                                   */
/* ------------------------------------------------------------------------ */

uint64_t orc_sum(const uint64_t *a, uint64_t n) {
  uint64_t s = 0;
  for (uint64_t i = 0; i < n; i++) s += a[i];
  return s;
}

uint64_t orc_possum(const uint64_t *a, uint64_t n) {
  uint64_t s = 0;
  for (uint64_t i = 0; i < n; i++) s += (i + 1) * a[i];
  return s;
}

static uint64_t pow2_ceil(uint64_t v) {
  uint64_t p = 1;
  while (p < v) p <<= 1;
  return p;
}

/* LSD radix sort of u64 values: std::sort on plain integers has exactly one
 * possible result, so any correct sort is bit-identical to DataGen.hpp:43,60. */
static void sort_u64(uint64_t *a, uint64_t n) {
  if (n < 2) return;
  uint64_t *tmp = (uint64_t *)malloc(n * sizeof(uint64_t));
  uint64_t *src = a, *dst = tmp;
  for (int pass = 0; pass < 8; pass++) {
    uint64_t cnt[257];
    memset(cnt, 0, sizeof(cnt));
    int shift = pass * 8;
    for (uint64_t i = 0; i < n; i++) cnt[((src[i] >> shift) & 0xff) + 1]++;
    if (cnt[1] == n) continue; /* all zero digit: nothing moves */
    for (int d = 0; d < 256; d++) cnt[d + 1] += cnt[d];
    for (uint64_t i = 0; i < n; i++) dst[cnt[(src[i] >> shift) & 0xff]++] = src[i];
    uint64_t *t = src; src = dst; dst = t;
  }
  if (src != a) memcpy(a, src, n * sizeof(uint64_t));
  free(tmp);
}

/* ------------------------------------------------------------------------ */
/* inputs: include/DataGen.hpp:26-122                                        */
/* ------------------------------------------------------------------------ */

/* DataGen.hpp:46-53 (also :63-70, :107-115): walk i = 0..n-2; positions that
 * already received a swapped-in element are skipped WITHOUT consuming rand();
 * swap distance = rand() % min(range, n - i). */
static void local_shuffle(uint64_t *a, uint64_t n, int range) {
  if (n < 2) return;
  unsigned char *moved = (unsigned char *)calloc(n, 1);
  for (uint64_t i = 0; i + 1 < n; i++) {
    if (moved[i]) continue;
    int64_t left = (int64_t)(n - i);
    int lim = left < (int64_t)range ? (int)left : range;
    int d = rand() % lim;
    uint64_t t = a[i];
    a[i] = a[i + d];
    a[i + d] = t;
    moved[i + d] = 1;
  }
  free(moved);
}

int orc_generate_data(const char *dist, uint64_t n, uint64_t distinct_keys,
                      int local_shuffle_range, uint64_t *out) {
  srand(0); /* DataGen.hpp:27 */
  uint32_t mod_mask = (uint32_t)distinct_keys - 1; /* :28 */
  if (strcmp(dist, "uniform") == 0) { /* :30-54 */
    for (uint64_t i = 0; i < n; i++) out[i] = ((uint32_t)rand() & mod_mask) + 1;
    sort_u64(out, n);
    local_shuffle(out, n, local_shuffle_range);
  } else if (strcmp(dist, "random") == 0) { /* :55-71 */
    for (uint64_t i = 0; i < n; i++) {
      out[i] = (uint64_t)rand();
      while (out[i] == 0) out[i] = (uint64_t)rand();
    }
    sort_u64(out, n);
    local_shuffle(out, n, local_shuffle_range);
  } else if (strcmp(dist, "sorted") == 0) { /* :78-85 */
    for (uint64_t i = 0; i < n; i++) out[i] = i + 1;
  } else if (strcmp(dist, "shuffle") == 0) { /* :86-95 */
    for (uint64_t i = 0; i < n; i++) out[i] = i + 1;
    /* std::random_shuffle (libstdc++ bits/stl_algo.h): for i = 1..n-1 swap
     * a[i] with a[rand() % (i+1)] (no swap when equal, rand still consumed) */
    for (uint64_t i = 1; i < n; i++) {
      uint64_t j = (uint64_t)rand() % (i + 1);
      if (i != j) {
        uint64_t t = out[i];
        out[i] = out[j];
        out[j] = t;
      }
    }
  } else if (strcmp(dist, "local_shuffle") == 0) { /* :96-115 */
    for (uint64_t i = 0; i < n; i++) out[i] = i + 1;
    local_shuffle(out, n, local_shuffle_range);
  } else {
    return -1; /* :117-120: reference prints "Unknown distribution", exit(1) */
  }
  return 0;
}

/* mc/src/generator.c:20 RAND_RANGE, :84-93 knuth_shuffle, :117-126 */
static void mc_unique_gen(uint64_t *out, uint64_t n) {
  for (uint64_t i = 0; i < n; i++) out[i] = i + 1;
  for (int64_t i = (int64_t)n - 1; i > 0; i--) {
    int32_t j = (int32_t)((double)rand() / ((double)RAND_MAX + 1) * (double)i);
    uint64_t t = out[i];
    out[i] = out[j];
    out[j] = t;
  }
}

void orc_mc_pk(uint64_t *out, uint64_t n, unsigned seed) {
  srand(seed); /* mc/src/main.c:362 seed_generator */
  mc_unique_gen(out, n);
}

void orc_mc_fk(uint64_t *out, uint64_t m, uint64_t maxid, unsigned seed) {
  srand(seed);
  uint64_t iters = m / maxid; /* generator.c:424-430 */
  for (uint64_t i = 0; i < iters; i++) mc_unique_gen(out + maxid * i, maxid);
  uint64_t rem = m % maxid; /* :433-438 */
  if (rem) mc_unique_gen(out + maxid * iters, rem);
}

/* ------------------------------------------------------------------------ */
/* global table: AtomicHashBuild.hpp:37-85 / NoCCHashBuild.hpp:37-81         */
/* ------------------------------------------------------------------------ */

int orc_gt_build_probe_serial(const uint64_t *R, uint64_t n, const uint64_t *S,
                              uint64_t m, uint32_t scale_output,
                              uint32_t num_partitions, uint32_t probe_length,
                              int nocc, orc_gt_result *res) {
  (void)scale_output; /* parsed but ignored: AtomicHashBuild.hpp:21 */
  (void)num_partitions;
  if (!res || n == 0) return -1;
  uint64_t slots = 2 * n;
  uint64_t mask = slots - 1;
  uint64_t *table = (uint64_t *)calloc(slots + probe_length + 1, 8);
  if (!table) return -2;
  memset(res, 0, sizeof(*res));

  /* build: one thread, so the CAS at :50 always succeeds on an empty slot */
  for (uint64_t i = 0; i < n; i++) {
    uint64_t slot = R[i] & mask; /* :44 */
    uint32_t budget = probe_length;
    while (budget) {
      if (table[slot] == 0) {
        table[slot] = R[i];
        break;
      }
      slot = (slot + 1) & mask; /* :56-57 */
      budget--;
    }
    if (!budget) { /* :62-63 */
      res->conflicts++;
      res->overflow_sum += R[i];
    }
    res->input_sum += R[i];
  }

  /* probe: :75-83.  curSlot++ is NOT masked there; the padded tail of
   * `table` stays zero, which is what an in-bounds read past the end would
   * most plausibly see -- never exercised with keys <= n. */
  if (S) {
    for (uint64_t i = 0; i < m; i++) {
      uint64_t slot = S[i] & mask;
      uint32_t budget = probe_length;
      while (budget-- && table[slot] != 0) {
        if (table[slot] == S[i]) res->total_matches++;
        slot++;
      }
    }
  }

  for (uint64_t i = 0; i < slots; i++) res->table_sum += table[i];
  if (nocc) {
    /* NoCCHashBuild.hpp:94 sums only the first rSize slots; its conflict sum
     * (:104-113) has the same indexing as atomic's. */
    uint64_t s = 0;
    for (uint64_t i = 0; i < n; i++) s += table[i];
    res->ref_output_sum = s + res->overflow_sum;
  } else {
    res->ref_output_sum = res->table_sum + res->overflow_sum;
  }
  /* NOTE: AtomicHashBuild.hpp:113 reads conflicts[] at outputPartitionSize*i
   * although :63 wrote at inputPartitionSize*i, so with conflicts > 0 and
   * numPartitions > 1 the PRINTED outputSum is short.  ref_output_sum above
   * is the intended value (table + overflow); the printed one is obtained
   * from oracle/_ref when needed. */
  free(table);
  return 0;
}

/* ------------------------------------------------------------------------ */
/* true equi-join: counting as mc/src/no_partitioning_join.c:270-310         */
/* ------------------------------------------------------------------------ */

typedef struct {
  uint64_t key, val;
} kv_t;

static int kv_cmp(const void *a, const void *b) {
  uint64_t x = ((const kv_t *)a)->key, y = ((const kv_t *)b)->key;
  return (x > y) - (x < y);
}

void orc_true_join(const uint64_t *R, uint64_t n, const uint64_t *S, uint64_t m,
                   uint64_t key_mask, uint64_t *matches, uint64_t *checksum) {
  kv_t *r = (kv_t *)malloc((n ? n : 1) * sizeof(kv_t));
  kv_t *s = (kv_t *)malloc((m ? m : 1) * sizeof(kv_t));
  for (uint64_t i = 0; i < n; i++) { r[i].key = R[i] & key_mask; r[i].val = R[i]; }
  for (uint64_t i = 0; i < m; i++) { s[i].key = S[i] & key_mask; s[i].val = S[i]; }
  qsort(r, n, sizeof(kv_t), kv_cmp);
  qsort(s, m, sizeof(kv_t), kv_cmp);
  uint64_t i = 0, j = 0, cnt = 0, sum = 0;
  while (i < n && j < m) {
    if (r[i].key < s[j].key) i++;
    else if (r[i].key > s[j].key) j++;
    else {
      uint64_t k = r[i].key, cr = 0, cs = 0, sr = 0, ss = 0;
      while (i < n && r[i].key == k) { cr++; sr += r[i].val; i++; }
      while (j < m && s[j].key == k) { cs++; ss += s[j].val; j++; }
      cnt += cr * cs;
      sum += sr * cs + ss * cr; /* sum over pairs of (r.value + s.value) */
    }
  }
  free(r);
  free(s);
  if (matches) *matches = cnt;
  if (checksum) *checksum = sum;
}

/* The same join on mc's KEY_8B tuple, {int64 key; int64 payload} (mc/src/types.h:22-37): true equi-join
 * count (no_partitioning_join.c:270-310 counts `matches++` per equal key) and the checksum extended to the
 * four words of a pair.  T = n tuples of two words {key, payload}. */
void orc_true_join16(const uint64_t *R, uint64_t n, const uint64_t *S, uint64_t m,
                     uint64_t key_mask, uint64_t *matches, uint64_t *checksum) {
  kv_t *r = (kv_t *)malloc((n ? n : 1) * sizeof(kv_t));
  kv_t *s = (kv_t *)malloc((m ? m : 1) * sizeof(kv_t));
  uint64_t nr = 0, ns = 0;
  for (uint64_t i = 0; i < n; i++)
    if (R[2 * i] & key_mask) { r[nr].key = R[2 * i] & key_mask; r[nr].val = R[2 * i] + R[2 * i + 1]; nr++; }
  for (uint64_t i = 0; i < m; i++)
    if (S[2 * i] & key_mask) { s[ns].key = S[2 * i] & key_mask; s[ns].val = S[2 * i] + S[2 * i + 1]; ns++; }
  qsort(r, nr, sizeof(kv_t), kv_cmp);
  qsort(s, ns, sizeof(kv_t), kv_cmp);
  uint64_t i = 0, j = 0, cnt = 0, sum = 0;
  while (i < nr && j < ns) {
    if (r[i].key < s[j].key) i++;
    else if (r[i].key > s[j].key) j++;
    else {
      uint64_t k = r[i].key, cr = 0, cs = 0, sr = 0, ss = 0;
      while (i < nr && r[i].key == k) { cr++; sr += r[i].val; i++; }
      while (j < ns && s[j].key == k) { cs++; ss += s[j].val; j++; }
      cnt += cr * cs;
      sum += sr * cs + ss * cr;
    }
  }
  free(r);
  free(s);
  if (matches) *matches = cnt;
  if (checksum) *checksum = sum;
}

/* ------------------------------------------------------------------------ */
/* radix path: mc/src/parallel_radix_join.c                                  */
/* ------------------------------------------------------------------------ */

void orc_radix_partition(const uint64_t *in, uint64_t n, uint32_t shift,
                         uint32_t bits, uint64_t *out, uint64_t *offsets_out) {
  uint64_t fan = 1ull << bits;
  uint64_t msk = (fan - 1) << shift; /* :59 HASH_BIT_MODULO(K, MASK, R) */
  uint64_t *off = (uint64_t *)calloc(fan + 1, 8);
  for (uint64_t i = 0; i < n; i++) off[((in[i] & msk) >> shift) + 1]++; /* :586-589 */
  for (uint64_t p = 0; p < fan; p++) off[p + 1] += off[p];              /* :592-617 */
  if (offsets_out) memcpy(offsets_out, off, (fan + 1) * 8);
  for (uint64_t i = 0; i < n; i++) out[off[(in[i] & msk) >> shift]++] = in[i]; /* :622-626 */
  free(off);
}

/* Partition on the low radix_bits in `passes` passes: pass 1 on bits
 * [0, b/passes), pass 2 on the rest inside every pass-1 partition
 * (prj_params.h:76-78, parallel_radix_join.c:870-871,960-961).  The final
 * partition of a key is the pair (p1, p2); offsets are in p1-major order. */
static void partition_multi(const uint64_t *in, uint64_t n, uint32_t radix_bits,
                            uint32_t passes, uint64_t *out, uint64_t *offsets) {
  uint64_t fan = 1ull << radix_bits;
  if (passes <= 1 || radix_bits < 2) {
    orc_radix_partition(in, n, 0, radix_bits, out, offsets);
    return;
  }
  uint32_t b1 = radix_bits / 2, b2 = radix_bits - b1;
  uint64_t f1 = 1ull << b1, f2 = 1ull << b2;
  uint64_t *tmp = (uint64_t *)malloc((n ? n : 1) * 8);
  uint64_t *o1 = (uint64_t *)malloc((f1 + 1) * 8);
  uint64_t *o2 = (uint64_t *)malloc((f2 + 1) * 8);
  orc_radix_partition(in, n, 0, b1, tmp, o1);
  for (uint64_t p = 0; p < f1; p++) {
    orc_radix_partition(tmp + o1[p], o1[p + 1] - o1[p], b1, b2, out + o1[p], o2);
    for (uint64_t q = 0; q < f2; q++) offsets[p * f2 + q] = o1[p] + o2[q];
  }
  offsets[fan] = n;
  free(tmp); free(o1); free(o2);
}

uint64_t orc_pro_checksum(const uint64_t *R, uint64_t n, uint32_t radix_bits) {
  uint64_t fan = 1ull << radix_bits;
  uint64_t *out = (uint64_t *)malloc((n ? n : 1) * 8);
  uint64_t *off = (uint64_t *)malloc((fan + 1) * 8);
  orc_radix_partition(R, n, 0, radix_bits, out, off);
  uint64_t total = 0;
  for (uint64_t p = 0; p < fan; p++) {
    uint64_t cnt = off[p + 1] - off[p];
    if (!cnt) continue;
    uint32_t N = (uint32_t)pow2_ceil(cnt);                 /* :239-243 */
    uint32_t msk = (uint32_t)((N - 1) << radix_bits);      /* :245 */
    for (uint64_t i = off[p]; i < off[p + 1]; i++)          /* :252-259 */
      total += ((uint32_t)out[i] & msk) >> radix_bits;
  }
  free(out); free(off);
  return total;
}

/* histogram_join, mc/src/parallel_radix_join.c:306-369, on one partition */
static uint64_t hist_join(const uint64_t *Rp, uint64_t nr, const uint64_t *Sp,
                          uint64_t ns, uint32_t radix_bits, uint64_t *checksum) {
  if (!nr || !ns) return 0;
  uint64_t nh = pow2_ceil(nr) >> 2;                        /* :296-301 */
  if (nh < 4) nh = 4;
  uint32_t msk = (uint32_t)((nh - 1) << radix_bits);
  uint64_t *hist = (uint64_t *)calloc(nh + 2, 8);
  uint64_t *tmp = (uint64_t *)malloc(nr * 8);
  for (uint64_t i = 0; i < nr; i++) hist[(((uint32_t)Rp[i] & msk) >> radix_bits) + 2]++;
  for (uint64_t i = 2, s = 0; i <= nh + 1; i++) { s += hist[i]; hist[i] = s; }
  for (uint64_t i = 0; i < nr; i++) {
    uint64_t idx = (((uint32_t)Rp[i] & msk) >> radix_bits) + 1;
    tmp[hist[idx]++] = Rp[i];
  }
  uint64_t match = 0, sum = 0;
  for (uint64_t i = 0; i < ns; i++) {
    uint64_t idx = ((uint32_t)Sp[i] & msk) >> radix_bits;
    for (uint64_t j = hist[idx]; j < hist[idx + 1]; j++)
      if ((uint32_t)Sp[i] == (uint32_t)tmp[j]) { /* int32 key compare, types.h:31-37 */
        match++;
        sum += Sp[i] + tmp[j];
      }
  }
  free(hist); free(tmp);
  if (checksum) *checksum += sum;
  return match;
}

uint64_t orc_radix_join(const uint64_t *R, uint64_t n, const uint64_t *S,
                        uint64_t m, uint32_t radix_bits, uint32_t passes,
                        uint64_t *checksum) {
  uint64_t fan = 1ull << radix_bits;
  uint64_t *pr = (uint64_t *)malloc((n ? n : 1) * 8);
  uint64_t *ps = (uint64_t *)malloc((m ? m : 1) * 8);
  uint64_t *or_ = (uint64_t *)malloc((fan + 1) * 8);
  uint64_t *os = (uint64_t *)malloc((fan + 1) * 8);
  partition_multi(R, n, radix_bits, passes, pr, or_);
  partition_multi(S, m, radix_bits, passes, ps, os);
  uint64_t total = 0, sum = 0;
  for (uint64_t p = 0; p < fan; p++)
    total += hist_join(pr + or_[p], or_[p + 1] - or_[p], ps + os[p],
                       os[p + 1] - os[p], radix_bits, &sum);
  free(pr); free(ps); free(or_); free(os);
  if (checksum) *checksum = sum;
  return total;
}

/* ------------------------------------------------------------------------ */
/* HTM bucket table, serial, no aborts: HTMHashBuild.hpp:41-45,61-62,122-128 */
/* ------------------------------------------------------------------------ */

uint64_t orc_htm_build_serial(const uint64_t *R, uint64_t n, uint64_t *table_sum,
                              uint64_t *conflict_sum) {
  uint64_t nb = pow2_ceil(n / 3 + 1); /* :61-62 */
  uint64_t mask = nb - 1;
  uint64_t *tup = (uint64_t *)calloc(nb * 3, 8);
  uint32_t *cnt = (uint32_t *)calloc(nb, 4);
  uint64_t conflicts = 0, ts = 0, cs = 0;
  for (uint64_t i = 0; i < n; i++) {
    uint64_t b = (R[i] / 3) & mask; /* :122 */
    if (cnt[b] != 3) {              /* :123-125 */
      tup[b * 3 + cnt[b]++] = R[i];
      ts += R[i];
    } else {                        /* :126-127 */
      conflicts++;
      cs += R[i];
    }
  }
  free(tup); free(cnt);
  if (table_sum) *table_sum = ts;
  if (conflict_sum) *conflict_sum = cs;
  return conflicts;
}

/* ------------------------------------------------------------------------ */
/* counter-based generator for sizes beyond the reference's uint32_t/rand()  */
/* (SURVEY.md 8d, config C5).  The reference has no generator of this kind;  */
/* this is the CPU statement the device generator (hj_generate) is checked   */
/* against: kind 0 = sorted slice, kind 1 = 4-round Feistel-style bijection  */
/* of [0, 2^k) per block of `domain` indices (one independent permutation    */
/* per block, like the iterations of mc/src/generator.c:424-430).            */
/* ------------------------------------------------------------------------ */
static uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull;
  x ^= x >> 33;
  return x;
}
static uint64_t feistel_perm(uint64_t x, uint32_t k, uint64_t seed) {
  uint32_t hb = k / 2, lb = k - hb;
  uint64_t l = x & ((1ull << lb) - 1), h = x >> lb;
  if (k < 2) return x;
  for (uint32_t r = 0; r < 4; r++) {
    h ^= mix64(l + seed + r * 0x9E3779B97F4A7C15ull) & ((1ull << hb) - 1);
    l ^= mix64(h + (seed ^ 0xD1B54A32D192ED03ull) + r * 0xBF58476D1CE4E5B9ull) & ((1ull << lb) - 1);
  }
  return (h << lb) | l;
}
void orc_feistel_gen(int kind, uint64_t first, uint64_t count, uint64_t domain,
                     uint64_t seed, uint64_t *out) {
  uint32_t k = 0;
  while ((1ull << k) < domain) k++;
  uint64_t dmask = (1ull << k) - 1;
  for (uint64_t i = 0; i < count; i++) {
    uint64_t g = first + i;
    out[i] = kind == 0 ? g + 1 : feistel_perm(g & dmask, k, seed + (g >> k)) + 1;
  }
}
