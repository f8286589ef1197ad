/*
 * hj_b200.h -- C ABI of the B200-native (sm_100a) build/probe hash-join path.
 *
 * This is the drop-in boundary for HTM-HashJoin's L2 "join operator" layer:
 * the bodies of AtomicHashBuild / NoCCHashBuild / HTMHashBuild (main binary)
 * and PRO (mc/ radix join) are replaced by calls through these entry points.
 * Plain C: pointers and sizes only, no CUDA / torch / C++ types.  Paths cited
 * below are relative to the reference repository root.
 *
 * Tuple format: one 8-byte word per tuple,
 *   low 32 bits = key, high 32 bits = payload
 * which is bit-compatible with the reference's bare uint64_t values
 * (include/DataGen.hpp:26-122, values <= 2^28) and with mc's
 * struct {int32 key; int32 payload;} (mc/src/types.h:31-37).  0 is the empty
 * slot sentinel, so keys must be non-zero (DataGen.hpp:41,58,83 guarantee it).
 *
 * There is no CPU fallback behind this ABI: every entry point that computes
 * needs a CUDA device and returns HJ_ECUDA when none is usable.
 */
#ifndef HJ_B200_H
#define HJ_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HJ_ABI_VERSION 1

/* ---- status codes (reference: none; it perror()+exit()s, HTMHashBuild.hpp:66-70) */
enum {
  HJ_OK = 0,
  HJ_EINVAL = -1,    /* bad argument / bad config                        */
  HJ_ENOMEM = -2,    /* device or host allocation failed                 */
  HJ_ECUDA = -3,     /* CUDA runtime error (see hj_last_error)           */
  HJ_EOVERFLOW = -4, /* a partition does not fit the planned buffers     */
  HJ_ESTATE = -5     /* call order violated (probe before build, ...)    */
};

/* ---- which path (reference: --algo, main.cpp:99-122) */
enum {
  HJ_ALGO_ATOMIC = 0,  /* global table, CAS inserts   (AtomicHashBuild.hpp:37-67)   */
  HJ_ALGO_NOCC = 1,    /* global table, plain stores  (NoCCHashBuild.hpp:37-62)     */
  HJ_ALGO_RADIX = 2,   /* radix join                  (mc/src/parallel_radix_join.c)*/
  HJ_ALGO_ADAPTIVE = 3,/* sample, then global or radix (HTMHashBuild.hpp:100-156)   */
  HJ_ALGO_SORTMERGE = 4/* partitioned sort-merge join, the comparator of SortMerge.cpp:8-39:
                          radix-partition both sides, sort each partition pair on chip,
                          two-pointer merge; hj_join only, needs S           */
};

/* path_taken values in hj_result */
enum { HJ_PATH_NONE = 0, HJ_PATH_GLOBAL = 1, HJ_PATH_RADIX = 2, HJ_PATH_SORTMERGE = 3 };

/* Everything that is a command-line flag (main.cpp:47-62) or a compile-time
 * macro (config.h:4-16, mc/src/prj_params.h:15-22) in the reference. */
typedef struct hj_config {
  uint32_t struct_size;      /* = sizeof(hj_config), ABI check                          */
  int32_t algo;              /* HJ_ALGO_*                                               */
  int32_t device;            /* CUDA device ordinal                                     */
  uint32_t probe_length;     /* --probeLength, default 4 (main.cpp:81)                  */
  uint32_t scale_output;     /* --scaleOutput, default 2; table = pow2(scale*|R|)       */
  uint32_t num_partitions;   /* --numPartitions: CPU work split for atomic/nocc
                                (ignored on the GPU); radix FAN-OUT for radix/adaptive
                                (power of two, 0 = choose so a partition fits smem)     */
  uint32_t passes;           /* radix passes 1|2 (NUM_PASSES, prj_params.h:20), 0=auto  */
  uint32_t transaction_size; /* --transactionSize: tuples a thread inserts per batch
                                (HTM's tSize -> CAS batch / loads in flight), 0 = 2     */
  uint32_t enable_probe;     /* ENABLE_PROBE (config.h:4)                               */
  uint32_t sample_rounds;    /* K of HTMHashBuild.hpp:48-52 (K*16384 tuples / slice)    */
  uint64_t key_mask;         /* key = tuple & key_mask; 0 = all 64 bits (main binary),
                                0xffffffff for mc tuples                                */
  double adaptive_threshold; /* contention above this -> radix path; <=0 = default     */
  uint64_t pass1_bits;       /* 2-pass plans: radix bits of pass 1; 0 = floor(bits/2) as
                                prj_params.h:76-78.  The fused multi-GPU exchange uses a
                                smaller pass 1 so that G << pass1_bits stays ~128 bins    */
  uint64_t tuple_bytes;      /* 0 / 8: one 8-byte word per tuple (default).  16: mc's KEY_8B tuple
                                (mc/src/types.h:22-37) = {uint64_t key; uint64_t payload}: R / S point to
                                2 words per tuple, sizes still count tuples; key = word 0 & key_mask, key 0 =
                                empty; checksum = sum over pairs of all four words.  HJ_ALGO_RADIX and
                                HJ_ALGO_ATOMIC through hj_join / hj_build / hj_probe / hj_checksums only */
  uint64_t reserved[2];
} hj_config;

/* Per-call outcome.  The first block mirrors the JSON the reference prints
 * (AtomicHashBuild.hpp:133-152, HTMHashBuild.hpp:417-449). */
typedef struct hj_result {
  uint64_t matches;        /* true equi-join cardinality sum_k cntR(k)*cntS(k)           */
  uint64_t checksum;       /* sum over matching pairs of (r + s) mod 2^64               */
  uint64_t window_matches; /* matches inside the probe window only = the reference's
                              "totalMatches" (AtomicHashBuild.hpp:75-83,126-131)        */
  uint64_t input_sum;      /* "inputSum"  (AtomicHashBuild.hpp:90-98)                   */
  uint64_t table_sum;      /* sum of table slots (:100-107)                             */
  uint64_t overflow_sum;   /* sum of tuples that left the window (:109-119, fixed idx)  */
  uint64_t conflicts;      /* "conflicts": tuples whose probe budget ran out (:62-63)   */
  uint64_t cas_retries;    /* lost CAS races (the HTM-abort stand-in)                   */
  uint64_t probe_steps;    /* slots skipped because occupied, build side                */
  uint64_t sampled;        /* tuples in the sampling round (0 if none)                  */
  uint64_t sectors;        /* distinct 32 B table sectors touched per warp, summed      */
  double contention;       /* adaptive signal, see DESIGN.md                            */
  int32_t path_taken;      /* HJ_PATH_*                                                 */
  uint32_t radix_bits;     /* log2 fan-out actually used                                */
  uint32_t passes;         /* passes actually used                                      */
  uint32_t kernel_launches;/* kernels launched by this call                             */
  /* device time of each phase, microseconds, CUDA events on the handle's stream */
  float us_total, us_h2d, us_clear, us_sample, us_build, us_probe;
  float us_hist, us_scan, us_scatter1, us_scatter2, us_join;
  float us_reserved[5];    /* [0]: hj_xchg: device time spent waiting for the peers' copies */
  uint64_t hbm_bytes;      /* algorithmic HBM bytes of the path taken       */
  uint64_t max_partition;  /* largest R partition seen by the radix join                */
  uint64_t zero_keys;      /* input tuples equal to 0 (R and S): 0 is the empty-slot sentinel, such tuples
                              take no part in the join (mc --non-unique can emit them: bias keys by +1) */
  uint64_t reserved[5];
} hj_result;

typedef struct hj_handle hj_handle;

/* ---- lifetime -------------------------------------------------------------- */
void hj_config_default(hj_config *cfg);  /* main.cpp:78-85 defaults */
int hj_create(const hj_config *cfg, hj_handle **out);
void hj_destroy(hj_handle *h);
int hj_abi_version(void);
int hj_device_count(void);
const char *hj_strerror(int code);
const char *hj_last_error(const hj_handle *h); /* last CUDA error text */

/* Pre-size every device buffer for |R| <= max_r, |S| <= max_s so that no
 * allocation happens inside a timed call (the reference allocates before its
 * timer too, AtomicHashBuild.hpp:25-34). */
int hj_reserve(hj_handle *h, uint64_t max_r, uint64_t max_s);
/* The same for the H2D staging buffers of HOST inputs (hj_reserve cannot know where the
 * inputs will live): after both, a join of <= max_r / max_s host tuples allocates nothing. */
int hj_reserve_staging(hj_handle *h, uint64_t max_r, uint64_t max_s);
/* Device allocations the handle has made so far (a reserved handle must not add any in hj_join). */
uint64_t hj_alloc_count(const hj_handle *h);

/* ---- the hot path ------------------------------------------------------------
 * R / S may be host pointers (pageable or pinned; staged H2D inside the call)
 * or device pointers on cfg.device (used in place); detected per pointer.  The
 * radix kernels read tuples in 16-byte pairs: a device pointer that is only
 * 8-byte aligned (a slice starting at an odd tuple) is first copied to an
 * aligned staging buffer by hj_join / hj_join_materialize; the global-table
 * path reads 8 bytes at a time and takes any tuple-aligned pointer in place.
 * The building-block entry points further down (hj_radix_partition, hj_peer_*,
 * hj_join_pass1, hj_join_regions) require 16-byte aligned device pointers and
 * return HJ_EINVAL otherwise. */

/* AtomicHashBuild.hpp:37-67 / NoCCHashBuild.hpp:37-62: (re)build the global
 * table from R.  res may be NULL. */
int hj_build(hj_handle *h, const uint64_t *R, uint64_t n, hj_result *res);
/* AtomicHashBuild.hpp:69-85: probe the table built last with S. */
int hj_probe(hj_handle *h, const uint64_t *S, uint64_t m, hj_result *res);
/* Whole join, path chosen by cfg.algo; S == NULL / m == 0 means build only
 * (ENABLE_PROBE=0).  For HJ_ALGO_RADIX this is PRO (parallel_radix_join.c:1305)
 * with the probe of histogram_join (:306-369) restored. */
int hj_join(hj_handle *h, const uint64_t *R, uint64_t n, const uint64_t *S,
            uint64_t m, hj_result *res);

/* Result materialisation.  The reference only counts ("TODO: we don't
 * materialize", mc/src/no_partitioning_join.c:301, parallel_radix_join.c:359); this emits the
 * pairs.  Radix join of R and S (host or device pointers) whose every matching pair is written
 * to out_pairs as two uint64_t {r tuple, s tuple}: one dense DEVICE array (16-byte aligned) of
 * *n_pairs entries; pairs of one radix partition are contiguous, order inside a partition is
 * unspecified.  Count / scan / write: with out_pairs == NULL only the count pass runs and
 * *n_pairs says how large the buffer must be; with capacity_pairs < *n_pairs nothing is written
 * and HJ_EOVERFLOW is returned.  Needs keys of at most 32 bits and a fan-out that puts an R
 * partition into shared memory (the default plan; else HJ_EINVAL).  res->matches / checksum as
 * hj_join; the handle may have any cfg.algo. */
int hj_join_materialize(hj_handle *h, const uint64_t *R, uint64_t n, const uint64_t *S, uint64_t m,
                        uint64_t *out_pairs, uint64_t capacity_pairs, uint64_t *n_pairs, hj_result *res);
/* HTMHashBuild.hpp:100-156 sampling round: insert K*16384 tuples of each of
 * the 64 slices of R into the (cleared) global table, measure contention.
 * The sampled prefix stays in the table, as in the reference (:170). */
int hj_sample(hj_handle *h, const uint64_t *R, uint64_t n, hj_result *res);
/* Verification sums (untimed in the reference, AtomicHashBuild.hpp:90-119):
 * fills input_sum (over R, may be NULL), table_sum, overflow_sum. */
int hj_checksums(hj_handle *h, const uint64_t *R, uint64_t n, hj_result *res);

/* ---- radix building blocks (exposed for the multi-GPU exchange and tests) ----
 * One partitioning pass over bits [shift, shift+bits) of the key:
 * parallel_radix_join.c:559-627 (histogram -> prefix -> scatter).  in/out/
 * offsets are DEVICE pointers; offsets receives 2^bits+1 exclusive offsets
 * (u64).  Order inside a partition is unspecified. */
int hj_radix_partition(hj_handle *h, const uint64_t *in, uint64_t n,
                       uint32_t shift, uint32_t bits, uint64_t *out,
                       uint64_t *offsets);
/* Destination histogram for the G-way exchange: dest = ((key-1) * G) >> domain_bits
 * clamp G-1.  counts: G u64 on the device. */
int hj_exchange_partition(hj_handle *h, const uint64_t *in, uint64_t n,
                          uint32_t n_dest, uint32_t domain_bits, uint64_t *out,
                          uint64_t *offsets);

/* ---- exchange fused with radix pass 1 (multi-GPU, one process per GPU) ----
 * Replaces "partition by destination, all-to-all, then partition again": every rank scatters
 * its slice by (owner GPU of the key range, low b1 key bits) straight into the owners' pass-1
 * buffers over NVLink (peer memory mapped with CUDA IPC), i.e. the exchange IS
 * parallel_radix_partition's first pass (parallel_radix_join.c:559-627) with remote output.
 * Protocol (the collectives are the caller's: torch.distributed in exchange.py):
 *   1. hj_radix_plan            -> b1 of the local join (same on every rank)
 *   2. hj_peer_hist             -> counts[(dest << b1) | low bits] of the local slice
 *   3. all-gather the counts; every rank derives, for every (dest, low bits), the first element
 *      of ITS exclusive range in dest's buffer (exchange.peer_cursors)
 *   4. hj_peer_scatter          -> plain stores into the peers' buffers; claims are local atomics
 *   5. barrier; hj_join_pass1   -> histogram + pass 2 + per-partition join of what arrived */
int hj_radix_plan(hj_handle *h, uint64_t n_build, uint32_t *bits, uint32_t *b1,
                  uint32_t *b2, uint32_t *passes);
/* 64-byte CUDA IPC handle of a buffer from hj_device_alloc / open it in another process. */
int hj_ipc_export(hj_handle *h, const void *dptr, void *handle64);
int hj_ipc_open(hj_handle *h, const void *handle64, void **dptr);
int hj_ipc_close(hj_handle *h, void *dptr);
int hj_peer_hist(hj_handle *h, const uint64_t *in, uint64_t n, uint32_t n_dest,
                 uint32_t domain_bits, uint32_t b1, uint64_t *counts);
int hj_peer_scatter(hj_handle *h, const uint64_t *in, uint64_t n, uint32_t n_dest,
                    uint32_t domain_bits, uint32_t b1, uint64_t *const *peer_bufs,
                    uint64_t *cursors);
/* hj_peer_scatter without the final wait (hj_sync waits for it) and on at most max_ctas CTAs
 * (0 = all SMs): the scatter is NVLink-bound, so it can share the GPU with another handle's
 * hj_join_pass1 of the previous chunk of S (a join is additive over disjoint parts of S).
 * Consecutive async scatters on one handle run in order. */
int hj_peer_scatter_async(hj_handle *h, const uint64_t *in, uint64_t n, uint32_t n_dest, uint32_t domain_bits,
                          uint32_t b1, uint64_t *const *peer_bufs, uint64_t *cursors, uint32_t max_ctas);
/* offR1/offS1: DEVICE arrays of the 2^b1 + 1 pass-1 partition offsets inside R1/S1 (or NULL:
 * the library histograms what arrived). */
int hj_join_pass1(hj_handle *h, const uint64_t *R1, uint64_t n, const uint64_t *offR1,
                  const uint64_t *S1, uint64_t m, const uint64_t *offS1, hj_result *res);

/* ---- histogram-free exchange (no hj_peer_hist, no counts before the scatter) -----------------------
 * Every (source, destination, low bits) triple owns a fixed-capacity region of the destination's pass-1
 * buffer: region low * n_dest + source, region_cap tuples (even; expected size + slack, the same number
 * on all ranks, e.g. htm_hashjoin_b200/exchange.py: region_cap).  hj_peer_scatter_regions: cursors
 * (DEVICE, n_dest << b1) holds on entry the first element of this rank's region for every
 * (dest << b1) | low, on return its fill level; all-gather the cursors -- that collective is also the
 * barrier -- and give every receiver the fill levels of its regions: hj_join_regions (curR / curS: DEVICE
 * arrays of (2^b1 * n_src) entries, index low * n_src + source; n / m: tuples that arrived) runs pass 2 +
 * the per-partition join.  *overflowed != 0 (skewed keys): the surplus went to the scratch zone at
 * trash_off (16384 + 2 tuples) of the destination buffer; redo the step with the exact protocol above.
 * n_src must be a power of two and 2^b1 * n_src <= 256. */
int hj_peer_scatter_regions(hj_handle *h, const uint64_t *in, uint64_t n, uint32_t n_dest, uint32_t domain_bits,
                            uint32_t b1, uint64_t *const *peer_bufs, uint64_t *cursors, uint32_t rank,
                            uint64_t region_cap, uint64_t trash_off, uint32_t *overflowed);
int hj_join_regions(hj_handle *h, const uint64_t *R1, uint64_t n, const uint64_t *curR, uint64_t capR,
                    const uint64_t *S1, uint64_t m, const uint64_t *curS, uint64_t capS, uint32_t n_src,
                    hj_result *res);

/* ---- pipelined exchange over the copy engines (multi-GPU, round 2) ---------------------------
 * The SMs never touch NVLink: every rank scatters a STAGE of its slices (stage 0 = R, stages 1..s_chunks =
 * equal chunks of S by position; a join is additive over disjoint parts of S) LOCALLY by (owner GPU, low
 * pass-1 key bits) into per-destination blocks of fixed-capacity regions; one copy stream per rank ships block
 * `dest` to rank dest's receive WINDOW with a plain device-to-device copy (copy engines over NVLink through
 * the CUDA-IPC mapping), followed by the block's fill levels and a flag word carrying the step's epoch -- stage
 * after stage, the peers of a stage in ring order (rank+1 first), and with more than two ranks the ring is kept
 * in lock step: before shipping round k+1 a rank's copy stream waits (cuStreamWaitValue64 on a word of its own
 * window) for its successor's "round k done" token, so no GPU ever receives from two senders at once.  Blocks
 * start on 2 MiB boundaries and every allocation is a whole number of 2 MiB pages (the copy engines write into a
 * small-page allocation at 530 instead of 760 GB/s).  The compute stream scatters all stages, then stage by stage
 * waits on the device (bounded: 10 s) for the stage's flags, runs radix pass 2 on what arrived
 * (parallel_radix_join.c:402-440 with one parent per (source, low bits) region) and joins the chunks in groups
 * (one group at 2 ranks, two at 4, one per chunk from 8 on) -- the copies of the later stages run under pass 2 +
 * join of the earlier ones, and nothing but the final counter read-back synchronises with the host.
 *   hj_xchg_create   plans the stages, allocates the send blocks and this rank's window (one allocation:
 *                    export it with hj_ipc_export); the handle must be HJ_ALGO_RADIX with an explicit
 *                    2-pass plan (num_partitions, passes = 2, pass1_bits) identical on all ranks,
 *                    n_ranks a power of two, n_ranks << pass1_bits <= 256.  n_local / m_local: tuples of
 *                    R / S every rank holds (m_local divisible by 2 * s_chunks).
 *   hj_xchg_connect  peer_windows[r] = rank r's window as mapped here (hj_ipc_open; entry `rank` ignored).
 *   hj_xchg_start    enqueues the whole step (R, S: device pointers to this rank's slices); returns at once.
 *   hj_xchg_finish   waits for it; res = this rank's share of {matches, checksum} (sum them over ranks).
 *                    HJ_EOVERFLOW: a region overflowed (skewed keys) -- redo the step with the exact protocol.
 * All ranks pass the same, strictly increasing epoch, and nobody may start epoch e+1 before every rank
 * finished epoch e (the all-reduce of the result is that barrier). */
int hj_xchg_create(hj_handle *h, uint32_t n_ranks, uint32_t rank, uint64_t n_local, uint64_t m_local,
                   uint32_t s_chunks, uint32_t domain_bits, void **window, uint64_t *window_bytes);
int hj_xchg_connect(hj_handle *h, void *const *peer_windows);
int hj_xchg_start(hj_handle *h, const uint64_t *R, const uint64_t *S, uint64_t epoch);
int hj_xchg_finish(hj_handle *h, hj_result *res);

/* ---- device-side input synthesis (sizes beyond uint32_t/rand()) --
 * Counter-based generators with a bit-identical CPU restatement in oracle/:
 *   kind 0: out[i] = first + i + 1                       (sorted slice)
 *   kind 1: out[i] = perm_{seed}((first + i) mod domain) + 1, perm = 4-round
 *           Feistel-style bijection on [0, domain), domain = 2^k  (PK / FK slice)
 * out is a device pointer. */
int hj_generate(hj_handle *h, int kind, uint64_t first, uint64_t count,
                uint64_t domain, uint64_t seed, uint64_t *out);

/* ---- a join that allocates nothing ---------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * The global-table join (the atomic build + probe of hj_join on an HJ_ALGO_ATOMIC handle with the default
 * configuration: scaleOutput 2, all 64 bits of a tuple are the key) with its table, claim bitmap, overflow
 * structures and counters in a static __device__ array of the library: 2.53 MiB that every GPU which loads the
 * library holds until the process exits.  No cudaMalloc / cudaFree / allocating routine is called, no handle is
 * needed, nothing synchronises with the host: the work is enqueued on `stream` (a cudaStream_t; NULL = the default
 * stream) and out[0] = matches, out[1] = checksum are valid once the stream reaches that point.
 *   R, S, out   device pointers (8-byte aligned); |R| = n in [1, HJ_NOALLOC_MAX_R], |S| = m (any size, 0 allowed)
 *   probe_length   as --probeLength (0 = the default, 4)
 *   err, err_cap   optional buffer for the message of a failure
 * n > HJ_NOALLOC_MAX_R is rejected with HJ_EINVAL.  Every call on a GPU shares the one array and finds the last
 * call's contents in it; each call clears what it needs at its start.  Calls that overlap on the same GPU (two
 * streams, two threads) race on the array: serialise them.  The first call on a device looks the array's address
 * up (cudaGetSymbolAddress) and is therefore not capturable; later calls only launch kernels, memsets and a
 * 16-byte device-to-device copy. */
#define HJ_NOALLOC_MAX_R (1u << 16)
int hj_join_noalloc(int device, void *stream, const uint64_t *R, uint64_t n, const uint64_t *S, uint64_t m,
                    uint32_t probe_length, uint64_t *out, char *err, uint32_t err_cap);

/* ---- memory helpers so a C host needs no CUDA headers ------------------------ */
int hj_device_alloc(hj_handle *h, uint64_t bytes, void **out);
int hj_device_free(hj_handle *h, void *p);
int hj_host_alloc(uint64_t bytes, void **out); /* pinned */
int hj_host_free(void *p);
int hj_copy(hj_handle *h, void *dst, const void *src, uint64_t bytes); /* any direction, synchronous */
int hj_flush_l2(hj_handle *h); /* writes a >L2 scratch buffer (bench hygiene) */
int hj_sync(hj_handle *h);

#ifdef __cplusplus
}
#endif
#endif /* HJ_B200_H */
