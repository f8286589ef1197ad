// gt_kernels.cuh -- global open-addressing table: build, overflow, probe, sample.
//
// Replaces the TBB loops of AtomicHashBuild.hpp:37-85 and NoCCHashBuild.hpp:37-81
// (and, for the sampling round, HTMHashBuild.hpp:100-156).  Same table geometry
// as the reference: pow2 slots (2*|R| at the default scaleOutput), identity hash
// slot = key & mask, linear probing inside a window of probeLength slots, 0 =
// empty.  What differs, on purpose:
//   * tuples whose budget runs out are not dropped from the join: they go to the
//     overflow list (the reference's conflicts[], :62-63) and from there into a
//     second, Fibonacci-hashed table that the probe consults only for keys whose
//     window holds no empty slot.  The join result is therefore the TRUE
//     equi-join for any key multiplicity, while "conflicts" and the window-only
//     match count keep the reference's meaning.
//   * a lost claim moves on to the next free slot (the reference burns an attempt
//     and stays, :50-54).
//   * the probe masks slot+1 (the reference does not, :78-82).
//
// B200 shape of the insert (round 2).  The reference's compare_exchange (:47-58)
// needs the slot's old value, so a CAS on the table itself drags every 32-byte
// sector of the (DRAM-resident, 2 GiB at |R| = 2^27) table through L2 before it
// is written: 24 B of DRAM traffic per tuple for 16 algorithmic, and one L2
// atomic per 8 bytes.  Here the atomic is taken on a CLAIM BITMAP instead -- one
// bit per slot, slots/8 bytes, 32 MB at 2^27: L2-resident -- and the tuple is then
// written with a plain store:
//   * the claim of a warp's load instruction is warp-cooperative: lanes whose
//     slots fall into the same 32-slot word (every lane of a sorted or locally
//     shuffled relation) are found with one shuffle + ballot, their bits are
//     OR-reduced (REDUX) and ONE atomicOr per word claims them all; its return
//     value tells every lane whether its bit was free;
//   * a lane that lost sees the word's whole occupancy and goes straight to the
//     next free slot of its window instead of stepping one CAS round trip at a
//     time;
//   * the stores of a warp to consecutive slots leave as full sectors, so the
//     table is written without ever being read: 16 B/tuple of DRAM traffic.
#pragma once
#include "hj_common.cuh"

namespace hj {

constexpr int kGtThreads = 256;

// Which tuples a launch covers: nslices segments of seg_len tuples, segment s
// starting at s*slice_stride + slice_off (the 64 work slices of
// AtomicHashBuild.hpp:37 / the K*16384 sample prefix of HTMHashBuild.hpp:100-113).
struct SegMap {
  u64 n;            // |R|
  u64 seg_len;      // tuples per segment
  u64 slice_stride; // distance between segment starts
  u64 slice_off;    // offset of the segment inside its slice
  u64 nslices;
};

__device__ __forceinline__ u32 ld_relaxed_u32(const u32* p) {
  u32 r;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(r) : "l"(p) : "memory");
  return r;
}

// Warp-cooperative claim of one slot per lane in the claim bitmap.  Lanes whose slots share a
// 32-slot word AND sit in consecutive lanes (a run) are served by one atomicOr from the run's
// first lane; a run whose members ask for the same bit twice (duplicate keys), and every lane
// that is alone with its word, claims for itself.  Aggregation is an optimisation only: any
// grouping gives the same table.  `view` = occupancy of the word right after the attempt.
// Must be called by all 32 lanes (want = false: nothing to claim).
__device__ __forceinline__ bool claim_slot(u32* __restrict__ claim, u64 slot, bool want, u32& view) {
  const u32 lane = lane_id();
  const u32 word = want ? (u32)(slot >> 5) : 0xffffffffu - lane;  // dummies never form a run
  const u32 req = want ? 1u << ((u32)slot & 31u) : 0u;
  const u32 prev = __shfl_up_sync(0xffffffffu, word, 1);
  const u32 heads = __ballot_sync(0xffffffffu, lane == 0 || word != prev);
  const u32 le = 0xffffffffu >> (31u - lane);              // lanes <= mine
  const u32 start = 31u - __clz(heads & le);               // first lane of my run
  const u32 above = heads & ~le;
  const u32 end = above ? __ffs(above) - 1u : 32u;         // one past its last lane
  const u32 peers = (end == 32u ? 0xffffffffu : (1u << end) - 1u) & ~((1u << start) - 1u);
  const u32 ormask = __reduce_or_sync(peers, req);
  u32 old;
  if (__popc(ormask) == __popc(peers)) {  // uniform over the run: every member wants its own bit
    old = 0;
    if (lane == start) old = atomicOr(claim + word, ormask);
    old = __shfl_sync(peers, old, start);
    view = old | ormask;  // (a run of this kind has no dummy lanes)
  } else if (want) {
    old = atomicOr(claim + word, req);
    view = old | req;
  } else {
    old = 0xffffffffu;  // `view` stays what the caller had (the sampling round's peek)
  }
  return want && !(old & req);
}

// IPT tuples per thread ("transaction size": how many inserts a thread has in
// flight before it looks at any outcome), loaded as IPT coalesced 8-byte loads so
// that consecutive lanes hold consecutive tuples (the claim runs above).
// STATS: the sampling round of the adaptive switch -- peeks at the word before
// claiming (only a load-then-claim can tell a lost race from an old occupant).
template <int IPT, bool STATS>
__global__ void __launch_bounds__(kGtThreads, 4)
gt_build_kernel(const u64* __restrict__ R, SegMap sm, u64* __restrict__ table, u32* __restrict__ claim, u64 mask,
                u64 key_mask, u32 L, u64* __restrict__ ovf_list, u64* __restrict__ ctr) {
  constexpr int TILE = kGtThreads * IPT;
  const u32 tid = threadIdx.x;
  const u64 chunks_per_seg = (sm.seg_len + TILE - 1) / TILE;
  const u64 nchunks = chunks_per_seg * sm.nslices;
  u32 retries = 0, steps = 0;
  u64 sectors = 0, sampled = 0, span = 0, chunks = 0;

  for (u64 c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const u64 slice = c / chunks_per_seg;
    const u64 w = c - slice * chunks_per_seg;
    const u64 seg_begin = slice * sm.slice_stride + sm.slice_off;
    u64 seg_end = seg_begin + sm.seg_len;
    if (seg_end > sm.n) seg_end = sm.n;
    const u64 base = seg_begin + w * TILE;

    u64 v[IPT];
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      const u64 i = base + (u64)j * kGtThreads + tid;
      v[j] = i < seg_end ? ld_stream(R + i) : 0ull;
    }

    if (STATS) {
      // working-set signal: how wide is the band of table slots one CTA chunk touches?
      __shared__ u64 smin[kGtThreads / 32], smax[kGtThreads / 32];
      u64 lo = ~0ull, hi = 0ull;
#pragma unroll
      for (int j = 0; j < IPT; j++)
        if (v[j]) {
          const u64 hm = (v[j] & key_mask) & mask;
          lo = hm < lo ? hm : lo;
          hi = hm > hi ? hm : hi;
        }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
      }
      if (lane_id() == 0) { smin[tid >> 5] = lo; smax[tid >> 5] = hi; }
      __syncthreads();
      if (tid == 0) {
        for (int w2 = 1; w2 < kGtThreads / 32; w2++) { lo = min(lo, smin[w2]); hi = max(hi, smax[w2]); }
        if (hi >= lo) { span += hi - lo + 1; chunks += 1; }
      }
      __syncthreads();
    }

    u32 att[IPT];  // offset from home of the next slot to try
    u32 pend = 0;  // bit j: tuple j still has to be placed
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      att[j] = 0;
      if (v[j]) pend |= 1u << j;
    }
    u32 failed_bits = 0;

    // Rounds: every pending tuple of the batch makes its next attempt.  AtomicHashBuild.hpp:46-60
    // with the claim bitmap standing in for the slot's old value.
#pragma unroll 1
    while (__any_sync(0xffffffffu, pend != 0)) {  // warp-uniform: the claim is a warp collective
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        bool want = pend >> j & 1u;
        if (!__any_sync(0xffffffffu, want)) continue;
        const u64 slot = (((v[j] & key_mask) & mask) + att[j]) & mask;
        u32 view = 0;
        bool occupied = false;
        if (STATS) {  // peek: an occupied slot costs a step, not a retry
          if (want) {
            view = ld_relaxed_u32(claim + (slot >> 5));
            occupied = view >> ((u32)slot & 31u) & 1u;
          }
        }
        const bool won = claim_slot(claim, slot, want && !occupied, view);
        if (STATS && want && !occupied && !won) ++retries;  // free when looked at, taken when claimed
        if (won) {
          table[slot] = v[j];
          pend &= ~(1u << j);
          want = false;
        }
        // losers: tuples of one load instruction that lost the SAME slot (duplicate keys sit next to
        // each other in sorted / locally shuffled inputs) would chase each other down the window one
        // round at a time; rank them and let rank r take the (r+1)-th free slot above.  The lower
        // ranks take the slots in between, so the probe's walk from home never meets a gap.
        if (__any_sync(0xffffffffu, want)) {
          // same home AND same attempt count: the same slot with the same budget left -- what the
          // lower ranks skip to, they also reach (or they overflow, and then so does every higher rank)
          const u32 same = __match_any_sync(
              0xffffffffu, want ? ((((v[j] & key_mask) & mask) << 11) | att[j]) : ~0ull - lane_id());
          if (want) {
            ++steps;
            const u32 b = (u32)slot & 31u;
            u32 freeb = b == 31u ? 0u : (~view & (0xffffffffu << (b + 1u)));
            u32 r = __popc(same & ((1u << lane_id()) - 1u));
            while (r && freeb) { freeb &= freeb - 1u; --r; }
            att[j] += freeb ? (u32)__ffs(freeb) - 1u - b : 32u - b;  // none left in this word: on to the next
            if (att[j] >= L) {  // AtomicHashBuild.hpp:62-63: budget exhausted
              pend &= ~(1u << j);
              failed_bits |= 1u << j;
            }
          }
        }
      }
    }

    if (STATS) {
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        // locality signal: distinct 32 B table sectors this warp's inserts start in
        const u64 sec = v[j] ? (((v[j] & key_mask) & mask) >> 2) : ~0ull;
        const u32 peers = __match_any_sync(0xffffffffu, sec);
        const bool leader = v[j] && (__ffs(peers) - 1 == (int)lane_id());
        const u32 leaders = __ballot_sync(0xffffffffu, leader);
        const u32 valid = __ballot_sync(0xffffffffu, v[j] != 0ull);
        if (lane_id() == 0) {
          sectors += __popc(leaders);
          sampled += __popc(valid);
        }
      }
    }
    // conflicts[] append: one atomicAdd per warp and tile claims the range (a claim per load
    // instruction put ~n/32 atomics on one address)
    if (__any_sync(0xffffffffu, failed_bits != 0)) {
      const u32 mine = __popc(failed_bits);
      u32 incl = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const u32 y = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)lane_id() >= o) incl += y;
      }
      u64 base2 = 0;
      if (lane_id() == 31) base2 = atomicAdd(ctr + C_CONFLICTS, (u64)incl);
      base2 = __shfl_sync(0xffffffffu, base2, 31);
      u64 dst = base2 + (incl - mine);
#pragma unroll
      for (int j = 0; j < IPT; j++)
        if (failed_bits >> j & 1u) ovf_list[dst++] = v[j];
    }
  }

  if (STATS) {
    u64 s[4] = {retries, steps, sectors, sampled};
    u64* const d[4] = {ctr + C_CAS_RETRIES, ctr + C_PROBE_STEPS, ctr + C_SECTORS, ctr + C_SAMPLED};
    block_sum_atomic<4, kGtThreads>(s, d);
    if (tid == 0 && chunks) {
      atomicAdd(ctr + C_SPAN, span);
      atomicAdd(ctr + C_CHUNKS, chunks);
    }
  } else {
    u64 s[1] = {steps};
    u64* const d[1] = {ctr + C_PROBE_STEPS};
    block_sum_atomic<1, kGtThreads>(s, d);
  }
}

// NoCCHashBuild.hpp:43-58: the same table without any synchronisation -- load the slot, store if
// it looked empty (racy by design; exact for unique keys under the identity hash).  Never touches
// the claim bitmap.
template <int IPT>
__global__ void __launch_bounds__(kGtThreads)
gt_nocc_build_kernel(const u64* __restrict__ R, SegMap sm, u64* __restrict__ table, u64 mask, u64 key_mask, u32 L,
                     u64* __restrict__ ovf_list, u64* __restrict__ ctr) {
  constexpr int TILE = kGtThreads * IPT;
  const u32 tid = threadIdx.x;
  const u64 chunks_per_seg = (sm.seg_len + TILE - 1) / TILE;
  const u64 nchunks = chunks_per_seg * sm.nslices;
  u32 steps = 0;
  for (u64 c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const u64 slice = c / chunks_per_seg;
    const u64 w = c - slice * chunks_per_seg;
    const u64 seg_begin = slice * sm.slice_stride + sm.slice_off;
    u64 seg_end = seg_begin + sm.seg_len;
    if (seg_end > sm.n) seg_end = sm.n;
    const u64 base = seg_begin + w * TILE;
    u64 v[IPT];
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      const u64 i = base + (u64)j * kGtThreads + tid;
      v[j] = i < seg_end ? ld_stream(R + i) : 0ull;
    }
    u32 att[IPT], pend = 0, failed_bits = 0;
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      att[j] = 0;
      if (v[j]) pend |= 1u << j;
    }
#pragma unroll 1
    while (pend) {
      u64 res[IPT];
#pragma unroll
      for (int j = 0; j < IPT; j++)
        if (pend >> j & 1u) res[j] = ld_relaxed(table + ((((v[j] & key_mask) & mask) + att[j]) & mask));
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        if (pend >> j & 1u) {
          if (res[j] == 0ull) {
            table[(((v[j] & key_mask) & mask) + att[j]) & mask] = v[j];
            pend &= ~(1u << j);
          } else {
            ++steps;
            if (++att[j] >= L) {
              pend &= ~(1u << j);
              failed_bits |= 1u << j;
            }
          }
        }
      }
    }
    if (failed_bits) {  // rare (never with unique keys): one claim per lane
      u64 dst = atomicAdd(ctr + C_CONFLICTS, (u64)__popc(failed_bits));
#pragma unroll
      for (int j = 0; j < IPT; j++)
        if (failed_bits >> j & 1u) ovf_list[dst++] = v[j];
    }
  }
  u64 s[1] = {steps};
  u64* const d[1] = {ctr + C_PROBE_STEPS};
  block_sum_atomic<1, kGtThreads>(s, d);
}

// ---- overflow table ---------------------------------------------------------
// Sized from the device-side conflict count so no host round trip is needed
// between build and probe; both kernels return at once when nothing overflowed.
// ovfmap: one bit per HOME slot, set when a tuple of that home overflowed -- the
// probe consults the overflow table only for those homes (on the reference's
// "uniform" input 85 % of the full windows have no overflowed tuple behind them).
__global__ void __launch_bounds__(256) gt_ovf_clear_kernel(u64* __restrict__ ovf_table, u32* __restrict__ ovfmap,
                                                            u64 map_words, const u64* __restrict__ ctr,
                                                            u64 capacity) {
  const u64 count = ctr[C_CONFLICTS];
  if (count == 0) return;
  u64 slots = ovf_slots_for(count);
  if (slots > capacity) slots = capacity;
  ulonglong2* t2 = reinterpret_cast<ulonglong2*>(ovf_table);
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < slots / 2; i += (u64)gridDim.x * blockDim.x)
    t2[i] = make_ulonglong2(0ull, 0ull);
  uint4* m4 = reinterpret_cast<uint4*>(ovfmap);
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < map_words / 4; i += (u64)gridDim.x * blockDim.x)
    m4[i] = make_uint4(0u, 0u, 0u, 0u);
}

constexpr int kOvfIPT = 4;
__global__ void __launch_bounds__(256) gt_ovf_insert_kernel(const u64* __restrict__ ovf_list,
                                                             u64* __restrict__ ovf_table, u32* __restrict__ ovfmap,
                                                             const u64* __restrict__ ctr, u64 mask, u64 key_mask,
                                                             u64 capacity) {
  const u64 count = ctr[C_CONFLICTS];
  if (count == 0) return;
  u64 slots = ovf_slots_for(count);
  if (slots > capacity) slots = capacity;
  const u64 omask = slots - 1;
  const u64 stride = (u64)gridDim.x * blockDim.x;
  for (u64 i0 = (u64)blockIdx.x * blockDim.x + threadIdx.x; i0 < count; i0 += stride * kOvfIPT) {
    u64 v[kOvfIPT], slot[kOvfIPT];
    u32 pend = 0;
#pragma unroll
    for (int j = 0; j < kOvfIPT; j++) {
      const u64 i = i0 + (u64)j * stride;
      v[j] = i < count ? ovf_list[i] : 0ull;
      if (v[j]) {
        pend |= 1u << j;
        const u64 home = (v[j] & key_mask) & mask;
        atomicOr(ovfmap + (home >> 5), 1u << ((u32)home & 31u));
        slot[j] = ovf_hash(v[j] & key_mask, omask);
      }
    }
    // rounds: kOvfIPT independent CAS in flight per thread (load <= 0.5: terminates)
#pragma unroll 1
    while (pend) {
      u64 res[kOvfIPT];
#pragma unroll
      for (int j = 0; j < kOvfIPT; j++)
        if (pend >> j & 1u) res[j] = atomicCAS(ovf_table + slot[j], 0ull, v[j]);
#pragma unroll
      for (int j = 0; j < kOvfIPT; j++)
        if (pend >> j & 1u) {
          if (res[j] == 0ull) pend &= ~(1u << j);
          else slot[j] = (slot[j] + 1) & omask;
        }
    }
  }
}

// Locality pre-check of the adaptive switch: the span signal of the sampling round WITHOUT the table
// (read-only, same chunk geometry as gt_build_kernel<2, true>).  If the slots of a CTA chunk
// already lie too far apart the radix join is taken at once and the table is never cleared -- at
// |R| = 2^27 the clear of the 2 GiB table alone cost 330 us of a 2.6 ms join.
__global__ void __launch_bounds__(kGtThreads)
gt_span_kernel(const u64* __restrict__ R, SegMap sm, u64 mask, u64 key_mask, u64* __restrict__ ctr) {
  constexpr int IPT = 2, TILE = kGtThreads * IPT;
  const u32 tid = threadIdx.x;
  const u64 chunks_per_seg = (sm.seg_len + TILE - 1) / TILE;
  const u64 nchunks = chunks_per_seg * sm.nslices;
  __shared__ u64 smin[kGtThreads / 32], smax[kGtThreads / 32];
  u64 span = 0, chunks = 0, sampled = 0;
  for (u64 c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const u64 slice = c / chunks_per_seg;
    const u64 w = c - slice * chunks_per_seg;
    const u64 seg_begin = slice * sm.slice_stride + sm.slice_off;
    u64 seg_end = seg_begin + sm.seg_len;
    if (seg_end > sm.n) seg_end = sm.n;
    const u64 base = seg_begin + w * TILE;
    u64 lo = ~0ull, hi = 0ull;
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      const u64 i = base + (u64)j * kGtThreads + tid;
      const u64 v = i < seg_end ? ld_stream(R + i) : 0ull;
      if (v) {
        const u64 hm = (v & key_mask) & mask;
        lo = hm < lo ? hm : lo;
        hi = hm > hi ? hm : hi;
        sampled++;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
      hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lane_id() == 0) { smin[tid >> 5] = lo; smax[tid >> 5] = hi; }
    __syncthreads();
    if (tid == 0) {
      for (int w2 = 1; w2 < kGtThreads / 32; w2++) { lo = min(lo, smin[w2]); hi = max(hi, smax[w2]); }
      if (hi >= lo) { span += hi - lo + 1; chunks += 1; }
    }
    __syncthreads();
  }
  u64 sv[1] = {sampled};
  u64* const d[1] = {ctr + C_SAMPLED};
  block_sum_atomic<1, kGtThreads>(sv, d);
  if (tid == 0 && chunks) {
    atomicAdd(ctr + C_SPAN, span);
    atomicAdd(ctr + C_CHUNKS, chunks);
  }
}

// ---- probe --------------------------------------------------------------------
// AtomicHashBuild.hpp:75-83 per S tuple, plus the overflow table.  Per-thread
// counts are reduced warp -> CTA -> one atomicAdd per CTA ("warp-aggregated").
//
// Two paths, chosen from the build's own device-side counters (no host round trip):
//   * home-only: the build displaced nothing (probe_steps == 0, conflicts == 0), so every R tuple
//     sits in its home slot and a key can only match there -- ONE coalesced 8-byte read per S
//     tuple: the 16 B/tuple of SURVEY 8d exactly.  Every unique-key input (sorted, shuffle,
//     local_shuffle) takes this path.
//   * general: the window is read as whole 32-byte sectors (one LDG.256 for the sector of the home
//     slot, a second one only if the walk runs over its end), then, for windows without an empty
//     slot whose home has overflowed tuples (ovfmap), the overflow table.
__device__ __forceinline__ void ld_sector(const u64* p, u64 (&c)[4]) {
  asm volatile("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];"
               : "=l"(c[0]), "=l"(c[1]), "=l"(c[2]), "=l"(c[3]) : "l"(p));
}

template <int IPT>
__global__ void __launch_bounds__(kGtThreads, 4)
gt_probe_kernel(const u64* __restrict__ S, u64 m, const u64* __restrict__ table, u64 mask, u64 key_mask, u32 L,
                const u64* __restrict__ ovf_table, u64 ovf_capacity, const u32* __restrict__ ovfmap,
                u64* __restrict__ ctr) {
  constexpr int TILE = kGtThreads * IPT;
  const u32 tid = threadIdx.x;
  const u64 ovf_count = ctr[C_CONFLICTS];
  const bool home_only = ovf_count == 0 && ctr[C_PROBE_STEPS] == 0;
  u64 omask = 0;
  if (ovf_count) {
    u64 slots = ovf_slots_for(ovf_count);
    if (slots > ovf_capacity) slots = ovf_capacity;
    omask = slots - 1;
  }
  u64 checksum = 0;
  u32 wmatches = 0, omatches = 0;  // per thread: < 2^32
  const u64 nchunks = (m + TILE - 1) / TILE;

  for (u64 c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const u64 base = c * TILE;
    u64 s[IPT];
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      const u64 i = base + (u64)j * kGtThreads + tid;
      s[j] = i < m ? ld_stream(S + i) : 0ull;
    }
    if (home_only) {
      u64 cur[IPT];
#pragma unroll
      for (int j = 0; j < IPT; j++) cur[j] = s[j] ? __ldg(table + ((s[j] & key_mask) & mask)) : 0ull;
#pragma unroll
      for (int j = 0; j < IPT; j++)
        if (cur[j] != 0ull && (cur[j] & key_mask) == (s[j] & key_mask)) { ++wmatches; checksum += cur[j] + s[j]; }
      continue;
    }
    // general path, in rounds of one sector per tuple still walking
    u32 pend = 0, full = 0;  // bit j: still walking / window exhausted without an empty slot
    u32 off[IPT];            // slots of the window already walked
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      off[j] = 0;
      if (s[j]) pend |= 1u << j;
    }
#pragma unroll 1
    while (pend) {
      u64 a[IPT][4];
#pragma unroll
      for (int j = 0; j < IPT; j++)
        if (pend >> j & 1u) ld_sector(table + (((((s[j] & key_mask) & mask) + off[j]) & mask) & ~3ull), a[j]);
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        if (pend >> j & 1u) {
          const u64 key = s[j] & key_mask;
          const u32 i0 = (u32)((key & mask) + off[j]) & 3u;
          bool stop = false;
#pragma unroll
          for (u32 i = 0; i < 4; i++) {
            if (i >= i0 && !stop) {
              if (off[j] >= L) { stop = true; full |= 1u << j; }
              else if (a[j][i] == 0ull) stop = true;
              else {
                if ((a[j][i] & key_mask) == key) { ++wmatches; checksum += a[j][i] + s[j]; }
                ++off[j];
              }
            }
          }
          if (!stop && off[j] >= L) { stop = true; full |= 1u << j; }
          if (stop) pend &= ~(1u << j);
        }
      }
    }
    // the window is full: tuples of this home slot may have overflowed
    if (ovf_count && full) {
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        if (full >> j & 1u) {
          const u64 key = s[j] & key_mask;
          const u64 home = key & mask;
          if (!(__ldg(ovfmap + (home >> 5)) >> ((u32)home & 31u) & 1u)) continue;
          u64 os = ovf_hash(key, omask);
#pragma unroll 1
          while (true) {
            const u64 o = __ldg(ovf_table + os);
            if (o == 0ull) break;
            if ((o & key_mask) == key) { ++omatches; checksum += o + s[j]; }
            os = (os + 1) & omask;
          }
        }
      }
    }
  }
  u64 v[3] = {(u64)wmatches + omatches, checksum, (u64)wmatches};
  u64* const d[3] = {ctr + C_MATCHES, ctr + C_CHECKSUM, ctr + C_WINDOW_MATCHES};
  block_sum_atomic<3, kGtThreads>(v, d);
}

// ---- verification sums (AtomicHashBuild.hpp:90-119; untimed there and here) -----
__global__ void __launch_bounds__(256) sum_u64_kernel(const u64* __restrict__ a, u64 n, u64* __restrict__ dst) {
  u64 s = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) s += a[i];
  u64 v[1] = {s};
  u64* const d[1] = {dst};
  block_sum_atomic<1, 256>(v, d);
}
// sum of the first ctr[C_CONFLICTS] entries of the overflow list
__global__ void __launch_bounds__(256) sum_ovf_kernel(const u64* __restrict__ ovf_list, const u64* __restrict__ ctr,
                                                       u64* __restrict__ dst) {
  const u64 n = ctr[C_CONFLICTS];
  u64 s = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) s += ovf_list[i];
  u64 v[1] = {s};
  u64* const d[1] = {dst};
  block_sum_atomic<1, 256>(v, d);
}

__global__ void __launch_bounds__(256) fill_zero_kernel(ulonglong2* __restrict__ p, u64 n2) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (u64)gridDim.x * blockDim.x)
    p[i] = make_ulonglong2(0ull, 0ull);
}

}  // namespace hj
