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
// is written -- 24 B of DRAM traffic per tuple for 16 algorithmic -- and costs one
// L2 atomic per 8 bytes: round 1's kernel ran at the SM's ATOMG issue rate
// (139 G atomics/s), not at HBM speed.  Here the atomic is taken on a CLAIM BITMAP
// instead -- one bit per slot, slots/8 bytes, 32 MB at 2^27: L2-resident -- and the
// tuple is then written with a plain store.  The claims of a CTA tile are
// write-combined in shared memory, the way rx_scatter combines its output runs:
//   * the tile's tuples set their bits in a shared-memory image of the band of the
//     bitmap the tile touches (home slots of a sorted / locally shuffled tile lie
//     within tile + 2W slots of each other); a bit that is already set there is a
//     duplicate request inside the tile and loses at once;
//   * ONE global atomicOr per non-empty 32-slot word claims everything the tile
//     wants in it -- 65 atomics for 2048 sorted tuples instead of 2048 -- and its
//     return value tells every requester whether its bit was free;
//   * a tuple that lost sees the word's whole occupancy and goes straight to the
//     next free slot of its window instead of stepping one round trip at a time;
//   * tuples whose window lies outside the band (fully shuffled inputs) claim their
//     bit with a global atomic of their own, all of a thread's in flight together;
//   * winners store their tuple: the stores of a tile to neighbouring slots meet
//     in L2 and leave as full sectors, so the table is written without ever being
//     read -- 16 B/tuple of DRAM traffic.
// Slot arithmetic is 32-bit: the table has at most 2^32 slots (the reference's sizes
// are uint32_t, AtomicHashBuild.hpp:15,21).
// Measured alternatives (profiles/r02_sweep_global_n27_v*.jsonl, |R| = 2^27 sorted / local_shuffle 2^8,
// build in us): CAS on the table (round 1) 964 / 1392; warp-aggregated claims with per-run member masks
// 1119 / 3083 (REDUX/SHFL with 32 different masks per instruction serialise); probing resolved in shared
// memory with a commit per word 1238 / 998; this kernel: see DESIGN.md.
#pragma once
#include "hj_common.cuh"

namespace hj {

constexpr int kGtThreads = 256;
constexpr int kGtBandWords = 1024;                 // shared-memory band of the claim bitmap: 32768 slots
constexpr u32 kGtBand = 32u * kGtBandWords;

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

// A tuple is "in band" when its whole window [home, home + L + 32) lies inside the tile's band and
// below the table end (no wrap-around): then every attempt it ever makes goes through shared memory
// and its position is one 32-bit offset.  Everything else ("far": fully shuffled inputs, the last
// few slots of the table) claims with global atomics of its own.
constexpr u32 kGtFar = 0xffffffffu;

// next attempt of a tuple that found bit b of its word taken: the first slot above b that `view` shows
// free, or the first slot of the next word.  Everything skipped is occupied (or asked for by this tile,
// hence occupied when the round ends), so the probe's walk from the home slot never meets a gap.
__device__ __forceinline__ u32 gt_advance(u32 view, u32 b) {
  const u32 freeb = b == 31u ? 0u : (~view & (0xffffffffu << (b + 1u)));
  return freeb ? (u32)__ffs(freeb) - 1u - b : 32u - b;
}

// IPT tuples per thread ("transaction size": how many inserts a thread has in
// flight before it looks at any outcome), loaded as IPT coalesced 8-byte loads.
// hmask = (u32)(key_mask & table mask): home slot = (u32)tuple & hmask.
// STATS: the sampling round of the adaptive switch: cas_retries = claims that found their slot taken
// (the HTM-abort stand-in), probe_steps = slots stepped over, span of slots per CTA chunk, sectors per warp.
template <int IPT, bool STATS>
__global__ void __launch_bounds__(kGtThreads, 4)
gt_build_kernel(const u64* __restrict__ R, SegMap sm, u64* __restrict__ table, u32* __restrict__ claim, u32 mask,
                u32 hmask, u32 L, u64* __restrict__ ovf_list, u64* __restrict__ ctr) {
  constexpr int TILE = kGtThreads * IPT;
  __shared__ __align__(16) u32 s_req[kGtBandWords * 8];  // one byte per slot of the band: asked for this round
  __shared__ u32 s_want[kGtBandWords];  // the same, packed: bits the tile asks for this round
  __shared__ u32 s_old[kGtBandWords];   // what the global word held before the tile's atomicOr
  __shared__ u32 s_min[kGtThreads / 32], s_max[kGtThreads / 32];
  const u32 tid = threadIdx.x, lane = lane_id();
  const u64 chunks_per_seg = (sm.seg_len + TILE - 1) / TILE;
  const u64 nchunks = chunks_per_seg * sm.nslices;
  u32 retries = 0, steps = 0, zeros = 0;
  u64 sectors = 0, sampled = 0, span = 0, chunks = 0;
  for (u32 i = tid; i < kGtBandWords * 8; i += kGtThreads) s_req[i] = 0u;  // (every commit scan re-zeroes what it read)
  __syncthreads();

  for (u64 c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const u64 slice = c / chunks_per_seg;
    const u64 w = c - slice * chunks_per_seg;
    const u64 seg_begin = slice * sm.slice_stride + sm.slice_off;
    u64 seg_end = seg_begin + sm.seg_len;
    if (seg_end > sm.n) seg_end = sm.n;
    const u64 tbase = seg_begin + w * TILE;

    u64 v[IPT];
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      const u64 i = tbase + (u64)j * kGtThreads + tid;
      v[j] = i < seg_end ? ld_stream(R + i) : 0ull;
    }

    // the band: 32768 slots starting at the tile's lowest home slot (rounded down to a bitmap word)
    u32 lo = 0xffffffffu, hi = 0u;
#pragma unroll
    for (int j = 0; j < IPT; j++)
      if (v[j]) {
        const u32 hm = (u32)v[j] & hmask;
        lo = min(lo, hm);
        hi = max(hi, hm);
      }
    lo = __reduce_min_sync(0xffffffffu, lo);
    hi = __reduce_max_sync(0xffffffffu, hi);
    if (lane == 0) { s_min[tid >> 5] = lo; s_max[tid >> 5] = hi; }
    __syncthreads();
#pragma unroll
    for (int w2 = 0; w2 < kGtThreads / 32; w2++) { lo = min(lo, s_min[w2]); hi = max(hi, s_max[w2]); }
    if (STATS && tid == 0 && hi >= lo) {  // working-set signal: how wide is the band of slots one CTA chunk touches?
      span += (u64)(hi - lo) + 1;
      chunks += 1;
    }
    const u32 base = lo & ~31u;  // (an empty tile has lo = ~0, hi = 0: nothing is pending, no round runs)
    u32* const gband = claim + (base >> 5);
    u32 nwords = 0;  // words of the band this tile can touch
    if (hi >= lo) {
      const u64 nw = (((u64)hi - base + L + 64) >> 5) + 1;
      nwords = nw < (u64)kGtBandWords ? (u32)nw : (u32)kGtBandWords;
    }

    u32 rel[IPT];  // home - base, or kGtFar
    u32 att[IPT];  // offset from home of the next slot to try
    u32 pend = 0;  // bit j: tuple j still has to be placed
    u32 farbits = 0, failed_bits = 0;
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      att[j] = 0;
      rel[j] = kGtFar;
      if (v[j]) {
        pend |= 1u << j;
        const u32 hm = (u32)v[j] & hmask;
        const u32 r = hm - base;
        if ((u64)r + L + 32 < kGtBand && (u64)hm + L + 32 <= mask) rel[j] = r;
        else farbits |= 1u << j;
      } else if (tbase + (u64)j * kGtThreads + tid < seg_end) {
        ++zeros;  // a tuple equal to 0, the empty-slot sentinel: not inserted, reported (hj_result.zero_keys)
      }
    }

    // Rounds: every pending tuple of the tile makes its next attempt.  AtomicHashBuild.hpp:46-60
    // with the claim bitmap standing in for the slot's old value.
    while (__syncthreads_or(pend != 0)) {  // (also: everybody is done with s_want / s_old / s_min)
      u32 far_old[IPT];  // far tuples: what their own atomicOr returned
      u32 pre = 0;       // bit j: somebody in the tile asked for the same slot first
      if (pend & farbits) {  // all of a thread's far claims in flight together
#pragma unroll
        for (int j = 0; j < IPT; j++)
          if (pend & farbits & (1u << j)) {
            const u32 slot = (((u32)v[j] & hmask) + att[j]) & mask;
            far_old[j] = atomicOr(claim + (slot >> 5), 1u << (slot & 31u));
          }
      }
      // One BYTE per slot in shared memory: the 32 lanes of a sorted row ask for 32 consecutive slots = 8
      // different words (4 lanes each) instead of 32 bits of ONE word -- a 32-fold serialised ATOMS otherwise
      // (ncu, bitmap image: 16 M bank conflicts in 20 M shared wavefronts, the pipe at 62 %).
#pragma unroll
      for (int j = 0; j < IPT; j++)
        if ((pend & ~farbits) >> j & 1u) {
          const u32 cur = rel[j] + att[j];
          const u32 sh = (cur & 3u) * 8u;
          if (atomicOr(&s_req[cur >> 2], 1u << sh) >> sh & 1u) pre |= 1u << j;  // somebody in the tile asked first
        }
      __syncthreads();
      // one global atomic per non-empty word of the band: 32 request bytes -> 32 bits (a byte is 0 or 1:
      // x * 0x01020408 collects the four of a word in bits 24..27), and the bytes are zeroed for the next round
      for (u32 wi = tid; wi < nwords; wi += kGtThreads) {
        uint4* const rq = reinterpret_cast<uint4*>(s_req) + 2u * wi;
        const uint4 a = rq[0], b = rq[1];
        const u32 want = (a.x * 0x01020408u) >> 24 | ((a.y * 0x01020408u) >> 24) << 4 | ((a.z * 0x01020408u) >> 24) << 8 |
                         ((a.w * 0x01020408u) >> 24) << 12 | ((b.x * 0x01020408u) >> 24) << 16 |
                         ((b.y * 0x01020408u) >> 24) << 20 | ((b.z * 0x01020408u) >> 24) << 24 |
                         ((b.w * 0x01020408u) >> 24) << 28;
        s_want[wi] = want;
        if (want) {
          s_old[wi] = atomicOr(gband + wi, want);
          rq[0] = make_uint4(0u, 0u, 0u, 0u);
          rq[1] = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      __syncthreads();
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        if (pend >> j & 1u) {
          u32 b, view, old;
          if (farbits >> j & 1u) {
            b = ((((u32)v[j] & hmask) + att[j]) & mask) & 31u;
            old = far_old[j];
            view = old | (1u << b);
          } else {
            const u32 cur = rel[j] + att[j];
            b = cur & 31u;
            old = s_old[cur >> 5];
            view = old | s_want[cur >> 5];
          }
          if (!(pre >> j & 1u) && !(old >> b & 1u)) {
            table[(((u32)v[j] & hmask) + att[j]) & mask] = v[j];
            pend &= ~(1u << j);
          } else {
            ++steps;
            if (STATS && !(pre >> j & 1u)) ++retries;
            att[j] += gt_advance(view, b);
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
        const u32 sec = v[j] ? ((u32)v[j] & hmask) >> 2 : 0xffffffffu - lane;
        const u32 peers = __match_any_sync(0xffffffffu, sec);
        const bool leader = v[j] && (__ffs(peers) - 1 == (int)lane);
        const u32 leaders = __ballot_sync(0xffffffffu, leader);
        const u32 valid = __ballot_sync(0xffffffffu, v[j] != 0ull);
        if (lane == 0) {
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
        if ((int)lane >= o) incl += y;
      }
      u64 base2 = 0;
      if (lane == 31) base2 = atomicAdd(ctr + C_CONFLICTS, (u64)incl);
      base2 = __shfl_sync(0xffffffffu, base2, 31);
      u64 dst = base2 + (incl - mine);
#pragma unroll
      for (int j = 0; j < IPT; j++)
        if (failed_bits >> j & 1u) ovf_list[dst++] = v[j];
    }
  }

  if (STATS) {
    u64 s[5] = {retries, steps, sectors, sampled, zeros};
    u64* const d[5] = {ctr + C_CAS_RETRIES, ctr + C_PROBE_STEPS, ctr + C_SECTORS, ctr + C_SAMPLED, ctr + C_ZERO_KEYS};
    block_sum_atomic<5, kGtThreads>(s, d);
    if (tid == 0 && chunks) {
      atomicAdd(ctr + C_SPAN, span);
      atomicAdd(ctr + C_CHUNKS, chunks);
    }
  } else {
    u64 s[2] = {steps, zeros};
    u64* const d[2] = {ctr + C_PROBE_STEPS, ctr + C_ZERO_KEYS};
    block_sum_atomic<2, kGtThreads>(s, d);
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
  u32 steps = 0, zeros = 0;
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
      else if (base + (u64)j * kGtThreads + tid < seg_end) ++zeros;
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
  u64 s[2] = {steps, zeros};
  u64* const d[2] = {ctr + C_PROBE_STEPS, ctr + C_ZERO_KEYS};
  block_sum_atomic<2, kGtThreads>(s, d);
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
    // rounds: kOvfIPT independent probes in flight per thread (load <= 0.5: terminates).  Look before
    // claiming: a plain load costs a fraction of an atomic, and duplicate keys (same hash) would otherwise
    // queue their CAS on one slot after the other.
#pragma unroll 1
    while (pend) {
      u64 res[kOvfIPT];
#pragma unroll
      for (int j = 0; j < kOvfIPT; j++)
        if (pend >> j & 1u) res[j] = ld_relaxed(ovf_table + slot[j]);
#pragma unroll
      for (int j = 0; j < kOvfIPT; j++)
        if ((pend >> j & 1u) && res[j] == 0ull) res[j] = atomicCAS(ovf_table + slot[j], 0ull, v[j]);
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
// A warp reads IPT rows of 32 consecutive S tuples (row j of warp w = tuples (w*IPT + j)*32 + lane of
// the tile), so that on a sorted S -- the only probe side the reference has, main.cpp:93 -- the home
// slots of a warp are IPT*32 consecutive table slots.
//
// Three paths, chosen from the build's own device-side counters and the warp's keys (no host round trip):
//   * home-only: the build displaced nothing (probe_steps == 0, conflicts == 0), so every R tuple
//     sits in its home slot and a key can only match there -- ONE coalesced 8-byte read per S
//     tuple: the 16 B/tuple of SURVEY 8d exactly.  Every unique-key input (sorted, shuffle,
//     local_shuffle) takes this path.
//   * dense (probeLength <= 4, the warp's homes are consecutive slots): slot home+k of lane l is the
//     home slot of lane l+k -- the window is assembled with shuffles from ONE coalesced read per row
//     (plus three slots past the last row) instead of probeLength reads per tuple.
//   * general: the window is read as whole 32-byte sectors (one LDG.256 for the sector of the home
//     slot, the next one only if the walk runs over its end).
// Windows without an empty slot whose home has overflowed tuples (ovfmap) then consult the overflow table.
__device__ __forceinline__ void ld_sector(const u64* p, u64 (&c)[4]) {
  asm volatile("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];"
               : "=l"(c[0]), "=l"(c[1]), "=l"(c[2]), "=l"(c[3]) : "l"(p));
}

// home-only path: launched in front of gt_probe_kernel, exactly one of the two does the work
template <int IPT>
__global__ void __launch_bounds__(kGtThreads)
gt_probe_home_kernel(const u64* __restrict__ S, u64 m, const u64* __restrict__ table, u64 mask, u64 key_mask,
                     u64* __restrict__ ctr) {
  if (ctr[C_CONFLICTS] != 0 || ctr[C_PROBE_STEPS] != 0) return;
  constexpr int TILE = kGtThreads * IPT;
  const u32 tid = threadIdx.x;
  u64 checksum = 0;
  u32 wmatches = 0, zeros = 0;
  const u64 nchunks = (m + TILE - 1) / TILE;
  for (u64 c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const u64 base = c * TILE + tid;
    u64 s[IPT], cur[IPT];
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      const u64 i = base + (u64)j * kGtThreads;
      s[j] = i < m ? ld_stream(S + i) : 0ull;
    }
    // Two batches of IPT loads in flight, S then the home slots.  ptxas otherwise interleaves them pairwise to save
    // registers (2-3 loads in flight: 337 -> 430 us at 2^27); a data dependency it cannot see through makes every
    // table address depend on ALL S loads and every compare on ALL table loads (`none` is 0 at run time: the table
    // has at most 2^32 slots).  The slot of key 0 is slot 0, so the table load needs no predicate, and a padding
    // lane (s = 0) can never match: a stored tuple has a non-zero key.
    const u64 none = 0ull - (mask >> 63);
    u64 dep = 0;
#pragma unroll
    for (int j = 0; j < IPT; j++) dep |= s[j];
    dep &= none;
#pragma unroll
    for (int j = 0; j < IPT; j++) cur[j] = __ldg(table + (((s[j] & key_mask) & mask) | dep));
    dep = 0;
#pragma unroll
    for (int j = 0; j < IPT; j++) dep |= cur[j];
    dep &= none;
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      const u64 t = cur[j] | dep;
      if (t != 0ull && (t & key_mask) == (s[j] & key_mask)) { ++wmatches; checksum += t + s[j]; }
      zeros += s[j] == 0ull;  // padding lanes included: taken out again below
    }
  }
  if (blockIdx.x == 0 && tid == 0) zeros -= (u32)(nchunks * TILE - m);  // (mod 2^32: the block sum below is mod 2^64)
  u64 v[4] = {(u64)wmatches, checksum, (u64)wmatches, (u64)(long long)(int)zeros};  // (sign-extended: may be negative here)
  u64* const d[4] = {ctr + C_MATCHES, ctr + C_CHECKSUM, ctr + C_WINDOW_MATCHES, ctr + C_ZERO_KEYS};
  block_sum_atomic<4, kGtThreads>(v, d);
}

template <int IPT>
__global__ void __launch_bounds__(kGtThreads, 3)
gt_probe_kernel(const u64* __restrict__ S, u64 m, const u64* __restrict__ table, u64 mask, u64 key_mask, u32 L,
                const u64* __restrict__ ovf_table, u64 ovf_capacity, const u32* __restrict__ ovfmap,
                u64* __restrict__ ctr) {
  constexpr int TILE = kGtThreads * IPT;
  const u32 tid = threadIdx.x, lane = lane_id(), warp = tid >> 5;
  const u64 ovf_count = ctr[C_CONFLICTS];
  if (ovf_count == 0 && ctr[C_PROBE_STEPS] == 0) return;  // gt_probe_home_kernel took the probe
  u64 omask = 0;
  if (ovf_count) {
    u64 slots = ovf_slots_for(ovf_count);
    if (slots > ovf_capacity) slots = ovf_capacity;
    omask = slots - 1;
  }
  u64 checksum = 0;
  u32 wmatches = 0, omatches = 0, zeros = 0;  // per thread: < 2^32
  const u64 nchunks = (m + TILE - 1) / TILE;
  __shared__ u64 s_q[TILE];  // S tuples of the tile that have to consult the overflow table
  __shared__ u32 s_qn;
  if (tid == 0) s_qn = 0;
  __syncthreads();

  for (u64 c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const u64 base = c * TILE + (u64)warp * (IPT * 32) + lane;
    u64 s[IPT];
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      const u64 i = base + (u64)j * 32;
      s[j] = i < m ? ld_stream(S + i) : 0ull;
    }
#pragma unroll
    for (int j = 0; j < IPT; j++)
      if (s[j] == 0ull && base + (u64)j * 32 < m) ++zeros;
    u32 full = 0;  // bit j: window exhausted without an empty slot
    // dense: homes of the warp = h0 .. h0 + IPT*32 - 1, no wrap-around, every tuple valid
    const u64 h0 = __shfl_sync(0xffffffffu, (s[0] & key_mask) & mask, 0);
    bool dense = L <= 4 && h0 + IPT * 32 + 4 <= mask;
#pragma unroll
    for (int j = 0; j < IPT; j++) dense = dense && s[j] != 0ull && ((s[j] & key_mask) & mask) == h0 + j * 32 + lane;
    if (__all_sync(0xffffffffu, dense)) {
      u64 w0[IPT + 1];
#pragma unroll
      for (int j = 0; j < IPT; j++) w0[j] = __ldg(table + h0 + j * 32 + lane);
      w0[IPT] = lane < 3 ? __ldg(table + h0 + IPT * 32 + lane) : 0ull;  // the slots past the last row
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        const u64 key = s[j] & key_mask;
        bool walking = true;
#pragma unroll
        for (u32 k = 0; k < 4; k++) {
          u64 cw = w0[j];
          if (k) {  // slot home+k: lane+k's word of this row, or lane+k-32's of the next
            const u64 a = __shfl_sync(0xffffffffu, w0[j], (lane + k) & 31u);
            const u64 b = __shfl_sync(0xffffffffu, w0[j + 1], (lane + k) & 31u);
            cw = lane + k < 32u ? a : b;
          }
          if (k < L && walking) {
            if (cw == 0ull) walking = false;
            else if ((cw & key_mask) == key) { ++wmatches; checksum += cw + s[j]; }
          }
        }
        if (walking) full |= 1u << j;
      }
    } else {
      // general path, in rounds of one sector per tuple still walking
      u32 pend = 0;
      u32 off[IPT];  // slots of the window already walked
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
    }
    // the window is full: tuples of this home slot may have overflowed.  Only ~15 % of the tuples of the
    // reference's "uniform" input get here, and their chains differ in length: walked in place, a warp would
    // run every round with 3-5 lanes busy (ncu: 54 % of the kernel's instructions at 2.5-5 threads each).
    // The CTA queues them in shared memory instead and walks the queue with all lanes.
    if (ovf_count) {  // (kernel-uniform)
      u32 act = 0;
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        if (full >> j & 1u) {
          const u64 home = (s[j] & key_mask) & mask;
          if (__ldg(ovfmap + (home >> 5)) >> ((u32)home & 31u) & 1u) act |= 1u << j;
        }
      }
      if (act) {
        u32 pos = atomicAdd(&s_qn, (u32)__popc(act));
#pragma unroll
        for (int j = 0; j < IPT; j++)
          if (act >> j & 1u) s_q[pos++] = s[j];
      }
      __syncthreads();
      const u32 qn = s_qn;
      for (u32 i = tid; i < qn; i += kGtThreads) {
        const u64 sv = s_q[i];
        const u64 key = sv & key_mask;
        u64 os = ovf_hash(key, omask);
#pragma unroll 1
        while (true) {
          const u64 o = __ldg(ovf_table + os);
          if (o == 0ull) break;
          if ((o & key_mask) == key) { ++omatches; checksum += o + sv; }
          os = (os + 1) & omask;
        }
      }
      __syncthreads();
      if (tid == 0) s_qn = 0;
      __syncthreads();
    }
  }
  u64 v[4] = {(u64)wmatches + omatches, checksum, (u64)wmatches, (u64)zeros};
  u64* const d[4] = {ctr + C_MATCHES, ctr + C_CHECKSUM, ctr + C_WINDOW_MATCHES, ctr + C_ZERO_KEYS};
  block_sum_atomic<4, kGtThreads>(v, d);
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
