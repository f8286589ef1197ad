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
//   * a lost CAS moves on to the next slot (the reference burns an attempt and
//     stays, :50-54).
//   * the probe masks slot+1 (the reference does not, :78-82).
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

// IPT tuples per thread ("transaction size": how many inserts a thread has in
// flight before it looks at any outcome).  IPT == 1 is the scalar variant used
// for unaligned inputs.
template <int IPT, bool NOCC, bool STATS>
__global__ void __launch_bounds__(kGtThreads)
gt_build_kernel(const u64* __restrict__ R, SegMap sm, u64* __restrict__ table, u64 mask, u64 key_mask, u32 L,
                u64* __restrict__ ovf_list, u64* __restrict__ ctr) {
  constexpr int TILE = kGtThreads * IPT;
  constexpr bool PEEK = STATS;  // only a load-then-CAS can tell a lost race from an old occupant
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
    if (IPT == 1) {
      const u64 i = base + tid;
      v[0] = i < seg_end ? ld_stream(R + i) : 0ull;
    } else {
#pragma unroll
      for (int j = 0; j < IPT / 2; j++) {
        const u64 i = base + 2ull * (j * kGtThreads + tid);
        if (i + 1 < seg_end) {
          ulonglong2 t = ld_stream_v2(R + i);
          v[2 * j] = t.x;
          v[2 * j + 1] = t.y;
        } else {
          v[2 * j] = i < seg_end ? ld_stream(R + i) : 0ull;
          v[2 * j + 1] = 0ull;
        }
      }
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

    u32 att[IPT];  // attempts used so far == offset from home of the next slot to try
    u32 pend = 0;  // bit j: tuple j still has to be placed
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      att[j] = 0;
      if (v[j]) pend |= 1u << j;
    }
    u32 failed_bits = 0;
    bool ranked = NOCC;  // NoCC keeps the reference's unsynchronised behaviour (NoCCHashBuild.hpp:45-49)

    // Rounds: every pending tuple of the batch issues its next attempt, all in flight
    // together (IPT independent L2 atomics per thread), then the outcomes are examined.
    // A lost CAS moves on to the next slot (the reference burns the attempt in place, 9-5).
#pragma unroll 1
    while (__any_sync(0xffffffffu, pend != 0)) {  // warp-uniform: the ranking below is a warp collective
      u64 res[IPT];
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        if (pend >> j & 1u) {
          u64* p = table + ((((v[j] & key_mask) & mask) + att[j]) & mask);
          if (NOCC || PEEK) res[j] = ld_relaxed(p);
          else res[j] = atomicCAS(p, 0ull, v[j]);
        }
      }
      if (PEEK) {  // second stage: CAS where the peek saw an empty slot
#pragma unroll
        for (int j = 0; j < IPT; j++) {
          if ((pend >> j & 1u) && res[j] == 0ull) {
            res[j] = atomicCAS(table + ((((v[j] & key_mask) & mask) + att[j]) & mask), 0ull, v[j]);
            if (res[j] != 0ull) ++retries;
          }
        }
      }
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        if (pend >> j & 1u) {
          if (res[j] == 0ull) {
            if (NOCC) table[(((v[j] & key_mask) & mask) + att[j]) & mask] = v[j];
            pend &= ~(1u << j);
          } else {
            ++steps;
            ++att[j];
          }
        }
      }
      // Warp-cooperative restart after the first lost round: lanes of one load instruction
      // whose tuples share a home slot (duplicate keys sit next to each other in sorted /
      // locally shuffled inputs) would chase each other down the window one CAS round trip at
      // a time.  Rank the losers of a home slot (__match_any_sync) and let rank i continue at
      // home+1+i with i more attempts spent: the lower ranks each try home+1..home+i first,
      // so those slots are occupied when the build ends and the probe's walk from home never
      // meets a gap.  Inputs without duplicates never get here.
      if (!ranked && __any_sync(0xffffffffu, pend != 0)) {
        ranked = true;
#pragma unroll
        for (int j = 0; j < IPT; j++) {
          const bool lost = pend >> j & 1u;
          const u64 home = lost ? ((v[j] & key_mask) & mask) : (~0ull - lane_id());
          const u32 peers = __match_any_sync(0xffffffffu, home);
          if (lost) att[j] += __popc(peers & ((1u << lane_id()) - 1u));
        }
      }
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        if ((pend >> j & 1u) && att[j] >= L) {
          pend &= ~(1u << j);
          failed_bits |= 1u << j;
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
    // AtomicHashBuild.hpp:62-63: budget exhausted -> conflicts[].  One atomicAdd per warp and
    // tile claims the range (a claim per load instruction put ~n/32 atomics on one address:
    // the uniform build spent most of its 1.1 ms queueing there).
    if (__any_sync(0xffffffffu, failed_bits != 0)) {
      const u32 mine = __popc(failed_bits);
      u32 incl = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const u32 y = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)lane_id() >= o) incl += y;
      }
      u64 base = 0;
      if (lane_id() == 31) base = atomicAdd(ctr + C_CONFLICTS, (u64)incl);
      base = __shfl_sync(0xffffffffu, base, 31);
      u64 dst = base + (incl - mine);
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

// ---- overflow table ---------------------------------------------------------
// Sized from the device-side conflict count so no host round trip is needed
// between build and probe; both kernels return at once when nothing overflowed.
__global__ void __launch_bounds__(256) gt_ovf_clear_kernel(u64* __restrict__ ovf_table, const u64* __restrict__ ctr,
                                                            u64 capacity) {
  const u64 count = ctr[C_CONFLICTS];
  if (count == 0) return;
  u64 slots = ovf_slots_for(count);
  if (slots > capacity) slots = capacity;
  ulonglong2* t2 = reinterpret_cast<ulonglong2*>(ovf_table);
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < slots / 2; i += (u64)gridDim.x * blockDim.x)
    t2[i] = make_ulonglong2(0ull, 0ull);
}

__global__ void __launch_bounds__(256) gt_ovf_insert_kernel(const u64* __restrict__ ovf_list,
                                                             u64* __restrict__ ovf_table,
                                                             const u64* __restrict__ ctr, u64 key_mask, u64 capacity) {
  const u64 count = ctr[C_CONFLICTS];
  if (count == 0) return;
  u64 slots = ovf_slots_for(count);
  if (slots > capacity) slots = capacity;
  const u64 omask = slots - 1;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (u64)gridDim.x * blockDim.x) {
    const u64 v = ovf_list[i];
    u64 slot = ovf_hash(v & key_mask, omask);
    while (atomicCAS(ovf_table + slot, 0ull, v) != 0ull) slot = (slot + 1) & omask;  // load <= 0.5: terminates
  }
}

// Locality pre-check of the adaptive switch: the span signal of the sampling round WITHOUT the table
// (read-only, same chunk geometry as gt_build_kernel<2, false, true>).  If the slots of a CTA chunk
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
template <int IPT>
__global__ void __launch_bounds__(kGtThreads)
gt_probe_kernel(const u64* __restrict__ S, u64 m, const u64* __restrict__ table, u64 mask, u64 key_mask, u32 L,
                const u64* __restrict__ ovf_table, u64 ovf_capacity, u64* __restrict__ ctr) {
  constexpr int TILE = kGtThreads * IPT;
  const u32 tid = threadIdx.x;
  const u64 ovf_count = ctr[C_CONFLICTS];
  u64 omask = 0;
  if (ovf_count) {
    u64 slots = ovf_slots_for(ovf_count);
    if (slots > ovf_capacity) slots = ovf_capacity;
    omask = slots - 1;
  }
  u64 matches = 0, checksum = 0, wmatches = 0, omatches = 0;
  const u64 nchunks = (m + TILE - 1) / TILE;

  for (u64 c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const u64 base = c * TILE;
    u64 s[IPT];
    if (IPT == 1) {
      const u64 i = base + tid;
      s[0] = i < m ? ld_stream(S + i) : 0ull;
    } else {
#pragma unroll
      for (int j = 0; j < IPT / 2; j++) {
        const u64 i = base + 2ull * (j * kGtThreads + tid);
        if (i + 1 < m) {
          ulonglong2 t = ld_stream_v2(S + i);
          s[2 * j] = t.x;
          s[2 * j + 1] = t.y;
        } else {
          s[2 * j] = i < m ? ld_stream(S + i) : 0ull;
          s[2 * j + 1] = 0ull;
        }
      }
    }
    // Window walk in rounds: slot home+r of every tuple still walking, IPT independent loads
    // in flight per thread (a per-tuple walk is a chain of dependent L2 round trips).
    u32 pend = 0, full = 0;  // bit j: still walking / window exhausted without an empty slot
#pragma unroll
    for (int j = 0; j < IPT; j++)
      if (s[j]) pend |= 1u << j;
#pragma unroll 1
    for (u32 r = 0; r < L && pend; r++) {
      u64 cur[IPT];
#pragma unroll
      for (int j = 0; j < IPT; j++)
        if (pend >> j & 1u) cur[j] = __ldg(table + ((((s[j] & key_mask) & mask) + r) & mask));
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        if (pend >> j & 1u) {
          if (cur[j] == 0ull) {
            pend &= ~(1u << j);
          } else {
            if ((cur[j] & key_mask) == (s[j] & key_mask)) { ++wmatches; checksum += cur[j] + s[j]; }
            if (r + 1 >= L) { pend &= ~(1u << j); full |= 1u << j; }
          }
        }
      }
    }
    // the window is full: tuples of this home slot may have overflowed
    if (ovf_count && full) {
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        if (full >> j & 1u) {
          const u64 key = s[j] & key_mask;
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
  matches = wmatches + omatches;
  u64 v[3] = {matches, checksum, wmatches};
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
