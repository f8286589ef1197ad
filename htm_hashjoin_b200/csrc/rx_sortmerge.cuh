// rx_sortmerge.cuh -- partitioned sort-merge join: the GPU counterpart of the reference's
// comparator SortMerge (SortMerge.cpp:8-39: tim_sort the build side, then a two-pointer merge
// that counts `left == right`, :23-33).  The reference sorts 64 slices of R only and relies on its
// probe side being generate_data("sorted") (main.cpp:93); here both sides are sorted, so any S works
// and duplicates on either side count as in the true join (sum_k cntR(k) * cntS(k)).
//
// Both relations arrive radix-partitioned on the same key bits (rx_kernels.cuh, either layout), so
// the join is per partition pair and no global order is needed (the m-way / MPSM arrangement):
//   one CTA per partition p
//     A = R_p sorted by key, kept in shared memory (chunks of capA if R_p is larger)
//     for every 4096-tuple tile of S_p: B = the tile sorted by key
//       merge: thread t owns 8 consecutive elements of B, finds its start in A with ONE binary
//       search (the merge-path split) and then walks A and its B elements with two pointers;
//       equal runs in A are rescanned for equal keys in B (mark / restore of a merge join).
// Tiles of up to 4096 tuples are sorted in REGISTERS (sm_sort_regs): 512 threads x 8 elements, element
// i = tid * 8 + e.  Of the 78 compare-exchange steps of a 4096-element bitonic network, those with partner
// distance j < 8 stay inside a thread, 8 <= j < 256 are warp shuffles, and only the 10 steps with j >= 256 cross
// warps through a 32 KB shared-memory exchange buffer (conflict-free e * 512 + tid layout).  The merge then
// runs straight from the registers: a thread's 8 elements ARE its 8 consecutive elements of B.  R chunks of more
// than 4096 tuples use the shared-memory network (sm_bitonic).
// Integer only; results are order-independent sums, so they are bit-exact against the oracle.
#pragma once
#include "hj_common.cuh"

namespace hj {

constexpr int kSmThreads = 512;

// strict order on (key, tuple): the padding value ~0 is a maximum, so real tuples sort in front of it
__device__ __forceinline__ bool sm_less(u64 x, u64 y, u64 kmask) {
  const u64 kx = x & kmask, ky = y & kmask;
  return kx < ky || (kx == ky && x < y);
}

// bitonic sort of X[0, L), L a power of two, ascending by sm_less; all threads of the CTA call it
__device__ __forceinline__ void sm_bitonic(u64* X, u32 L, u64 kmask) {
  const u32 tid = threadIdx.x;
  for (u32 k = 2; k <= L; k <<= 1) {
    for (u32 j = k >> 1; j > 0; j >>= 1) {
      // pair index q -> the lower element of the q-th compare-exchange of this step
      for (u32 q = tid; q < (L >> 1); q += kSmThreads) {
        const u32 i = ((q & ~(j - 1)) << 1) | (q & (j - 1));
        const u32 ixj = i | j;
        const u64 a = X[i], b = X[ixj];
        const bool up = (i & k) == 0;
        const bool sw = up ? sm_less(b, a, kmask) : sm_less(a, b, kmask);
        if (sw) { X[i] = b; X[ixj] = a; }
      }
      __syncthreads();
    }
  }
}

__device__ __forceinline__ u32 sm_pow2(u32 v) {
  u32 p = 2;
  while (p < v) p <<= 1;
  return p;
}


constexpr int kSmIPT = 8;                          // elements per thread of the register sort
constexpr u32 kSmTile = kSmThreads * kSmIPT;       // 4096

// The register sort works on ENCODED tuples so that, for the two usual key masks, the order (key, tuple) is the
// plain order of 64-bit words and a compare-exchange costs a 64-bit compare and two selects instead of the ~20
// instructions of sm_less (ncu of the first version: 1750 thread instructions per tuple, math-pipe bound):
//   MODE 1  key_mask = 0xffffffff (mc tuples: key low, payload high): halves swapped -- a register rename
//   MODE 2  key_mask = all 64 bits: the word itself
//   MODE 0  any other mask: the word itself, ordered with sm_less
// The padding value ~0 encodes to itself and stays the maximum.
template <int MODE>
__device__ __forceinline__ u64 sm_enc(u64 x) { return MODE == 1 ? (x << 32) | (x >> 32) : x; }
template <int MODE>
__device__ __forceinline__ u64 sm_dec(u64 w) { return MODE == 1 ? (w << 32) | (w >> 32) : w; }
template <int MODE>
__device__ __forceinline__ u64 sm_key(u64 w, u64 kmask) { return MODE == 1 ? w >> 32 : MODE == 2 ? w : w & kmask; }
template <int MODE>
__device__ __forceinline__ bool sm_lt(u64 a, u64 b, u64 kmask) { return MODE == 0 ? sm_less(a, b, kmask) : a < b; }

// compare-exchange of a (lower index) and b; up: ascending.  (b < a) == up also swaps equal words when
// descending, which changes nothing.
template <int MODE>
__device__ __forceinline__ void sm_cmpx(u64& a, u64& b, bool up, u64 kmask) {
  const bool sw = sm_lt<MODE>(b, a, kmask) == up;
  const u64 t = a;
  a = sw ? b : a;
  b = sw ? t : b;
}
// this thread keeps the smaller (keep_min) or larger of its v and the partner's o: one compare, one select
template <int MODE>
__device__ __forceinline__ u64 sm_keep(u64 v, u64 o, bool keep_min, u64 kmask) {
  return (sm_lt<MODE>(o, v, kmask) == keep_min) ? o : v;
}

// Stage k = 2^LK of the bitonic network over the 4096-slot sequence whose element tid * 8 + e sits in v[e] of
// thread tid.  Fully unrolled: every partner distance j is a compile-time constant, so each step is straight-line
// code on fixed registers (the first version looped over k and j at run time: 24 % of its instructions were
// register moves).  j < 8: inside the thread; 8 <= j < 256: warp shuffle; j >= 256: through X.
template <int MODE, int LK>
__device__ __forceinline__ void sm_stage(u64 (&v)[kSmIPT], u64* X, u64 kmask) {
  constexpr u32 k = 1u << LK;
  const u32 tid = threadIdx.x;
  const bool up = ((tid * kSmIPT) & k) == 0;  // k >= 8: the same for the thread's 8 elements
#pragma unroll
  for (int lj = LK - 1; lj >= 0; lj--) {
    const u32 j = 1u << lj;
    if (j >= 8 * 32) {
      const u32 tm = j >> 3, pt = tid ^ tm;
      const bool keep_min = ((tid & tm) == 0) == up;
      __syncthreads();  // earlier readers of X are done
#pragma unroll
      for (int e = 0; e < kSmIPT; e++) X[e * kSmThreads + tid] = v[e];
      __syncthreads();
#pragma unroll
      for (int e = 0; e < kSmIPT; e++) v[e] = sm_keep<MODE>(v[e], X[e * kSmThreads + pt], keep_min, kmask);
    } else if (j >= 8) {
      const u32 lm = j >> 3;
      const bool keep_min = ((tid & lm) == 0) == up;
#pragma unroll
      for (int e = 0; e < kSmIPT; e++)
        v[e] = sm_keep<MODE>(v[e], __shfl_xor_sync(0xffffffffu, v[e], lm), keep_min, kmask);
    } else {
      // direction of element e: (e & k) == 0 for k < 8
#pragma unroll
      for (int e = 0; e < kSmIPT; e++)
        if ((e & j) == 0) sm_cmpx<MODE>(v[e], v[e | j], k >= 8 ? up : (e & k) == 0, kmask);
    }
  }
}

// Bitonic sort, ascending, in blocks of L (a power of two <= 4096: only stages k <= L run; slots >= L must hold
// the padding value).  X: kSmTile u64 of shared memory for the cross-warp steps.  All threads of the CTA call it.
template <int MODE>
__device__ __forceinline__ void sm_sort_regs(u64 (&v)[kSmIPT], u64* X, u32 L, u64 kmask) {
  sm_stage<MODE, 1>(v, X, kmask);
  if (L < 4) return;
  sm_stage<MODE, 2>(v, X, kmask);
  if (L < 8) return;
  sm_stage<MODE, 3>(v, X, kmask);
  if (L < 16) return;
  sm_stage<MODE, 4>(v, X, kmask);
  if (L < 32) return;
  sm_stage<MODE, 5>(v, X, kmask);
  if (L < 64) return;
  sm_stage<MODE, 6>(v, X, kmask);
  if (L < 128) return;
  sm_stage<MODE, 7>(v, X, kmask);
  if (L < 256) return;
  sm_stage<MODE, 8>(v, X, kmask);
  if (L < 512) return;
  sm_stage<MODE, 9>(v, X, kmask);
  if (L < 1024) return;
  sm_stage<MODE, 10>(v, X, kmask);
  if (L < 2048) return;
  sm_stage<MODE, 11>(v, X, kmask);
  if (L < 4096) return;
  sm_stage<MODE, 12>(v, X, kmask);
}

// The shared-memory-only version (HJ_SM_VARIANT=0; kept for the A/B record).  capA / tileB: powers of two, shared memory = (capA + tileB) * 8 bytes
__global__ void __launch_bounds__(kSmThreads)
rx_sortmerge_kernel(const u64* __restrict__ Rp, const u64* __restrict__ offR, const u64* __restrict__ Sp,
                    const u64* __restrict__ offS, u32 nparts, u32 capA, u32 tileB, u64 key_mask,
                    u64* __restrict__ ctr, u64 capR, u64 capS) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  u64* A = reinterpret_cast<u64*>(smem_raw);
  u64* B = A + capA;
  const u32 tid = threadIdx.x;
  const u64 kmask = key_mask ? key_mask : ~0ull;
  u64 matches = 0, checksum = 0, maxpart = 0, chunked = 0;

  for (u32 p = blockIdx.x; p < nparts; p += gridDim.x) {
    // exact layout: partition p = [off[p], off[p+1]); histogram-free layout (cap != 0): [p*cap, min(off[p],
    // (p+1)*cap)) with off = the scatter's final cursors (as rx_join_kernel)
    const u64 r0 = capR ? p * capR : offR[p];
    const u64 r1 = capR ? (offR[p] < r0 + capR ? offR[p] : r0 + capR) : offR[p + 1];
    const u64 s0 = capS ? p * capS : offS[p];
    const u64 s1 = capS ? (offS[p] < s0 + capS ? offS[p] : s0 + capS) : offS[p + 1];
    const u64 nr = r1 - r0;
    maxpart = nr > maxpart ? nr : maxpart;
    if (nr == 0 || s1 == s0) continue;
    if (nr > capA) chunked++;
    for (u64 c0 = r0; c0 < r1; c0 += capA) {
      const u32 na = (u32)(c0 + capA < r1 ? capA : r1 - c0);
      const u32 LA = sm_pow2(na);
      __syncthreads();  // the previous merge is done with A and B
      for (u32 i = tid; i < LA; i += kSmThreads) A[i] = i < na ? ld_stream(Rp + c0 + i) : ~0ull;
      __syncthreads();
      sm_bitonic(A, LA, kmask);
      for (u64 t0 = s0; t0 < s1; t0 += tileB) {
        const u32 nb = (u32)(t0 + tileB < s1 ? tileB : s1 - t0);
        const u32 LB = sm_pow2(nb);
        __syncthreads();  // the previous tile's merge is done with B
        for (u32 i = tid; i < LB; i += kSmThreads) B[i] = i < nb ? ld_stream(Sp + t0 + i) : ~0ull;
        __syncthreads();
        sm_bitonic(B, LB, kmask);
        // merge: V consecutive B elements per thread, one binary search for the start in A
        const u32 V = (nb + kSmThreads - 1) / kSmThreads;
        const u32 b0 = tid * V;
        if (b0 < nb) {
          const u32 b1 = b0 + V < nb ? b0 + V : nb;
          const u64 kfirst = B[b0] & kmask;
          u32 lo = 0, hi = na;
          while (lo < hi) {
            const u32 mid = (lo + hi) >> 1;
            if ((A[mid] & kmask) < kfirst) lo = mid + 1; else hi = mid;
          }
          u32 ai = lo;
          for (u32 q = b0; q < b1; q++) {
            const u64 s = B[q], ks = s & kmask;
            while (ai < na && (A[ai] & kmask) < ks) ai++;
            for (u32 j = ai; j < na; j++) {
              const u64 r = A[j];
              if ((r & kmask) != ks) break;
              matches++;
              checksum += r + s;
            }
          }
        }
      }
    }
  }
  u64 v[2] = {matches, checksum};
  u64* const d[2] = {ctr + C_MATCHES, ctr + C_CHECKSUM};
  block_sum_atomic<2, kSmThreads>(v, d);
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 16));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 8));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 4));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 2));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 1));
  if (tid == 0) {
    atomicMax(ctr + C_MAX_PART, maxpart);
    if (chunked) atomicAdd(ctr + C_TOO_BIG, chunked);
  }
}

// merge of this thread's 8 sorted elements (slots tid * 8 + e < nb of the tile) with A[0, na)
template <int MODE>
__device__ __forceinline__ void sm_merge_regs(const u64 (&v)[kSmIPT], u32 nb, const u64* A, u32 na, u64 kmask,
                                              u64& matches, u64& checksum) {
  const u32 b0 = threadIdx.x * kSmIPT;
  if (b0 >= nb) return;
  const u64 kfirst = sm_key<MODE>(v[0], kmask);
  u32 lo = 0, hi = na;
  while (lo < hi) {
    const u32 mid = (lo + hi) >> 1;
    if (sm_key<MODE>(A[mid], kmask) < kfirst) lo = mid + 1; else hi = mid;
  }
  u32 ai = lo;
#pragma unroll
  for (int e = 0; e < kSmIPT; e++) {
    if (b0 + e < nb) {
      const u64 ks = sm_key<MODE>(v[e], kmask);
      while (ai < na && sm_key<MODE>(A[ai], kmask) < ks) ai++;
      for (u32 j = ai; j < na; j++) {
        const u64 r = A[j];
        if (sm_key<MODE>(r, kmask) != ks) break;
        matches++;
        checksum += sm_dec<MODE>(r) + sm_dec<MODE>(v[e]);
      }
    }
  }
}

// The register-sort version (default).  capA: 4096 or 8192; shared memory = (capA + 4096) * 8 bytes.  A holds
// encoded tuples.  CTAS = CTAs per SM the register budget is set for: 3 needs capA = 4096 (64 KB per CTA), which
// the host picks when no R partition is larger.
template <int MODE, int CTAS>
__global__ void __launch_bounds__(kSmThreads, CTAS)
rx_sortmerge_regs_kernel(const u64* __restrict__ Rp, const u64* __restrict__ offR, const u64* __restrict__ Sp,
                         const u64* __restrict__ offS, u32 nparts, u32 capA, u64 key_mask, u64* __restrict__ ctr,
                         u64 capR, u64 capS) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  u64* A = reinterpret_cast<u64*>(smem_raw);
  u64* X = A + capA;  // exchange buffer of the register sort
  const u32 tid = threadIdx.x;
  const u64 kmask = key_mask ? key_mask : ~0ull;
  u64 matches = 0, checksum = 0, maxpart = 0, chunked = 0;
  u64 v[kSmIPT];

  for (u32 p = blockIdx.x; p < nparts; p += gridDim.x) {
    // exact layout: partition p = [off[p], off[p+1]); histogram-free layout (cap != 0): [p*cap, min(off[p],
    // (p+1)*cap)) with off = the scatter's final cursors (as rx_join_kernel)
    const u64 r0 = capR ? p * capR : offR[p];
    const u64 r1 = capR ? (offR[p] < r0 + capR ? offR[p] : r0 + capR) : offR[p + 1];
    const u64 s0 = capS ? p * capS : offS[p];
    const u64 s1 = capS ? (offS[p] < s0 + capS ? offS[p] : s0 + capS) : offS[p + 1];
    const u64 nr = r1 - r0;
    maxpart = nr > maxpart ? nr : maxpart;
    if (nr == 0 || s1 == s0) continue;
    if (nr > capA) chunked++;
    for (u64 c0 = r0; c0 < r1; c0 += capA) {
      const u32 na = (u32)(c0 + capA < r1 ? capA : r1 - c0);
      const u32 LA = sm_pow2(na);
      __syncthreads();  // the previous merge is done with A
      if (LA <= kSmTile) {
        // the input order is irrelevant to a sort: coalesced loads (slot tid * 8 + e <- tuple e * 512 + tid)
        // when the chunk fills all slots, else slot i <- tuple i so that the padding sits behind slot na
#pragma unroll
        for (int e = 0; e < kSmIPT; e++) {
          const u32 i = na == kSmTile ? e * kSmThreads + tid : tid * kSmIPT + e;
          v[e] = i < na ? sm_enc<MODE>(ld_stream(Rp + c0 + i)) : ~0ull;
        }
        sm_sort_regs<MODE>(v, X, LA, kmask);
#pragma unroll
        for (int e = 0; e < kSmIPT; e++)
          if (tid * kSmIPT + e < LA) A[tid * kSmIPT + e] = v[e];
        __syncthreads();
      } else {
        // more than 4096 tuples: the shared-memory network on plain tuples, encoded afterwards (order kept)
        for (u32 i = tid; i < LA; i += kSmThreads) A[i] = i < na ? ld_stream(Rp + c0 + i) : ~0ull;
        __syncthreads();
        sm_bitonic(A, LA, kmask);
        for (u32 i = tid; i < LA; i += kSmThreads) A[i] = sm_enc<MODE>(A[i]);
        __syncthreads();
      }
      for (u64 t0 = s0; t0 < s1; t0 += kSmTile) {
        const u32 nb = (u32)(t0 + kSmTile < s1 ? kSmTile : s1 - t0);
#pragma unroll
        for (int e = 0; e < kSmIPT; e++) {
          const u32 i = nb == kSmTile ? e * kSmThreads + tid : tid * kSmIPT + e;
          v[e] = i < nb ? sm_enc<MODE>(ld_stream(Sp + t0 + i)) : ~0ull;
        }
        sm_sort_regs<MODE>(v, X, sm_pow2(nb), kmask);
        sm_merge_regs<MODE>(v, nb, A, na, kmask, matches, checksum);
      }
    }
  }
  u64 vs[2] = {matches, checksum};
  u64* const d[2] = {ctr + C_MATCHES, ctr + C_CHECKSUM};
  block_sum_atomic<2, kSmThreads>(vs, d);
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 16));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 8));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 4));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 2));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 1));
  if (tid == 0) {
    atomicMax(ctr + C_MAX_PART, maxpart);
    if (chunked) atomicAdd(ctr + C_TOO_BIG, chunked);
  }
}

}  // namespace hj
