// rx_sortmerge.cuh -- partitioned sort-merge join: the GPU counterpart of the reference's
// comparator SortMerge (SortMerge.cpp:8-39: tim_sort the build side, then a two-pointer merge
// that counts `left == right`, :23-33).  The reference sorts 64 slices of R only and relies on its
// probe side being generate_data("sorted") (main.cpp:93); here both sides are sorted, so any S works
// and duplicates on either side count as in the true join (sum_k cntR(k) * cntS(k)).
//
// Both relations arrive radix-partitioned on the same key bits (rx_kernels.cuh, exact layout), so
// the join is per partition pair and no global order is needed (the m-way / MPSM arrangement):
//   one CTA per partition p
//     A = R_p sorted by key in shared memory (bitonic network; chunks of capA if R_p is larger)
//     for every tile of S_p: B = tile sorted by key in shared memory
//       merge: thread t owns V consecutive elements of B, finds its start in A with ONE binary
//       search (the merge-path split) and then walks A and its B elements with two pointers;
//       equal runs in A are rescanned for equal keys in B (mark / restore of a merge join).
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

// capA / tileB: powers of two, shared memory = (capA + tileB) * 8 bytes
__global__ void __launch_bounds__(kSmThreads)
rx_sortmerge_kernel(const u64* __restrict__ Rp, const u64* __restrict__ offR, const u64* __restrict__ Sp,
                    const u64* __restrict__ offS, u32 nparts, u32 capA, u32 tileB, u64 key_mask,
                    u64* __restrict__ ctr) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  u64* A = reinterpret_cast<u64*>(smem_raw);
  u64* B = A + capA;
  const u32 tid = threadIdx.x;
  const u64 kmask = key_mask ? key_mask : ~0ull;
  u64 matches = 0, checksum = 0, maxpart = 0, chunked = 0;

  for (u32 p = blockIdx.x; p < nparts; p += gridDim.x) {
    const u64 r0 = offR[p], r1 = offR[p + 1], s0 = offS[p], s1 = offS[p + 1];
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

}  // namespace hj
