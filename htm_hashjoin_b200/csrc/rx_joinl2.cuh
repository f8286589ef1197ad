// rx_joinl2.cuh -- join of radix partitions that do NOT fit shared memory.
//
// BASELINE config 3 sweeps the fan-out from 2^7 to 2^14 at |R| = 2^27: below 2^13 an R
// partition (2^20 .. 2^15 tuples) is far beyond the 227 KB of one SM.  The reference joins
// any partition size with its cache-resident bucket chains (bucket_chaining_join,
// mc/src/parallel_radix_join.c:231-283: "cache-sized" there means the CPU's L2).  The B200
// counterpart of "cache-sized" is the 126 MB L2: partitions of up to a few MB are joined
// through an open-addressing table in GLOBAL memory that is laid out partition by
// partition, so that the slice of the table a wave of CTAs works on stays L2-resident:
//
//   The table is made of 4-slot buckets = one 32 B sector, the unit L2 and HBM move anyway:
//   window of partition p = buckets [B(p), B(p+1)), B(p) = offR[p]/2 + p, i.e. >= 2 slots per
//   R tuple (load <= 0.5) for ANY partition size -- the window comes from the partition's own
//   histogram count, so skew cannot overflow it.  bucket = B(p) + mulhi(hash(key >> radix_bits),
//   window size); a full bucket spills to the next one, wrapping inside the window.  A probe
//   is ONE 256-bit load (LDG.256) of the home bucket in 86 % of the cases; the first version
//   used single slots and needed ~5 dependent L2 round trips per thread batch (14 ms at C3).
//
// Both kernels walk the partitioned relation front to back, tiles handed out by a ticket
// counter: the tiles in flight are always the most recent ones, so all CTAs touch the windows
// of a few neighbouring partitions.  (A static grid-stride assignment lets CTAs drift apart:
// ncu showed every table sector fetched from HBM 4 times in the build.)  Inserts and probes
// run in rounds like the global table (gt_kernels.cuh): IPT independent L2 accesses in
// flight per thread and round.  Matches are counted for any key multiplicity (true
// equi-join); checksum = sum over pairs (r + s).
#pragma once
#include "hj_common.cuh"
#include "rx_kernels.cuh"

namespace hj {

constexpr int kL2Threads = 256;

__device__ __forceinline__ u32 l2_hash(u64 key, u32 radix_bits) {
  const u64 k = key >> radix_bits;
  return ((u32)k ^ (u32)(k >> 32)) * 0x9E3779B1u;
}

// window of the partition a tuple belongs to: first bucket and size (in buckets, >= 1)
__device__ __forceinline__ void l2_window(const u64* __restrict__ offR, const Digit& dg, u64 v, u64& base, u32& size) {
  const u32 p = digit_of(dg, v);
  const u64 b = (__ldg(offR + p) >> 1) + p, e = (__ldg(offR + p + 1) >> 1) + p + 1;
  base = b;
  size = (u32)(e - b);  // partitions are < 2^32 tuples
}
__device__ __forceinline__ void ld_bucket(const u64* p, u64 (&c)[4]) {
  asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0, %1, %2, %3}, [%4];"
               : "=l"(c[0]), "=l"(c[1]), "=l"(c[2]), "=l"(c[3]) : "l"(p));
}
// table slots needed for n tuples in nparts partitions
__host__ __device__ __forceinline__ u64 l2_table_slots(u64 n, u64 nparts) { return 4ull * ((n >> 1) + nparts + 1); }

template <int IPT>
__device__ __forceinline__ void l2_load_tile(const u64* __restrict__ in, u64 n, u64 base, u32 tid, u64 (&v)[IPT]) {
#pragma unroll
  for (int j = 0; j < IPT / 2; j++) {
    const u64 i = base + 2ull * (j * kL2Threads + tid);
    if (i + 1 < n) {
      const ulonglong2 t = ld_stream_v2(in + i);
      v[2 * j] = t.x;
      v[2 * j + 1] = t.y;
    } else {
      v[2 * j] = i < n ? ld_stream(in + i) : 0ull;
      v[2 * j + 1] = 0ull;
    }
  }
}

// Rp: R in partition order (n tuples, 16-byte aligned); offR: 2^bits + 1 partition offsets;
// dg: the digit function that produced the partition index; table: l2_table_slots(n, 2^bits)
// slots, zeroed, 32-byte aligned.
template <int IPT>
__global__ void __launch_bounds__(kL2Threads)
rx_l2_build_kernel(const u64* __restrict__ Rp, u64 n, const u64* __restrict__ offR, Digit dg, u32 radix_bits,
                   u64* __restrict__ table, u64* __restrict__ ticket) {
  constexpr int TILE = kL2Threads * IPT;
  const u32 tid = threadIdx.x;
  const u64 ntiles = (n + TILE - 1) / TILE;
  __shared__ u64 s_tile;
  while (true) {
    __syncthreads();
    if (tid == 0) s_tile = atomicAdd(ticket, 1ull);
    __syncthreads();
    const u64 t = s_tile;
    if (t >= ntiles) break;
    u64 v[IPT], wb[IPT];
    u32 ws[IPT], sl[IPT];
    l2_load_tile<IPT>(Rp, n, t * TILE, tid, v);
    u32 pend = 0;
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      wb[j] = 0; ws[j] = 1; sl[j] = 0;
      if (v[j]) {
        l2_window(offR, dg, v[j], wb[j], ws[j]);
        sl[j] = 4u * __umulhi(l2_hash(v[j] & dg.key_mask, radix_bits), ws[j]);  // slot index inside the window
        pend |= 1u << j;
      }
    }
    // slots of a bucket fill in order (slot i+1 is tried only after slot i was seen taken), so
    // "slot 3 taken" == "bucket full" for the probe
#pragma unroll 1
    while (pend) {
      u64 res[IPT];
#pragma unroll
      for (int j = 0; j < IPT; j++)
        if (pend >> j & 1u) res[j] = atomicCAS(table + 4ull * wb[j] + sl[j], 0ull, v[j]);
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        if (pend >> j & 1u) {
          if (res[j] == 0ull) pend &= ~(1u << j);
          else sl[j] = sl[j] + 1 == 4u * ws[j] ? 0u : sl[j] + 1;
        }
      }
    }
  }
}

// NARROW: every key fits 32 bits (ctr[C_KEY_HI] == 0, set by rx_hist) and the key is the low
// half of the tuple: one 32-bit compare per slot.  The two instantiations are launched back to
// back; the one whose precondition does not hold returns at once.
template <int IPT, bool NARROW>
__global__ void __launch_bounds__(kL2Threads, NARROW ? 4 : 3)
rx_l2_probe_kernel(const u64* __restrict__ Sp, u64 m, const u64* __restrict__ offR, Digit dg, u32 radix_bits,
                   const u64* __restrict__ table, u64* __restrict__ ctr, int narrow_possible) {
  if (NARROW != (narrow_possible && ctr[C_KEY_HI] == 0ull)) return;
  constexpr int TILE = kL2Threads * IPT;
  const u32 tid = threadIdx.x;
  const u64 ntiles = (m + TILE - 1) / TILE;
  u64 checksum = 0;
  u32 matches = 0;  // per thread: < 2^32 unless one key repeats ~2^16 times on both sides
  __shared__ u64 s_tile;
  while (true) {
    __syncthreads();
    if (tid == 0) s_tile = atomicAdd(ctr + C_TILE1, 1ull);
    __syncthreads();
    const u64 t = s_tile;
    if (t >= ntiles) break;
    u64 s[IPT], wb[IPT];
    u32 ws[IPT], sl[IPT];
    l2_load_tile<IPT>(Sp, m, t * TILE, tid, s);
    u32 pend = 0;
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      wb[j] = 0; ws[j] = 1; sl[j] = 0;
      if (s[j]) {
        l2_window(offR, dg, s[j], wb[j], ws[j]);
        sl[j] = __umulhi(l2_hash(s[j] & dg.key_mask, radix_bits), ws[j]);  // bucket index inside the window
        pend |= 1u << j;
      }
    }
#pragma unroll 1
    while (pend) {
      u64 cur[IPT][4];
#pragma unroll
      for (int j = 0; j < IPT; j++)
        if (pend >> j & 1u) ld_bucket(table + 4ull * (wb[j] + sl[j]), cur[j]);
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        if (pend >> j & 1u) {
          const u64 key = s[j] & dg.key_mask;
#pragma unroll
          for (int i = 0; i < 4; i++) {
            // NARROW: keys are non-zero in their low half, an empty slot is all zero
            const bool hit = NARROW ? (u32)cur[j][i] == (u32)s[j]
                                    : (cur[j][i] != 0ull && (cur[j][i] & dg.key_mask) == key);
            if (hit) { ++matches; checksum += cur[j][i] + s[j]; }
          }
          if (cur[j][3] == 0ull) pend &= ~(1u << j);  // bucket not full: nothing spilled further
          else sl[j] = sl[j] + 1 == ws[j] ? 0u : sl[j] + 1;
        }
      }
    }
  }
  u64 v[2] = {(u64)matches, checksum};
  u64* const d[2] = {ctr + C_MATCHES, ctr + C_CHECKSUM};
  block_sum_atomic<2, kL2Threads>(v, d);
}

}  // namespace hj
