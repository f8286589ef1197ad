// rx_join32.cuh -- per-partition shared-memory join, 32-bit-key fast path.
//
// Same job as rx_join_kernel (rx_kernels.cuh): the per-partition build + probe of
// bucket_chaining_join / histogram_join (mc/src/parallel_radix_join.c:231-283,
// :306-369), one CTA per final partition.  What is different is the table, laid
// out for the probe, which is 8/9 of the tuples at |S| = 8|R|:
//
//   * every in-scope tuple is {key: low 32 bits, payload: high 32 bits} (mc's
//     tuple_t, types.h:31-36) or a bare value < 2^32 (DataGen.hpp).  rx_hist ORs the
//     key bits above bit 31 of every tuple into ctr[C_KEY_HI]; if that is zero this
//     kernel runs, otherwise it returns at once and rx_join_kernel does the work.
//   * all tuples of partition p share their low radix bits, so the table stores
//     k = key >> radix_bits (32 bits, 0xffffffff = empty) next to the payload half.
//   * buckets of W keys (W = 2: LDS.64, W = 4: LDS.128) in one array, the payloads in a
//     second array at the same index (interleaving them inside the bucket was measured:
//     the key loads then only ever touch half of the banks and the join got 40 % slower).
//     ONE vector LDS fetches the bucket's keys, W compares resolve it, one LDS at a
//     compare-selected offset fetches the payload.  At load ~0.3 almost every key sits
//     in its home bucket, so the probe is straight-line code instead of 32 divergent
//     linear-probe walks (ncu r01_ncu_radix_v2: rx_join_kernel spends 65 instructions
//     per tuple, issue-bound at 62 %).
//   * shared memory is addressed with 32-bit shared-space addresses through ld.shared /
//     st.shared / atom.shared PTX: the first version of this kernel went through generic
//     pointers and the compiler rebuilt the shared window base (S2UR CgaCtaId, ULEA ...)
//     around every access -- 60 instructions per tuple (ncu j32_v0).
//   * the build notes whether any key occurred twice in the chunk.  Unique R_p (every
//     PK/FK join): the probe stops at the first match.  Otherwise it counts every match
//     and walks on while buckets are full -- the true equi-join either way.
//   * the bucket count need not be a power of two (index = mulhi(hash, nbuckets)), so
//     the table is sized to what shared memory gives at the chosen CTAs/SM.
//
// checksum = sum over pairs (r + s) = 2*sum(key) + ((sum(pay_r + pay_s)) << 32)  (mod 2^64)
#pragma once
#include "hj_common.cuh"

namespace hj {

constexpr u32 kJ32Empty = 0xffffffffu;  // k = key >> radix_bits < 2^31 for radix_bits >= 1
constexpr u32 kJ32Mul = 0x9E3779B1u;    // Fibonacci hashing

__device__ __forceinline__ u32 lds_u32(u32 a) {
  u32 r;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(a));
  return r;
}
__device__ __forceinline__ uint2 lds_v2(u32 a) {
  uint2 r;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(a));
  return r;
}
__device__ __forceinline__ uint4 lds_v4(u32 a) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a));
  return r;
}
__device__ __forceinline__ void sts_u32(u32 a, u32 v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void sts_v4(u32 a, u32 x, u32 y, u32 z, u32 w) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
__device__ __forceinline__ u32 atoms_cas_u32(u32 a, u32 cmp, u32 val) {
  u32 r;
  asm volatile("atom.shared.cas.b32 %0, [%1], %2, %3;" : "=r"(r) : "r"(a), "r"(cmp), "r"(val) : "memory");
  return r;
}

// Bucket = W consecutive keys (4*W bytes); payload of key slot x is at x + pay_delta.
template <int W>
__device__ __forceinline__ void j32_load_keys(u32 a, u32 (&k)[W]) {
  if (W == 4) {
    const uint4 t = lds_v4(a);
    k[0] = t.x; k[1] = t.y; k[W - 2] = t.z; k[W - 1] = t.w;
  } else {
    const uint2 t = lds_v2(a);
    k[0] = t.x; k[1] = t.y;
  }
}
// byte offset of the payload of the (first) slot holding key, from the bucket address; true = some slot matches
template <int W>
__device__ __forceinline__ bool j32_match(const u32 (&k)[W], u32 key, u32& pay_off) {
  if (W == 4) {
    const bool m0 = k[0] == key, m1 = k[1] == key, m2 = k[W - 2] == key, m3 = k[W - 1] == key;
    pay_off = m0 ? 0u : m1 ? 4u : m2 ? 8u : 12u;
    return m0 | m1 | m2 | m3;
  } else {
    const bool m0 = k[0] == key, m1 = k[1] == key;
    pay_off = m0 ? 0u : 4u;
    return m0 | m1;
  }
}

struct J32Acc {
  u64 key_sum;  // sum of matched keys (once per pair)
  u32 pay_sum;  // sum of (pay_r + pay_s) over pairs, mod 2^32 (it is shifted left by 32)
  u32 matches;  // per chunk; folded into a u64 by the caller
};

// Everything the probe needs, in registers.
struct J32Tab {
  u32 base;       // shared-space address of key bucket 0
  u32 pay_delta;  // payload array starts this many bytes after the key array (= nbuckets * 4 * W)
  u32 nbuckets;
  u32 radix_bits;
};
template <int W>
__device__ __forceinline__ u32 j32_home(const J32Tab& t, u32 k) {
  return t.base + __umulhi(k * kJ32Mul, t.nbuckets) * (4u * W);
}

// The key was not in its home bucket and that bucket is full: walk on (rare).
// Returns 1<<32 | payload when found, 0 otherwise (by value: a reference would pin the
// caller's accumulators to local memory).
template <int W>
__device__ __noinline__ u64 j32_probe_walk(u32 base, u32 pay_delta, u32 a, u32 k) {
  const u32 end = base + pay_delta;
#pragma unroll 1
  while (true) {
    a += 4u * W;
    if (a == end) a = base;
    u32 kk[W], off;
    j32_load_keys<W>(a, kk);
    if (j32_match<W>(kk, k, off)) return (1ull << 32) | lds_u32(a + pay_delta + off);
    if (kk[W - 1] == kJ32Empty) return 0ull;
  }
}

// G tuples against a chunk whose keys are unique: all bucket loads first, then the
// compares and payload loads, then the sums; stop at the first match.
// NOPAY: no tuple of either relation carries a payload (ctr[C_PAY_HI] == 0): the payload array is never read --
// the random LDS.32 per match was 4.6 of the ~10 shared-memory wavefronts a probe cost (ncu, round 1).
template <int W, int G, bool CHECK, bool NOPAY>
__device__ __forceinline__ void j32_probe_unique(const J32Tab& t, const u64* __restrict__ s, J32Acc& acc) {
  u32 a[G], kk[G][W];
#pragma unroll
  for (int g = 0; g < G; g++) {
    a[g] = j32_home<W>(t, (u32)s[g] >> t.radix_bits);
    j32_load_keys<W>(a[g], kk[g]);
  }
#pragma unroll
  for (int g = 0; g < G; g++) {
    const u32 lo = (u32)s[g], hi = (u32)(s[g] >> 32);
    const u32 k = lo >> t.radix_bits;
    u32 off;
    const bool found = j32_match<W>(kk[g], k, off);
    if (CHECK && s[g] == 0ull) continue;
    if (found) {
      if (!NOPAY) acc.pay_sum += lds_u32(a[g] + t.pay_delta + off) + hi;
      acc.key_sum += lo;
      acc.matches += 1;
    } else if (kk[g][W - 1] != kJ32Empty) {
      const u64 r = j32_probe_walk<W>(t.base, t.pay_delta, a[g], k);
      if (r) {
        if (!NOPAY) acc.pay_sum += (u32)r + hi;
        acc.key_sum += lo;
        acc.matches += 1;
      }
    }
  }
}

// Duplicate keys in the chunk: count every match, walk on while buckets are full.
// Returns matches << 32 | (sum of the matching R payloads mod 2^32).
template <int W>
__device__ __noinline__ u64 j32_probe_multi(u32 base, u32 nbuckets, u32 k) {
  const u32 pay_delta = nbuckets * (4u * W), end = base + pay_delta;
  u32 a = base + __umulhi(k * kJ32Mul, nbuckets) * (4u * W);
  u32 cnt = 0, pay = 0;
#pragma unroll 1
  while (true) {
    u32 kk[W];
    j32_load_keys<W>(a, kk);
#pragma unroll
    for (int i = 0; i < W; i++)
      if (kk[i] == k) { cnt++; pay += lds_u32(a + pay_delta + 4u * i); }
    if (kk[W - 1] == kJ32Empty) break;
    a += 4u * W;
    if (a == end) a = base;
  }
  return ((u64)cnt << 32) | pay;
}

// Skew handling (the reference's SKEW_HANDLING task splitting, parallel_radix_join.c:958-1055): the
// unit of work is a TASK = (partition p, slice k of S_p), S_p cut into slices of `slice` tuples, so a
// heavy-hitter partition (zipf foreign keys) is probed by many CTAs, each with its own copy of R_p's
// table, instead of serialising the whole join behind one CTA (z = 1.0: join 2.05 ms of a 3.3 ms
// total before).  With balanced partitions every partition is one task and nothing changes.
// One CTA builds tasks[t] = p | k << 32, partition-major, and ctr[C_NTASKS].
__global__ void __launch_bounds__(1024)
rx_tasks_kernel(const u64* __restrict__ offS, u32 nparts, u64 capS, u64 slice, u64* __restrict__ tasks,
                u64 max_tasks, u64* __restrict__ ctr) {
  __shared__ u32 warp_tot[32];
  __shared__ u32 s_carry;
  const u32 tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (u32 b0 = 0; b0 < nparts; b0 += 1024u) {
    const u32 p = b0 + tid;
    u32 nt = 0;
    if (p < nparts) {
      const u64 s0 = capS ? p * capS : offS[p];
      const u64 s1 = capS ? (offS[p] < s0 + capS ? offS[p] : s0 + capS) : offS[p + 1];
      const u64 ns = s1 - s0;
      nt = ns <= slice ? 1u : (u32)((ns + slice - 1) / slice);
    }
    u32 incl = nt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const u32 y = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane >= o) incl += y;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    if (wid == 0) {
      const u32 w = warp_tot[lane];
      u32 wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const u32 y = __shfl_up_sync(0xffffffffu, wi, o);
        if ((int)lane >= o) wi += y;
      }
      warp_tot[lane] = wi - w;
    }
    __syncthreads();
    const u32 carry = s_carry;
    u64 first = (u64)carry + warp_tot[wid] + incl - nt;
    for (u32 k = 0; k < nt; k++)
      if (first + k < max_tasks) tasks[first + k] = (u64)p | ((u64)k << 32);
    __syncthreads();
    if (tid == 1023) s_carry = carry + warp_tot[wid] + incl;
    __syncthreads();
  }
  if (tid == 0) ctr[C_NTASKS] = (u64)s_carry < max_tasks ? (u64)s_carry : max_tasks;
}

// Dynamic shared memory: nbuckets * 8 * W bytes.
template <int THREADS, int BATCH, int MINB, int W>
__global__ void __launch_bounds__(THREADS, MINB)
rx_join32_kernel(const u64* __restrict__ Rp, const u64* __restrict__ offR, const u64* __restrict__ Sp,
                 const u64* __restrict__ offS, u32 nparts, u32 radix_bits, u32 nbuckets, u64* __restrict__ ctr,
                 u64 capR, u64 capS, const u64* __restrict__ tasks, u64 slice, u32 nsrc, u64 src_stride,
                 u64 off_stride) {
  // nsrc > 1 (pipelined exchange): S arrived in nsrc chunks, each partitioned on its own -- source c's partitions
  // live at Sp + c * src_stride with fill levels offS + c * off_stride.  R_p's table is built ONCE and probed
  // with S_p of every source (a join is additive over disjoint parts of S); no S slicing then.
  static_assert(BATCH % 2 == 0, "BATCH tuples = BATCH/2 16-byte loads");
  static_assert(W == 2 || W == 4, "bucket of 2 or 4 keys");
  if (ctr[C_KEY_HI] != 0ull) return;  // keys wider than 32 bits: rx_join_kernel handles this join
  const bool nopay = ctr[C_PAY_HI] == 0ull;  // (kernel-uniform) nobody carries a payload: the payload array stays untouched
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ u32 s_dup;
  J32Tab tab;
  // pinned in registers: left alone, the compiler rebuilds the shared window base (4 uniform
  // instructions) and reloads the parameters from the constant bank around every access
  tab.base = (u32)__cvta_generic_to_shared(smem_raw);
  asm volatile("mov.u32 %0, %0;" : "+r"(tab.base));
  tab.pay_delta = nbuckets * (4u * W);
  asm volatile("mov.u32 %0, %0;" : "+r"(tab.pay_delta));
  tab.nbuckets = nbuckets;
  asm volatile("mov.u32 %0, %0;" : "+r"(tab.nbuckets));
  tab.radix_bits = radix_bits;
  const u32 tab_end = tab.base + tab.pay_delta;  // end of the key array

  const u32 tid = threadIdx.x;
  const u32 chunk_cap = (W * nbuckets / 4u) * 3u;  // <= 0.75 load in the worst case; ~0.3 at the planned fan-out
  constexpr u64 STEP = (u64)THREADS * BATCH;
  u64 matches = 0, key_sum = 0, maxpart = 0, chunked = 0;
  u32 pay_sum = 0;

  const u32 ntasks = (u32)ctr[C_NTASKS];
  for (u32 t = blockIdx.x; t < ntasks; t += gridDim.x) {
    const u64 task = tasks[t];
    const u32 p = (u32)task;
    const u64 k = task >> 32;
    // exact layout: partition p = [off[p], off[p+1]); histogram-free layout (cap != 0, see ScOpt
    // in rx_kernels.cuh): [p*cap, min(off[p], (p+1)*cap)) with off = the scatter's final cursors
    const u64 r0 = capR ? p * capR : offR[p];
    const u64 r1 = capR ? (offR[p] < r0 + capR ? offR[p] : r0 + capR) : offR[p + 1];
    const u64 nr = r1 - r0;
    if (k == 0) maxpart = nr > maxpart ? nr : maxpart;
    if (nr == 0) continue;
    if (nsrc == 1) {  // nothing to probe with: skip the build too
      const u64 q0 = (capS ? p * capS : offS[p]) + k * slice;
      const u64 q1 = capS ? (offS[p] < p * capS + capS ? offS[p] : p * capS + capS) : offS[p + 1];
      if (q1 <= q0) continue;
    }
    if (nr > chunk_cap && k == 0) chunked++;
    // an R partition beyond the table's capacity is cut into EQUAL chunks (two chunks at load 0.6 probe
    // faster than one at 0.75 and one at 0.4: full buckets are what costs)
    const u64 nchunk = (nr + chunk_cap - 1) / chunk_cap;
    const u64 cstep = nchunk > 1 ? (nr + nchunk - 1) / nchunk : chunk_cap;
    for (u64 c0 = r0; c0 < r1; c0 += cstep) {
      const u64 c1 = c0 + cstep < r1 ? c0 + cstep : r1;
      __syncthreads();  // the previous probe is over before the table is cleared
      // nbuckets * W is a multiple of 4 (host): the key array is whole 16-byte words
      for (u32 a = tab.base + 16u * tid; a < tab_end; a += 16u * THREADS)
        sts_v4(a, kJ32Empty, kJ32Empty, kJ32Empty, kJ32Empty);
      if (tid == 0) s_dup = 0;
      __syncthreads();

      // ---- build: BATCH loads in flight per thread, then the CAS inserts ----
      for (u64 b0 = c0; b0 < c1; b0 += STEP) {
        u64 rv[BATCH];
#pragma unroll
        for (int j = 0; j < BATCH; j++) {
          const u64 i = b0 + (u64)j * THREADS + tid;
          rv[j] = i < c1 ? ld_stream(Rp + i) : 0ull;
        }
        bool dup = false;
#pragma unroll
        for (int j = 0; j < BATCH; j++) {
          if (rv[j] == 0ull) continue;
          const u32 k = (u32)rv[j] >> radix_bits;
          u32 a = j32_home<W>(tab, k);  // bucket address; slot i of the bucket at a + 4*i
          u32 i = 0;
#pragma unroll 1
          while (true) {
            const u32 old = atoms_cas_u32(a + 4u * i, kJ32Empty, k);
            if (old == kJ32Empty) break;
            dup |= old == k;
            if (++i == W) {
              i = 0;
              a += 4u * W;
              if (a == tab_end) a = tab.base;
            }
          }
          if (!nopay) sts_u32(a + tab.pay_delta + 4u * i, (u32)(rv[j] >> 32));
        }
        if (dup) s_dup = 1;
      }
      __syncthreads();
      const bool unique = s_dup == 0;

      J32Acc acc;
      acc.key_sum = 0; acc.pay_sum = 0; acc.matches = 0;
      for (u32 src = 0; src < nsrc; src++) {
      const u64* __restrict__ Sp_ = Sp + (u64)src * src_stride;
      const u64* __restrict__ offS_ = offS + (u64)src * off_stride;
      u64 s0 = capS ? p * capS : offS_[p];
      u64 s1 = capS ? (offS_[p] < s0 + capS ? offS_[p] : s0 + capS) : offS_[p + 1];
      s0 += k * slice;  // slice k of S_p (k == 0 and one slice: the whole partition)
      if (s1 > s0 && s1 - s0 > slice) s1 = s0 + slice;
      if (s1 <= s0) continue;
      if (!unique) {
        for (u64 i = s0 + tid; i < s1; i += THREADS) {
          const u64 sv = ld_stream(Sp_ + i);
          const u32 lo = (u32)sv, hi = (u32)(sv >> 32);
          const u64 r = j32_probe_multi<W>(tab.base, nbuckets, lo >> radix_bits);
          const u32 cnt = (u32)(r >> 32);
          acc.matches += cnt;
          acc.key_sum += (u64)cnt * lo;
          if (!nopay) acc.pay_sum += (u32)r + cnt * hi;  // (the payload array is uninitialised otherwise)
        }
      } else {
        // ---- probe: batches of 16-byte loads from the first even index, two register batches
        // in flight (the next one is loading while this one is probed); the last, partial
        // batch and the odd head tuple go through the checked variant ----
        const u64 i0 = s0 + (s0 & 1ull);
        const u64 nfull = i0 < s1 ? (s1 - i0) / STEP : 0;
        const u64 tail0 = i0 + nfull * STEP;
        const u64 nb = nfull + (tail0 < s1 ? 1 : 0);
        u64 A[BATCH], B[BATCH];
        auto load = [&](u64 (&v)[BATCH], u64 bi) {
          if (bi < nfull) {
            const u64* srcp = Sp_ + i0 + bi * STEP + 2u * tid;
#pragma unroll
            for (int j = 0; j < BATCH / 2; j++) {
              const ulonglong2 x = ld_stream_v2(srcp + 2u * j * THREADS);
              v[2 * j] = x.x;
              v[2 * j + 1] = x.y;
            }
          } else {
#pragma unroll
            for (int j = 0; j < BATCH; j++) {
              const u64 i = tail0 + (u64)j * THREADS + tid;
              v[j] = i < s1 ? ld_stream(Sp_ + i) : 0ull;
            }
          }
        };
        auto process = [&](u64 (&v)[BATCH], u64 bi) {
          constexpr int G = BATCH < 4 ? BATCH : (BATCH % 4 == 0 ? 4 : 2);  // bucket loads in flight per thread
          if (bi < nfull) {
            if (nopay) {
#pragma unroll
              for (int g = 0; g < BATCH; g += G) j32_probe_unique<W, G, false, true>(tab, v + g, acc);
            } else {
#pragma unroll
              for (int g = 0; g < BATCH; g += G) j32_probe_unique<W, G, false, false>(tab, v + g, acc);
            }
          } else if (nopay) {
#pragma unroll
            for (int g = 0; g < BATCH; g += G) j32_probe_unique<W, G, true, true>(tab, v + g, acc);
          } else {
#pragma unroll
            for (int g = 0; g < BATCH; g += G) j32_probe_unique<W, G, true, false>(tab, v + g, acc);
          }
        };
        if (tid == 0 && (s0 & 1ull)) {
          u64 h[1] = {ld_stream(Sp_ + s0)};
          if (nopay) j32_probe_unique<W, 1, true, true>(tab, h, acc);
          else j32_probe_unique<W, 1, true, false>(tab, h, acc);
        }
        if (nb) {
          u64 bi = 0;
          load(A, 0);
          while (true) {
            if (bi + 1 < nb) load(B, bi + 1);
            process(A, bi);
            if (++bi >= nb) break;
            if (bi + 1 < nb) load(A, bi + 1);
            process(B, bi);
            if (++bi >= nb) break;
          }
        }
      }
      }  // sources
      matches += acc.matches;  // < 2^32 per thread per chunk
      key_sum += acc.key_sum;
      pay_sum += acc.pay_sum;
    }
  }
  u64 v[2] = {matches, 2ull * key_sum + ((u64)pay_sum << 32)};
  u64* const d[2] = {ctr + C_MATCHES, ctr + C_CHECKSUM};
  block_sum_atomic<2, THREADS>(v, d);
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 16));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 8));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 4));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 2));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 1));
  if (lane_id() == 0 && maxpart) atomicMax(ctr + C_MAX_PART, maxpart);
  if (tid == 0 && chunked) atomicAdd(ctr + C_TOO_BIG, chunked);
}

}  // namespace hj
