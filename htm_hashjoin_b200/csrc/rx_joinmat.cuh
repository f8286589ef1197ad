// rx_joinmat.cuh -- result materialisation for the radix join.
//
// The reference stops at counting ("TODO: we don't materialize",
// mc/src/no_partitioning_join.c:301, parallel_radix_join.c:359); an engine needs the pairs.
// hj_join_materialize emits every matching pair as {r tuple, s tuple} (16 bytes) into ONE
// dense device array of exactly `matches` entries, in two passes over the partitioned
// relations -- the textbook count / scan / write scheme, chosen over a single global cursor
// because the output size is known before anything is written (no overflow protocol, the
// caller can size the buffer from the count pass) and every claim is a shared-memory atomic:
//
//   count  rx_join32_mat_kernel<.., false>: per-partition join, matches of partition p ->
//          part_cnt[p]
//   scan   rx_mat_scan_kernel: exclusive prefix of part_cnt in place, total -> ctr[C_MAT_TOTAL]
//   write  rx_join32_mat_kernel<.., true>: the same join again; partition p owns
//          out[part_cnt[p] .. part_cnt[p+1]).  Inside it a warp handles 32*G probe tuples per
//          step: the home-bucket matches of the step are ranked with ballots, ONE shared-memory
//          atomicAdd per warp and step claims their slots, and the 16-byte pair stores of a
//          step are consecutive (coalesced).  Matches beyond the home bucket (a full bucket,
//          duplicate build keys) claim one slot each.
//
// Same table as rx_join32 (rx_join32.cuh): 32-bit keys, SoA buckets of W keys, one CTA per
// partition, R partitions larger than the table are processed in chunks.  Pairs of one
// partition are contiguous; their order inside the partition is not deterministic.
// Algorithmic bytes: 2 x 8 B read per tuple (two join passes) + 16 B written per pair.
#pragma once
#include "hj_common.cuh"
#include "rx_join32.cuh"

namespace hj {

template <int THREADS, int W, bool WRITE>
__global__ void __launch_bounds__(THREADS, 1)
rx_join32_mat_kernel(const u64* __restrict__ Rp, const u64* __restrict__ offR, const u64* __restrict__ Sp,
                     const u64* __restrict__ offS, u32 nparts, u32 radix_bits, u32 nbuckets, u64* __restrict__ ctr,
                     u64 capR, u64 capS, u64* __restrict__ part_cnt, ulonglong2* __restrict__ out) {
  constexpr int G = 4;  // probe tuples per lane and step
  constexpr int BATCH = 4;
  if (ctr[C_KEY_HI] != 0ull) return;  // host reports HJ_EINVAL: 32-bit keys only
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ unsigned long long s_cur;  // pairs of this partition claimed so far
  J32Tab tab;
  tab.base = (u32)__cvta_generic_to_shared(smem_raw);
  tab.pay_delta = nbuckets * (4u * W);
  tab.nbuckets = nbuckets;
  tab.radix_bits = radix_bits;
  const u32 tab_end = tab.base + tab.pay_delta;
  const u32 tid = threadIdx.x, lane = tid & 31u;
  const u32 lt = (1u << lane) - 1u;
  const u32 chunk_cap = (W * nbuckets / 4u) * 3u;
  u64 matches = 0, key_sum = 0;
  u32 pay_sum = 0;

  for (u32 p = blockIdx.x; p < nparts; p += gridDim.x) {
    const u64 r0 = capR ? p * capR : offR[p];
    const u64 r1 = capR ? (offR[p] < r0 + capR ? offR[p] : r0 + capR) : offR[p + 1];
    const u64 s0 = capS ? p * capS : offS[p];
    const u64 s1 = capS ? (offS[p] < s0 + capS ? offS[p] : s0 + capS) : offS[p + 1];
    const u64 out0 = WRITE ? part_cnt[p] : 0ull;
    __syncthreads();  // s_cur of the previous partition has been read
    if (tid == 0) s_cur = 0ull;
    const u64 nr = r1 - r0;
    const u64 nchunk = (nr + chunk_cap - 1) / chunk_cap;  // equal chunks, as rx_join32_kernel
    const u64 cstep = nchunk > 1 ? (nr + nchunk - 1) / nchunk : chunk_cap;
    for (u64 c0 = r0; c0 < r1 && s1 > s0; c0 += cstep) {
      const u64 c1 = c0 + cstep < r1 ? c0 + cstep : r1;
      __syncthreads();
      for (u32 a = tab.base + 16u * tid; a < tab_end; a += 16u * THREADS)
        sts_v4(a, kJ32Empty, kJ32Empty, kJ32Empty, kJ32Empty);
      __syncthreads();
      // ---- build (as rx_join32_kernel) ----
      for (u64 b0 = c0; b0 < c1; b0 += (u64)THREADS * BATCH) {
        u64 rv[BATCH];
#pragma unroll
        for (int j = 0; j < BATCH; j++) {
          const u64 i = b0 + (u64)j * THREADS + tid;
          rv[j] = i < c1 ? ld_stream(Rp + i) : 0ull;
        }
#pragma unroll
        for (int j = 0; j < BATCH; j++) {
          if (rv[j] == 0ull) continue;
          const u32 k = (u32)rv[j] >> radix_bits;
          u32 a = j32_home<W>(tab, k);
          u32 i = 0;
#pragma unroll 1
          while (atoms_cas_u32(a + 4u * i, kJ32Empty, k) != kJ32Empty) {
            if (++i == W) {
              i = 0;
              a += 4u * W;
              if (a == tab_end) a = tab.base;
            }
          }
          sts_u32(a + tab.pay_delta + 4u * i, (u32)(rv[j] >> 32));
        }
      }
      __syncthreads();

      // ---- probe: whole warps stay in the loop (ballots), G tuples per lane and step ----
      const u64 ns = s1 - s0;
      const u64 steps = (ns + (u64)THREADS * G - 1) / ((u64)THREADS * G);
      for (u64 st = 0; st < steps; st++) {
        u64 sv[G];
        u32 a[G], kk[G][W];
#pragma unroll
        for (int g = 0; g < G; g++) {
          const u64 i = s0 + (st * G + g) * THREADS + tid;
          sv[g] = i < s1 ? ld_stream(Sp + i) : 0ull;
        }
#pragma unroll
        for (int g = 0; g < G; g++) {
          a[g] = j32_home<W>(tab, (u32)sv[g] >> radix_bits);
          j32_load_keys<W>(a[g], kk[g]);
        }
        // home-bucket matches: every slot of the bucket that holds the key is a pair
        u32 hit[G], mask[G][W], total = 0;
#pragma unroll
        for (int g = 0; g < G; g++) {
          const u32 k = (u32)sv[g] >> radix_bits;
          hit[g] = 0;
#pragma unroll
          for (int w = 0; w < W; w++) {
            const bool m = sv[g] != 0ull && kk[g][w] == k;
            hit[g] |= m ? 1u << w : 0u;
            mask[g][w] = __ballot_sync(0xffffffffu, m);
            total += __popc(mask[g][w]);
          }
        }
        u64 base = 0;
        if (WRITE) {
          if (lane == 0 && total) base = atomicAdd(&s_cur, (unsigned long long)total);
          base = __shfl_sync(0xffffffffu, base, 0) + out0;
        }
        u32 before = 0;
#pragma unroll
        for (int g = 0; g < G; g++) {
          const u32 lo = (u32)sv[g], hi = (u32)(sv[g] >> 32);
#pragma unroll
          for (int w = 0; w < W; w++) {
            if (hit[g] >> w & 1u) {
              const u32 rp = lds_u32(a[g] + tab.pay_delta + 4u * w);
              pay_sum += rp + hi;
              key_sum += lo;
              matches++;
              if (WRITE) out[base + before + __popc(mask[g][w] & lt)] = make_ulonglong2(((u64)rp << 32) | lo, sv[g]);
            }
            before += __popc(mask[g][w]);
          }
        }
        // a full home bucket: the key may continue in the following buckets (rare)
#pragma unroll
        for (int g = 0; g < G; g++) {
          if (sv[g] == 0ull || kk[g][W - 1] == kJ32Empty) continue;
          const u32 lo = (u32)sv[g], hi = (u32)(sv[g] >> 32), k = lo >> radix_bits;
          u32 b = a[g];
#pragma unroll 1
          while (true) {
            b += 4u * W;
            if (b == tab_end) b = tab.base;
            if (b == a[g]) break;  // went round a completely full table
            u32 k2[W];
            j32_load_keys<W>(b, k2);
#pragma unroll
            for (int w = 0; w < W; w++) {
              if (k2[w] == k) {
                const u32 rp = lds_u32(b + tab.pay_delta + 4u * w);
                pay_sum += rp + hi;
                key_sum += lo;
                matches++;
                if (WRITE) out[out0 + atomicAdd(&s_cur, 1ull)] = make_ulonglong2(((u64)rp << 32) | lo, sv[g]);
              }
            }
            if (k2[W - 1] == kJ32Empty) break;
          }
        }
      }
    }
    if (!WRITE) {
      // count pass: the partition's total through the same shared counter
      u32 c = (u32)matches;  // matches is reset per partition in this mode
      c += __shfl_xor_sync(0xffffffffu, c, 16);
      c += __shfl_xor_sync(0xffffffffu, c, 8);
      c += __shfl_xor_sync(0xffffffffu, c, 4);
      c += __shfl_xor_sync(0xffffffffu, c, 2);
      c += __shfl_xor_sync(0xffffffffu, c, 1);
      if (lane == 0 && c) atomicAdd(&s_cur, (unsigned long long)c);
      matches = 0;
      __syncthreads();
      if (tid == 0) part_cnt[p] = s_cur;
    }
  }
  if (WRITE) {
    u64 v[2] = {matches, 2ull * key_sum + ((u64)pay_sum << 32)};
    u64* const d[2] = {ctr + C_MATCHES, ctr + C_CHECKSUM};
    block_sum_atomic<2, THREADS>(v, d);
  }
}

// Exclusive prefix of cnt[0..n) in place (one CTA), cnt[n] = total, total also to *total_out.
__global__ void __launch_bounds__(1024) rx_mat_scan_kernel(u64* __restrict__ cnt, u32 n, u64* __restrict__ total_out) {
  __shared__ u64 warp_tot[32];
  __shared__ u64 s_carry;
  const u32 tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
  if (tid == 0) s_carry = 0ull;
  __syncthreads();
  for (u32 b0 = 0; b0 < n; b0 += 1024u) {
    const u32 i = b0 + tid;
    const u64 x = i < n ? cnt[i] : 0ull;
    u64 incl = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const u64 y = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane >= o) incl += y;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    if (wid == 0) {
      const u64 w = warp_tot[lane];
      u64 wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const u64 y = __shfl_up_sync(0xffffffffu, wi, o);
        if ((int)lane >= o) wi += y;
      }
      warp_tot[lane] = wi - w;
    }
    __syncthreads();
    const u64 carry = s_carry;
    if (i < n) cnt[i] = carry + warp_tot[wid] + incl - x;
    __syncthreads();
    if (tid == 1023) s_carry = carry + warp_tot[wid] + incl;
    __syncthreads();
  }
  if (tid == 0) {
    cnt[n] = s_carry;
    *total_out = s_carry;
  }
}

}  // namespace hj
