// rx_wide.cuh -- the join paths on 16-byte tuples {u64 key, u64 payload}.
//
// The reference's mc/ code switches its tuple from {int32 key, int32 payload} to {int64 key, int64 payload}
// with -DKEY_8B (mc/src/types.h:22-37); every loop of parallel_radix_join.c / no_partitioning_join.c is the
// same on the wider tuple.  So are these kernels: the same steps as rx_kernels.cuh / gt_kernels.cuh
// (histogram -> scan -> tile-sorting scatter x passes -> per-partition shared-memory build + probe;
// bitmap-claimed global table with an overflow table), moving 16 bytes per tuple -- half the tuples per
// second at the same HBM roofline, which is what SURVEY 8f-4 asks to be shown.  Key 0 is the empty slot,
// keys are full 64-bit values (cfg.key_mask applies to the key word), the payload word travels untouched.
//   checksum = sum over matching pairs of (r.key + r.payload + s.key + s.payload) mod 2^64.
// Exact (histogram-based) layout only: the histogram-free regions, the TMA flush and the 32-bit-key join of
// the 8-byte path are tuned for the headline workload and are not repeated here.
#pragma once
#include "hj_common.cuh"
#include "gt_kernels.cuh"
#include "rx_kernels.cuh"

namespace hj {

typedef ulonglong2 tup16;  // x = key, y = payload

__device__ __forceinline__ tup16 ld_stream_t16(const tup16* p) {
  tup16 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(r.x), "=l"(r.y) : "l"(p));
  return r;
}

// ---- histogram over every radix bit of the plan (rx_hist_kernel on 16-byte tuples) ---------------------
constexpr int kWHistThreads = 512, kWHistIPT = 4;
__global__ void __launch_bounds__(kWHistThreads)
w_hist_kernel(const tup16* __restrict__ in, u64 n, Digit dg, u32 nbins, u64* __restrict__ ghist, u64* __restrict__ ctr) {
  extern __shared__ u32 shist[];
  const u32 tid = threadIdx.x;
  u32 zeros = 0;
  for (u32 i = tid; i < nbins; i += kWHistThreads) shist[i] = 0;
  __syncthreads();
  constexpr int TILE = kWHistThreads * kWHistIPT;
  const u64 ntiles = (n + TILE - 1) / TILE;
  for (u64 t = blockIdx.x; t < ntiles; t += gridDim.x) {
    tup16 v[kWHistIPT];
#pragma unroll
    for (int j = 0; j < kWHistIPT; j++) {
      const u64 i = t * TILE + (u64)j * kWHistThreads + tid;
      v[j] = i < n ? ld_stream_t16(in + i) : make_ulonglong2(0ull, 0ull);
      zeros += i < n && (v[j].x & dg.key_mask) == 0ull;
    }
#pragma unroll
    for (int j = 0; j < kWHistIPT; j++)
      if (v[j].x & dg.key_mask) atomicAdd(shist + digit_of(dg, v[j].x), 1u);
  }
  __syncthreads();
  for (u32 i = tid; i < nbins; i += kWHistThreads) {
    const u32 c = shist[i];
    if (c) atomicAdd(ghist + i, (u64)c);
  }
  zeros = __reduce_add_sync(0xffffffffu, zeros);
  if (zeros && lane_id() == 0) atomicAdd(ctr + C_ZERO_KEYS, (u64)zeros);
}

// ---- scatter: the digit-sorted CTA tile of rx_scatter_kernel, 16 bytes per element -----------------------
// parent_off / parent_tiles / cursors: what rx_scan_kernel wrote (it only sees counts).  Tile = 256 threads x
// 8 tuples = 2048 tuples (32 KB); runs leave as 16-byte stores of consecutive lanes to consecutive addresses.
// Dynamic shared memory: tile (TILE*16) | gdelta (F*8) | hist (F*4) | lbase (F*4).
constexpr int kWScThreads = 256, kWScIPT = 8;
__global__ void __launch_bounds__(kWScThreads, 4)
w_scatter_kernel(const tup16* __restrict__ in, tup16* __restrict__ out, Digit dg, u32 F, u32 nparent,
                 const u64* __restrict__ parent_off, const u64* __restrict__ parent_tiles, u64* __restrict__ cursors) {
  constexpr int TILE = kWScThreads * kWScIPT;
  extern __shared__ __align__(16) unsigned char smem_w[];
  tup16* tile = reinterpret_cast<tup16*>(smem_w);
  u64* gdelta = reinterpret_cast<u64*>(tile + TILE);
  u32* hist = reinterpret_cast<u32*>(gdelta + F);
  u32* lbase = hist + F;
  __shared__ u32 warp_tot[kWScThreads / 32];
  __shared__ u32 s_total;
  const u32 tid = threadIdx.x;
  for (u32 i = tid; i < F; i += kWScThreads) hist[i] = 0;
  __syncthreads();
  const u32 total_tiles = (u32)parent_tiles[nparent];
  const u32 per_thread = (F + kWScThreads - 1) / kWScThreads;
  for (u32 t = blockIdx.x; t < total_tiles; t += gridDim.x) {
    u32 lo = 0, hi = nparent;  // last p with parent_tiles[p] <= t
    while (hi - lo > 1) {
      const u32 mid = (lo + hi) >> 1;
      if (parent_tiles[mid] <= t) lo = mid; else hi = mid;
    }
    const u32 p = lo;
    const u64 begin = parent_off[p] + (u64)(t - parent_tiles[p]) * TILE;
    const u64 pend = parent_off[p + 1];
    const u32 cnt = (u32)(pend - begin < (u64)TILE ? pend - begin : (u64)TILE);
    tup16 v[kWScIPT];
    u32 dr[kWScIPT];
#pragma unroll
    for (int j = 0; j < kWScIPT; j++) {
      const u32 i = j * kWScThreads + tid;
      v[j] = i < cnt ? ld_stream_t16(in + begin + i) : make_ulonglong2(0ull, 0ull);
    }
#pragma unroll
    for (int j = 0; j < kWScIPT; j++)
      if (v[j].x & dg.key_mask) {
        const u32 d = digit_of(dg, v[j].x);
        dr[j] = (d << 16) | atomicAdd(hist + d, 1u);
      }
    __syncthreads();  // (also: the previous tile is flushed)
    {
      const u32 b0 = tid * per_thread;
      u32 local = 0;
      for (u32 k = 0; k < per_thread; k++)
        if (b0 + k < F) local += hist[b0 + k];
      u32 incl = local;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const u32 y = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)lane_id() >= o) incl += y;
      }
      if (lane_id() == 31) warp_tot[tid >> 5] = incl;
      __syncthreads();
      if (tid < 32) {
        const u32 w = tid < kWScThreads / 32 ? warp_tot[tid] : 0u;
        u32 wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const u32 y = __shfl_up_sync(0xffffffffu, wi, o);
          if ((int)tid >= o) wi += y;
        }
        if (tid < kWScThreads / 32) warp_tot[tid] = wi - w;
        if (tid == kWScThreads / 32 - 1) s_total = wi;  // tuples of the tile with a non-zero key
      }
      __syncthreads();
      u32 run = warp_tot[tid >> 5] + incl - local;
      for (u32 k = 0; k < per_thread; k++) {
        const u32 b = b0 + k;
        if (b < F) {
          const u32 c = hist[b];
          hist[b] = 0;
          lbase[b] = run;
          if (c) gdelta[b] = atomicAdd(cursors + (u64)p * F + b, (u64)c) - run;
          run += c;
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < kWScIPT; j++)
      if (v[j].x & dg.key_mask) tile[lbase[dr[j] >> 16] + (dr[j] & 0xffffu)] = v[j];
    __syncthreads();
    const u32 total = s_total;
    for (u32 i = tid; i < total; i += kWScThreads) {
      const tup16 x = tile[i];
      out[gdelta[digit_of(dg, x.x)] + i] = x;
    }
  }
}

// ---- per-partition join: bucket_chaining_join / histogram_join (parallel_radix_join.c:231-283,306-369) ----
// One CTA per partition, open-addressing table of {key, payload} slots in shared memory (slots = power of two),
// R_p inserted in chunks of slots/2, S_p streamed once per chunk.  Dynamic shared memory: slots * 16 bytes.
constexpr int kWJoinThreads = 512;
__global__ void __launch_bounds__(kWJoinThreads, 2)
w_join_kernel(const tup16* __restrict__ Rp, const u64* __restrict__ offR, const tup16* __restrict__ Sp,
              const u64* __restrict__ offS, u32 nparts, u32 radix_bits, u32 slots, u64 key_mask, u64* __restrict__ ctr) {
  extern __shared__ __align__(16) unsigned char smem_w[];
  u64* tkey = reinterpret_cast<u64*>(smem_w);
  u64* tpay = tkey + slots;
  const u32 tid = threadIdx.x, smask = slots - 1, chunk_cap = slots / 2;
  const u32 hshift = 32 - (31 - __clz(slots));
  u64 matches = 0, checksum = 0, maxpart = 0;
  for (u32 p = blockIdx.x; p < nparts; p += gridDim.x) {
    const u64 r0 = offR[p], r1 = offR[p + 1], s0 = offS[p], s1 = offS[p + 1];
    maxpart = max(maxpart, r1 - r0);
    if (r1 == r0 || s1 == s0) continue;
    for (u64 c0 = r0; c0 < r1; c0 += chunk_cap) {
      const u64 c1 = c0 + chunk_cap < r1 ? c0 + chunk_cap : r1;
      __syncthreads();
      for (u32 i = tid; i < slots; i += kWJoinThreads) tkey[i] = 0ull;
      __syncthreads();
      for (u64 i = c0 + tid; i < c1; i += kWJoinThreads) {
        const tup16 r = ld_stream_t16(Rp + i);
        const u64 key = r.x & key_mask;
        u32 slot = join_slot(key, radix_bits, hshift);
        while (atomicCAS(tkey + slot, 0ull, key) != 0ull) slot = (slot + 1) & smask;
        tpay[slot] = r.x - key + r.y;  // everything of the tuple but the key, summed (only sums leave the kernel)
      }
      __syncthreads();
      for (u64 i = s0 + tid; i < s1; i += kWJoinThreads) {
        const tup16 s = ld_stream_t16(Sp + i);
        const u64 key = s.x & key_mask;
        u32 slot = join_slot(key, radix_bits, hshift);
        u64 k;
        while ((k = tkey[slot]) != 0ull) {
          if (k == key) { ++matches; checksum += k + tpay[slot] + s.x + s.y; }
          slot = (slot + 1) & smask;
        }
      }
    }
  }
  u64 v[2] = {matches, checksum};
  u64* const d[2] = {ctr + C_MATCHES, ctr + C_CHECKSUM};
  block_sum_atomic<2, kWJoinThreads>(v, d);
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 16));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 8));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 4));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 2));
  maxpart = max(maxpart, __shfl_xor_sync(0xffffffffu, maxpart, 1));
  if (lane_id() == 0 && maxpart) atomicMax(ctr + C_MAX_PART, maxpart);
}

// ---- global table on 16-byte slots (AtomicHashBuild.hpp:37-85 on the KEY_8B tuple) ------------------------
// The claim goes to the slot bitmap (gt_kernels.cuh), one global atomicOr per attempt, the winner stores its
// 16-byte tuple; a tuple whose budget runs out goes to the overflow list.  table / ovf_list / ovf_table hold tup16.
constexpr int kWGtIPT = 4;
__global__ void __launch_bounds__(kGtThreads)
gw_build_kernel(const tup16* __restrict__ R, u64 n, tup16* __restrict__ table, u32* __restrict__ claim, u32 mask,
                u64 key_mask, u32 L, tup16* __restrict__ ovf_list, u64* __restrict__ ctr) {
  constexpr int TILE = kGtThreads * kWGtIPT;
  const u32 tid = threadIdx.x;
  u32 steps = 0, zeros = 0;
  const u64 nchunks = (n + TILE - 1) / TILE;
  for (u64 c = blockIdx.x; c < nchunks; c += gridDim.x) {
    tup16 v[kWGtIPT];
    u32 att[kWGtIPT], pend = 0, failed = 0;
#pragma unroll
    for (int j = 0; j < kWGtIPT; j++) {
      const u64 i = c * TILE + (u64)j * kGtThreads + tid;
      v[j] = i < n ? ld_stream_t16(R + i) : make_ulonglong2(0ull, 0ull);
      att[j] = 0;
      if (v[j].x & key_mask) pend |= 1u << j;
      else zeros += i < n;
    }
#pragma unroll 1
    while (pend) {
      u32 old[kWGtIPT];
#pragma unroll
      for (int j = 0; j < kWGtIPT; j++)
        if (pend >> j & 1u) {
          const u32 slot = (((u32)(v[j].x & key_mask) & mask) + att[j]) & mask;
          old[j] = atomicOr(claim + (slot >> 5), 1u << (slot & 31u));
        }
#pragma unroll
      for (int j = 0; j < kWGtIPT; j++)
        if (pend >> j & 1u) {
          const u32 slot = (((u32)(v[j].x & key_mask) & mask) + att[j]) & mask;
          const u32 b = slot & 31u;
          if (!(old[j] >> b & 1u)) {
            table[slot] = v[j];
            pend &= ~(1u << j);
          } else {
            ++steps;
            att[j] += gt_advance(old[j] | (1u << b), b);
            if (att[j] >= L) { pend &= ~(1u << j); failed |= 1u << j; }
          }
        }
    }
    if (failed) {
      u64 dst = atomicAdd(ctr + C_CONFLICTS, (u64)__popc(failed));
#pragma unroll
      for (int j = 0; j < kWGtIPT; j++)
        if (failed >> j & 1u) ovf_list[dst++] = v[j];
    }
  }
  u64 s[2] = {steps, zeros};
  u64* const d[2] = {ctr + C_PROBE_STEPS, ctr + C_ZERO_KEYS};
  block_sum_atomic<2, kGtThreads>(s, d);
}

// overflow table of 16-byte entries, sized from the device-side count like gt_ovf_*: the key word is claimed
// with a CAS, the payload word follows (nobody probes before the kernel ends)
__global__ void __launch_bounds__(256) gw_ovf_clear_kernel(tup16* __restrict__ ovf_table, const u64* __restrict__ ctr,
                                                            u64 capacity) {
  const u64 count = ctr[C_CONFLICTS];
  if (count == 0) return;
  u64 slots = ovf_slots_for(count);
  if (slots > capacity) slots = capacity;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < slots; i += (u64)gridDim.x * blockDim.x)
    ovf_table[i] = make_ulonglong2(0ull, 0ull);
}
__global__ void __launch_bounds__(256) gw_ovf_insert_kernel(const tup16* __restrict__ ovf_list, tup16* __restrict__ ovf_table,
                                                             const u64* __restrict__ ctr, u64 key_mask, u64 capacity) {
  const u64 count = ctr[C_CONFLICTS];
  if (count == 0) return;
  u64 slots = ovf_slots_for(count);
  if (slots > capacity) slots = capacity;
  const u64 omask = slots - 1;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (u64)gridDim.x * blockDim.x) {
    const tup16 v = ovf_list[i];
    u64 slot = ovf_hash(v.x & key_mask, omask);
    while (atomicCAS(&ovf_table[slot].x, 0ull, v.x) != 0ull) slot = (slot + 1) & omask;
    ovf_table[slot].y = v.y;
  }
}

__global__ void __launch_bounds__(kGtThreads)
gw_probe_kernel(const tup16* __restrict__ S, u64 m, const tup16* __restrict__ table, u32 mask, u64 key_mask, u32 L,
                const tup16* __restrict__ ovf_table, u64 ovf_capacity, u64* __restrict__ ctr) {
  const u64 ovf_count = ctr[C_CONFLICTS];
  u64 omask = 0;
  if (ovf_count) {
    u64 slots = ovf_slots_for(ovf_count);
    if (slots > ovf_capacity) slots = ovf_capacity;
    omask = slots - 1;
  }
  u64 checksum = 0;
  u32 wmatches = 0, omatches = 0, zeros = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (u64)gridDim.x * blockDim.x) {
    const tup16 s = ld_stream_t16(S + i);
    const u64 key = s.x & key_mask;
    if (key == 0ull) { ++zeros; continue; }
    const u32 home = (u32)key & mask;
    bool full = true;
    for (u32 k = 0; k < L; k++) {  // AtomicHashBuild.hpp:75-83
      const tup16 t = table[(home + k) & mask];
      if ((t.x & key_mask) == 0ull) { full = false; break; }
      if ((t.x & key_mask) == key) { ++wmatches; checksum += t.x + t.y + s.x + s.y; }
    }
    if (full && ovf_count) {
      u64 os = ovf_hash(key, omask);
      while (true) {
        const tup16 o = ovf_table[os];
        if ((o.x & key_mask) == 0ull) break;
        if ((o.x & key_mask) == key) { ++omatches; checksum += o.x + o.y + s.x + s.y; }
        os = (os + 1) & omask;
      }
    }
  }
  u64 v[4] = {(u64)wmatches + omatches, checksum, (u64)wmatches, (u64)zeros};
  u64* const d[4] = {ctr + C_MATCHES, ctr + C_CHECKSUM, ctr + C_WINDOW_MATCHES, ctr + C_ZERO_KEYS};
  block_sum_atomic<4, kGtThreads>(v, d);
}

// verification sums over 16-byte tuples: sum of key + payload words
__global__ void __launch_bounds__(256) sum_t16_kernel(const tup16* __restrict__ a, u64 n, u64* __restrict__ dst) {
  u64 s = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) s += a[i].x + a[i].y;
  u64 v[1] = {s};
  u64* const d[1] = {dst};
  block_sum_atomic<1, 256>(v, d);
}
__global__ void __launch_bounds__(256) sum_ovf16_kernel(const tup16* __restrict__ ovf_list, const u64* __restrict__ ctr,
                                                         u64* __restrict__ dst) {
  const u64 n = ctr[C_CONFLICTS];
  u64 s = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x)
    s += ovf_list[i].x + ovf_list[i].y;
  u64 v[1] = {s};
  u64* const d[1] = {dst};
  block_sum_atomic<1, 256>(v, d);
}

}  // namespace hj
