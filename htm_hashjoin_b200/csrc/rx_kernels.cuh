// rx_kernels.cuh -- radix join: histogram, scan, scatter, per-partition join.
//
// Replaces mc/src/parallel_radix_join.c: parallel_radix_partition (:559-627,
// histogram :586-589, prefix :592-617, scatter :622-626), its software
// write-combining variant (:701-795), serial_radix_partition/radix_cluster
// (:516-549, :402-440) for pass 2, and the per-partition joins
// bucket_chaining_join (:231-283) / histogram_join (:306-369).
//
// B200 shape of the same steps:
//   rx_hist     ONE read of the relation builds the histogram over ALL radix
//               bits of all passes (shared-memory privatised, merged with
//               global atomics) -- the reference re-reads per pass.
//   rx_scan     one CTA: exclusive offsets of every final partition; the
//               pass-1 partitions are every F2-th entry of the same array.
//   rx_scatter  a CTA stages a tile of tuples in shared memory, ranks them by
//               digit there, claims one global range per (tile, digit) with a
//               single atomicAdd and writes each digit's run contiguously:
//               the write-combining buffer is the digit-sorted tile, flushed
//               as runs of T/F tuples (full 128 B lines at the default fan-out).
//   rx_join     one CTA per final partition: open-addressing table of R_p in
//               shared memory, S_p streamed through it; R partitions larger
//               than the table are processed in chunks.
#pragma once
#include "hj_common.cuh"

namespace hj {

// How a key maps to a bin.
//   mode 0: (key >> shift) & (2^bits - 1)                    one radix pass
//   mode 1: p1-major index over two passes: ((key & (F1-1)) << b2) | ((key >> b1) & (F2-1))
//           (pass 1 = low b1 bits, pass 2 = next b2 bits: prj_params.h:76-78,
//            parallel_radix_join.c:870-871,960-961)
//   mode 2: range split for the G-way exchange: min(G-1, ((key-1)*G) >> domain_bits)
//   mode 3: exchange fused with radix pass 1: (dest << b1) | (key & (F1-1)), dest as in mode 2
struct Digit {
  u32 mode, shift, bits, b1, b2, ndest, domain_bits, pad;
  u64 key_mask;
};
__device__ __forceinline__ u32 digit_of(const Digit& d, u64 v) {
  const u64 key = v & d.key_mask;
  if (d.mode == 0) return (u32)((key >> d.shift) & ((1u << d.bits) - 1u));
  if (d.mode == 1) return (u32)(((key & ((1u << d.b1) - 1u)) << d.b2) | ((key >> d.b1) & ((1u << d.b2) - 1u)));
  u64 q = ((key - 1) * (u64)d.ndest) >> d.domain_bits;
  const u32 dest = (u32)(q < d.ndest ? q : d.ndest - 1);
  if (d.mode == 2) return dest;
  return (dest << d.b1) | ((u32)key & ((1u << d.b1) - 1u));
}

// ---- histogram ---------------------------------------------------------------
constexpr int kHistIPT = 8;  // threads per CTA: 512, or 1024 when the privatised histogram allows one CTA per SM only

// nbins*4 bytes of dynamic shared memory when SMEM, else straight global atomics
template <bool SMEM, int kHistThreads>
__global__ void __launch_bounds__(kHistThreads)
rx_hist_kernel(const u64* __restrict__ in, u64 n, Digit dg, u32 nbins, u64* __restrict__ ghist,
               u64* __restrict__ ctr) {
  extern __shared__ u32 shist[];
  const u32 tid = threadIdx.x;
  u32 key_hi = 0;  // key bits above bit 31, ORed over every tuple: picks the join kernel (rx_join32.cuh)
  u32 zeros = 0;   // tuples equal to 0 (the empty-slot sentinel) are ignored by the join: counted for the caller
  u32 pay_hi = 0;  // upper halves ORed, unmasked: 0 = nobody carries a payload (rx_join32 then never touches them)
  if (SMEM) {
    for (u32 i = tid; i < nbins; i += kHistThreads) shist[i] = 0;
    __syncthreads();
  }
  constexpr int TILE = kHistThreads * kHistIPT;
  const u64 ntiles = (n + TILE - 1) / TILE;
  for (u64 t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const u64 base = t * TILE;
    u64 v[kHistIPT];
#pragma unroll
    for (int j = 0; j < kHistIPT / 2; j++) {
      const u64 i = base + 2ull * (j * kHistThreads + tid);
      if (i + 1 < n) {
        ulonglong2 x = ld_stream_v2(in + i);
        v[2 * j] = x.x;
        v[2 * j + 1] = x.y;
        zeros += (x.x == 0ull) + (x.y == 0ull);
      } else {
        v[2 * j] = i < n ? ld_stream(in + i) : 0ull;
        v[2 * j + 1] = 0ull;
        zeros += i < n && v[2 * j] == 0ull;
      }
    }
#pragma unroll
    for (int j = 0; j < kHistIPT; j++) {
      const bool valid = v[j] != 0ull;
      key_hi |= (u32)((v[j] & dg.key_mask) >> 32);
      pay_hi |= (u32)(v[j] >> 32);
      const u32 d = valid ? digit_of(dg, v[j]) : 0xffffffffu;
      // sorted input on high bits puts a whole warp in one bin: add once
      const u32 d0 = __shfl_sync(0xffffffffu, d, 0);
      if (__all_sync(0xffffffffu, d == d0)) {
        if (lane_id() == 0 && valid) {
          if (SMEM) atomicAdd(shist + d, 32u);
          else atomicAdd(ghist + d, 32ull);
        }
      } else if (valid) {
        if (SMEM) atomicAdd(shist + d, 1u);
        else atomicAdd(ghist + d, 1ull);
      }
    }
  }
  if (SMEM) {
    __syncthreads();
    for (u32 i = tid; i < nbins; i += kHistThreads) {
      const u32 c = shist[i];
      if (c) atomicAdd(ghist + i, (u64)c);
    }
  }
  key_hi = __reduce_or_sync(0xffffffffu, key_hi);
  if (key_hi && lane_id() == 0) atomicOr(ctr + C_KEY_HI, (u64)key_hi);
  pay_hi = __reduce_or_sync(0xffffffffu, pay_hi);
  if (pay_hi && lane_id() == 0) atomicOr(ctr + C_PAY_HI, (u64)pay_hi);
  zeros = __reduce_add_sync(0xffffffffu, zeros);
  if (zeros && lane_id() == 0) atomicAdd(ctr + C_ZERO_KEYS, (u64)zeros);
}

// ---- scan ----------------------------------------------------------------------
// One CTA.  offsets[0..nbins] = exclusive prefix of ghist; cursors[i] = offsets[i]
// (claimed by the final scatter pass).  When f2 > 1 also emits the pass-1 view:
// parent_off[p1] = offsets[p1*f2] (nparent = nbins/f2 entries + 1),
// parent_cursor[p1] = parent_off[p1], and for both passes the tile prefix
// tiles[p] = number of scatter tiles of `tile` tuples before parent p.
// one[0..3] = the single-parent view {0, n} / {0, ceil(n/tile)} used by pass 1.
constexpr int kScanThreads = 1024;
__global__ void __launch_bounds__(kScanThreads)
rx_scan_kernel(const u64* __restrict__ ghist, u32 nbins, u64* __restrict__ offsets, u64* __restrict__ cursors,
               u32 f2, u64* __restrict__ parent_off, u64* __restrict__ parent_cursor, u64* __restrict__ parent_tiles,
               u64* __restrict__ one, u32 tile, u64* __restrict__ max_bin, u64 n_input) {
  // n_input: tuples in the relation the first pass READS.  It can exceed the histogram's total: tuples equal to 0
  // (the empty-slot sentinel) are counted nowhere, but the pass that reads the raw input must still walk over them.
  __shared__ u64 wsum[kScanThreads / 32];
  __shared__ u64 carry_s;
  const u32 tid = threadIdx.x;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  u64 mx = 0;
  for (u32 base = 0; base < nbins; base += kScanThreads) {
    const u32 i = base + tid;
    const u64 x = i < nbins ? ghist[i] : 0ull;
    mx = x > mx ? x : mx;
    u64 incl = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const u64 y = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane_id() >= o) incl += y;
    }
    if (lane_id() == 31) wsum[tid >> 5] = incl;
    __syncthreads();
    if (tid < 32) {
      u64 w = wsum[tid];
      u64 wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const u64 y = __shfl_up_sync(0xffffffffu, wi, o);
        if ((int)tid >= o) wi += y;
      }
      wsum[tid] = wi - w;  // exclusive warp offsets
    }
    __syncthreads();
    const u64 carry = carry_s;
    const u64 excl = carry + wsum[tid >> 5] + incl - x;
    if (i < nbins) {
      offsets[i] = excl;
      cursors[i] = excl;
    }
    __syncthreads();
    if (tid == kScanThreads - 1) carry_s = excl + x;
    __syncthreads();
  }
  if (tid == 0) {
    offsets[nbins] = carry_s;
    if (one) {  // one-parent view of the whole relation: offsets {0, n}, tiles {0, ceil(n/tile)}
      one[0] = 0; one[1] = n_input; one[2] = 0; one[3] = (n_input + tile - 1) / tile;
    }
  }
  if (max_bin) {
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    if (lane_id() == 0) atomicMax(max_bin, mx);
  }
  __syncthreads();
  // parent (pass-1) view and tile prefixes; nparent <= 4096, serial per thread chunk is fine
  if (parent_off) {
    const u32 nparent = nbins / f2;
    for (u32 p = tid; p <= nparent; p += kScanThreads) {
      u64 o = offsets[p * f2];  // p == nparent reads offsets[nbins]
      if (nparent == 1 && p == 1) o = n_input;  // single pass: the one parent is the raw input
      parent_off[p] = o;
      if (p < nparent) parent_cursor[p] = o;
    }
    __syncthreads();
    if (tid == 0) {
      u64 t = 0;
      for (u32 p = 0; p < nparent; p++) {
        parent_tiles[p] = t;
        t += (parent_off[p + 1] - parent_off[p] + tile - 1) / tile;
      }
      parent_tiles[nparent] = t;
    }
  }
}

// ---- scatter ---------------------------------------------------------------------
// A launch re-partitions every "parent" range [parent_off[p], parent_off[p+1]) of
// `in` by the digit `dg` into `out`, claiming space from cursors[p*F + digit].
// Pass 1 / single pass: one parent covering the whole relation.
// Dynamic shared memory: tile (TILE*8) | gdelta (F*8) | hist (F*4) | lbase (F*4).
constexpr int kScThreads = 512;
constexpr int kScParentCache = 256;  // parents whose offset/tile tables are cached in shared memory

// MODE: compile-time digit function (0 = one radix pass on bits [shift, shift+bits),
// which must lie below bit 32; 2 = generic digit_of, used for the G-way range split).
// 3 = exchange digit for 32-bit keys, a power-of-two number of ranks and a key domain of 2^domain_bits:
//     (min(G-1, (key-1) >> shift) << b1) | (key & (2^b1 - 1)) with shift = domain_bits - log2 G (host) -- digit_of's
//     mode 3 without the 64-bit multiply (the local scatter of the pipelined exchange).
template <int MODE>
__device__ __forceinline__ u32 sc_digit(const Digit& dg, u32 fmask, u64 v) {
  if (MODE == 0) return (u32)(v >> dg.shift) & fmask;
  if (MODE == 3) {
    const u32 k = (u32)v;
    return (min((k - 1u) >> dg.shift, dg.ndest - 1u) << dg.b1) | (k & ((1u << dg.b1) - 1u));
  }
  return digit_of(dg, v);
}

// PEER: the run of digit d goes to peer_out[d >> dg.b1] (the pass-1 buffer of the GPU that owns
// the key range, mapped through CUDA IPC: plain stores over NVLink) instead of `out`; cursors[d]
// then holds this rank's own, exclusive range inside that buffer, so the claims stay local atomics.
constexpr int kScMaxPeers = 16;

// OPT ("optimistic", histogram-free partitioning): the output is not laid out from exact
// histogram offsets but in fixed-capacity regions, region q = [q*region_cap, (q+1)*region_cap)
// with q = parent*F + digit, claimed from cursors[q] (initialised to q*region_cap); the regions'
// fill levels (the cursors) are what the next stage reads.  A claim that would run over its
// region is diverted to a scratch area at trash_off and sets ctr[C_OPT_OVF]: the host then redoes
// the join on the exact, histogram-based path.  Saves the full read of both relations that
// rx_hist costs (8 of the 48 B/tuple the 2-pass join moves).  Parents are [parent_off[p],
// parent_end[p]) (gaps between regions), and since rx_hist is gone the key-width flag
// ctr[C_KEY_HI] is gathered here.
// low_mask (!PEER): parents that share an output partition by their LOW bits (source-major blocks of the
// pipelined exchange, hj_xchg_*): cursor index ((parent >> parent_shift) & low_mask) * F + digit.
// parent_shift: parents that share an output partition -- after a histogram-free EXCHANGE the pass-1
// partition `low` arrives as one region per source rank, parent = low * n_src + src, so the cursor
// index is (parent >> parent_shift) * F + digit.  PEER && OPT (the sending side of that exchange):
// digit = (dest << b1) | low goes to region low * ndest + rank of dest's buffer; cursors[digit] is
// initialised to that region's start by the caller and stays local, the final values are what the
// receivers are told.
struct ScOpt {
  const u64* parent_end;
  u64 region_cap, trash_off;
  u64* ctr;
  u32 parent_shift, low_mask, ndest, rank;
};

// OPT: 0 = exact layout, 1 = histogram-free + key-width flag (pass 1), 2 = histogram-free (pass 2)
// ---- TMA (bulk async copy) input staging: mbarrier + cp.async.bulk global -> shared ----------
__device__ __forceinline__ void mbar_init(u32 a, u32 count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u32 a, u32 bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(u32 dst, const void* src, u32 bytes, u32 mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(mbar)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(u32 a, u32 parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}" ::"r"(a),
      "r"(parity)
      : "memory");
}

// PEER launches use big tiles (IPT = 32: 16384 tuples, 128 KB) and one CTA per SM: NVLink, not the SM,
// bounds them (same time with 1 or 3 CTAs/SM), and 4x longer runs mean 4x larger NVLink writes.
// BULK (with PEER): a run leaves through ONE bulk async copy (cp.async.bulk shared -> global, the
// TMA engine) instead of 8-byte stores from 32 lanes.  Bulk copies need 16-byte aligned source,
// destination and size: every digit gets one spare slot in the shared-memory tile so its run can
// start on the parity of its global offset; an odd head / tail tuple goes out as a scalar store.
// TMAIN: full, 16-byte aligned input tiles are fetched by ONE bulk async copy (TMA) into a
// shared-memory staging buffer, issued a whole tile ahead (right after the scan of the previous tile,
// when every thread holds its tuples in registers): no registers are tied up by loads in flight and
// the copy has the placement and the flush to complete.  Measured at C3: 18 % SLOWER than the
// register loads (profiles/r01_scatter_tmain_vs_regs.jsonl) -- the extra shared-memory write + read
// per tuple costs more than the exposed load latency it removes; kept opt-in (HJ_SC_TMAIN=1).
template <int IPT, int MODE, bool PEER = false, int OPT = 0, bool BULK = false, bool TMAIN = false>
__global__ void __launch_bounds__(kScThreads, IPT >= 32 ? 1 : (PEER || IPT >= 16) ? 2 : 3)
rx_scatter_kernel(const u64* __restrict__ in, u64* __restrict__ out, Digit dg, u32 F, u32 nparent,
                  const u64* __restrict__ parent_off, const u64* __restrict__ parent_tiles,
                  u64* __restrict__ cursors, u64* const* __restrict__ peer_out, ScOpt opt) {
  __shared__ u64* s_peer[PEER ? kScMaxPeers : 1];
  if (PEER && threadIdx.x < dg.ndest) s_peer[threadIdx.x] = peer_out[threadIdx.x];
  __shared__ u64 s_pend[OPT ? kScParentCache + 1 : 1];
  u32 key_hi = 0;  // OPT == 1 (the pass that sees the input first, no rx_hist): key width (and zero tuples, below)
  constexpr int TILE = kScThreads * IPT;
  static_assert(TILE <= 65536, "rank is packed into 16 bits");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  u64* inbuf = reinterpret_cast<u64*>(smem_raw);  // TMAIN: staging buffer of the next input tile
  u64* tile = inbuf + (TMAIN ? TILE : 0);
  u64* gdelta = tile + TILE + (BULK ? F + 2 : 0);  // BULK: one spare slot per digit (F even)
  __shared__ __align__(8) u64 s_mbar;
  const u32 mbar = (u32)__cvta_generic_to_shared(&s_mbar);
  if (TMAIN && threadIdx.x == 0) {
    mbar_init(mbar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  u32* hist = reinterpret_cast<u32*>(gdelta + F);
  u32* lbase = hist + F;
  u32* rcnt = lbase + F;  // BULK only: run lengths of the tile being flushed
  __shared__ u32 warp_tot[kScThreads / 32];

  // the tile -> parent search would be a chain of dependent global loads per tile:
  // keep both small tables in shared memory when they fit
  __shared__ u64 s_ptiles[kScParentCache + 1], s_poff[kScParentCache + 1];

  const u32 tid = threadIdx.x;
  const u32 fmask = F - 1;
  const bool cached = nparent <= kScParentCache;
  if (cached) {
    for (u32 i = tid; i <= nparent; i += kScThreads) {
      s_ptiles[i] = parent_tiles[i];
      s_poff[i] = parent_off[i];
    }
  }
  if (OPT)
    for (u32 i = tid; i < nparent; i += kScThreads) s_pend[i] = opt.parent_end[i];
  for (u32 i = tid; i < F; i += kScThreads) hist[i] = 0;
  __syncthreads();
  const u64* pt = cached ? s_ptiles : parent_tiles;
  const u64* po = cached ? s_poff : parent_off;
  const u32 total_tiles = (u32)pt[nparent];  // < 2^32 tiles of 4096 tuples
  const u32 per_thread = (F + kScThreads - 1) / kScThreads;  // bins each thread scans

  // locate: tile t -> (parent, first element, length).  fetch: its loads, either register loads
  // (issued right after tile k has been moved to shared memory, so they fly while tile k is flushed)
  // or, TMAIN, one bulk copy issued by thread 0 a whole phase earlier.
  u64 v[IPT];
  u64 begin = 0;
  auto locate = [&](u32 t, u32& p_out, u32& cnt_out) {
    u32 lo = 0, hi = nparent;  // last p with parent_tiles[p] <= t
    while (hi - lo > 1) {
      const u32 mid = (lo + hi) >> 1;
      if (pt[mid] <= t) lo = mid; else hi = mid;
    }
    begin = po[lo] + (t - pt[lo]) * TILE;
    const u64 pend = OPT ? s_pend[lo] : po[lo + 1];  // OPT: nparent <= kScParentCache (host)
    cnt_out = (u32)(pend - begin < (u64)TILE ? pend - begin : (u64)TILE);
    p_out = lo;
  };
  auto fetch_regs = [&](u32 cnt) {
    const u64* src = in + begin;
    if (cnt == TILE && (begin & 1ull) == 0) {  // full, 16 B aligned tile: no per-element checks
#pragma unroll
      for (int j = 0; j < IPT / 2; j++) {
        const ulonglong2 x = ld_stream_v2(src + 2u * (j * kScThreads + tid));
        v[2 * j] = x.x;
        v[2 * j + 1] = x.y;
      }
    } else {
#pragma unroll
      for (int j = 0; j < IPT / 2; j++) {
        const u32 i = 2u * (j * kScThreads + tid);
        v[2 * j] = i < cnt ? ld_stream(src + i) : 0ull;
        v[2 * j + 1] = i + 1 < cnt ? ld_stream(src + i + 1) : 0ull;
      }
    }
  };
  auto tma_ok = [&](u32 cnt) { return TMAIN && cnt == TILE && (begin & 1ull) == 0; };
  auto fetch_tma = [&]() {
    if (tid == 0) {
      mbar_expect_tx(mbar, TILE * 8u);
      tma_load_1d((u32)__cvta_generic_to_shared(inbuf), in + begin, TILE * 8u, mbar);
    }
  };

  u32 t = blockIdx.x;
  u32 p = 0, cnt = 0, ph = 0;
  bool via_tma = false;
  if (t < total_tiles) {
    locate(t, p, cnt);
    via_tma = tma_ok(cnt);
    if (via_tma) fetch_tma(); else fetch_regs(cnt);
  }
  while (t < total_tiles) {
    if (TMAIN && via_tma) {
      mbar_wait(mbar, ph);
      ph ^= 1u;
#pragma unroll
      for (int j = 0; j < IPT / 2; j++) {
        const ulonglong2 x = *reinterpret_cast<const ulonglong2*>(inbuf + 2u * (j * kScThreads + tid));
        v[2 * j] = x.x;
        v[2 * j + 1] = x.y;
      }
    }
    u32 dr[IPT];  // digit << 16 | rank  (rank < TILE <= 65536, digit < 4096)
#pragma unroll
    for (int j = 0; j < IPT; j++) {
      if (v[j] != 0ull) {
        const u32 d = sc_digit<MODE>(dg, fmask, v[j]);
        dr[j] = (d << 16) | atomicAdd(hist + d, 1u);
        if (OPT == 1) key_hi |= (u32)(v[j] >> 32);
      }
    }
    // BULK: the previous tile's bulk copies have finished READING shared memory
    if (BULK) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    __syncthreads();  // also: every thread has finished flushing the previous tile

    // exclusive scan of hist -> lbase; one global claim per non-empty digit
    {
      const u32 b0 = tid * per_thread;
      u32 local = 0;
      for (u32 k = 0; k < per_thread; k++) {
        const u32 b = b0 + k;
        if (b < F) local += hist[b];
      }
      u32 incl = local;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const u32 y = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)lane_id() >= o) incl += y;
      }
      if (lane_id() == 31) warp_tot[tid >> 5] = incl;
      __syncthreads();
      if (tid < 32) {
        const u32 w = tid < kScThreads / 32 ? warp_tot[tid] : 0u;
        u32 wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const u32 y = __shfl_up_sync(0xffffffffu, wi, o);
          if ((int)tid >= o) wi += y;
        }
        if (tid < kScThreads / 32) warp_tot[tid] = wi - w;
      }
      __syncthreads();
      u32 run = warp_tot[tid >> 5] + incl - local;
      for (u32 k = 0; k < per_thread; k++) {
        const u32 b = b0 + k;
        if (b < F) {
          const u32 c = hist[b];
          hist[b] = 0;  // ready for the next tile (nobody reads hist after this scan)
          lbase[b] = run;
          if (c) {
            // (!PEER: low_mask, when set, keeps only the low bits of the parent index -- the blocks of the
            // pipelined exchange are source-major: parent = src * F1 + low, partition = low)
            u32 pq = OPT ? p >> opt.parent_shift : p;
            if (OPT && !PEER && opt.low_mask) pq &= opt.low_mask;
            const u64 q = (u64)pq * F + b;
            u64 g = atomicAdd(cursors + q, (u64)c);
            const u64 rend = (PEER && OPT) ? ((u64)(b & opt.low_mask) * opt.ndest + opt.rank + 1) * opt.region_cap
                                           : (q + 1) * opt.region_cap;
            if (OPT && g + c > rend) {  // region full: divert, flag, host falls back
              g = opt.trash_off;
              opt.ctr[C_OPT_OVF] = 1ull;
            }
            if (BULK) {
              const u32 sm = run + b + (((run + b) ^ (u32)g) & 1u);  // same parity as the global offset
              lbase[b] = sm;
              gdelta[b] = g;
            } else {
              gdelta[b] = g - run;
            }
          }
          if (BULK) rcnt[b] = c;
          run += c;
          // tuples equal to 0 (the empty-slot sentinel) were not ranked: what the tile's histogram lacks of its
          // length is their number (reported, hj_result.zero_keys; no per-thread counter: registers are full)
          if (OPT == 1 && b == F - 1 && run != cnt) atomicAdd(opt.ctr + C_ZERO_KEYS, (u64)(cnt - run));
        }
      }
    }
    __syncthreads();

    // the scan is done with p: locate the next tile; TMAIN: its bulk copy starts now (the staging
    // buffer is free: every thread passed the barrier above with its tuples in registers)
    const u32 cnt_cur = cnt;
    t += gridDim.x;
    if (TMAIN && t < total_tiles) {
      locate(t, p, cnt);
      via_tma = tma_ok(cnt);
      if (via_tma) fetch_tma();
    }

#pragma unroll
    for (int j = 0; j < IPT; j++)
      if (v[j] != 0ull) tile[lbase[dr[j] >> 16] + (dr[j] & 0xffffu)] = v[j];

    // v is dead: start the next tile's register loads before flushing this one.  (Without TMAIN the
    // tile is located only now: locating it before the placement keeps three more registers alive
    // across it, which spilled under the 40-register cap and cost 25 % in the 512-way pass.)
    if (t < total_tiles) {
      if (!TMAIN) locate(t, p, cnt);
      if (!TMAIN || !via_tma) fetch_regs(cnt);
    }
    __syncthreads();

    if (BULK) {
      // flush: one thread per run; generic-proxy writes of the tile made visible to the async proxy
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      for (u32 b = tid; b < F; b += kScThreads) {
        u32 c = rcnt[b];
        if (c == 0) continue;
        u32 sm = lbase[b];
        u64 g = gdelta[b];
        u64* o = PEER ? s_peer[b >> dg.b1] : out;
        if (g & 1ull) { o[g] = tile[sm]; ++sm; ++g; --c; }
        const u32 n2 = c & ~1u;
        if (n2) {
          const u32 src = (u32)__cvta_generic_to_shared(tile + sm);
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(o + g), "r"(src),
                       "r"(n2 * 8u)
                       : "memory");
        }
        if (c & 1u) o[g + n2] = tile[sm + n2];
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    } else
    // flush: element i of the digit-sorted tile goes to gdelta[digit] + i
    if (cnt_cur == TILE) {
#pragma unroll
      for (int j = 0; j < IPT; j++) {
        const u32 i = j * kScThreads + tid;
        const u64 x = tile[i];
        const u32 d = sc_digit<MODE>(dg, fmask, x);
        u64* o = PEER ? s_peer[d >> dg.b1] : out;
        o[gdelta[d] + i] = x;
      }
    } else {
      for (u32 i = tid; i < cnt_cur; i += kScThreads) {
        const u64 x = tile[i];
        const u32 d = sc_digit<MODE>(dg, fmask, x);
        u64* o = PEER ? s_peer[d >> dg.b1] : out;
        o[gdelta[d] + i] = x;
      }
    }
    // no barrier here: the one after the next tile's ranking orders this flush before the
    // next writes to lbase/gdelta/tile
  }
  if (BULK) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  if (OPT == 1) {
    key_hi = __reduce_or_sync(0xffffffffu, key_hi);  // (raw upper halves so far)
    if (key_hi && lane_id() == 0) atomicOr(opt.ctr + C_PAY_HI, (u64)key_hi);
    key_hi &= (u32)(dg.key_mask >> 32);
    if (key_hi && lane_id() == 0) atomicOr(opt.ctr + C_KEY_HI, (u64)key_hi);
  }
}

// One pass into more partitions than a shared-memory tile can hold a run for (2^13 .. 2^15 digits;
// the reference builds NUM_RADIX_BITS=14, NUM_PASSES=1 with the plain scatter of
// parallel_radix_join.c:622-626): the same loop, one global claim per tuple.  Equal digits inside a
// warp (sorted or skewed inputs) are claimed together.
__global__ void __launch_bounds__(256) rx_scatter_direct_kernel(const u64* __restrict__ in, u64 n,
                                                                 u64* __restrict__ out, Digit dg,
                                                                 u64* __restrict__ cursors) {
  const u64 stride = (u64)gridDim.x * blockDim.x;
  const u64 rounds = (n + stride - 1) / stride;  // all lanes stay in the loop: the claim is a warp collective
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  for (u64 r = 0; r < rounds; r++, i += stride) {
    const u64 v = i < n ? ld_stream(in + i) : 0ull;
    const u32 d = v ? digit_of(dg, v) : 0xffffffffu - lane_id();
    const u32 peers = __match_any_sync(0xffffffffu, d);
    const u32 leader = __ffs(peers) - 1;
    u64 base = 0;
    if (v && lane_id() == leader) base = atomicAdd(cursors + d, (u64)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (v) out[base + __popc(peers & ((1u << lane_id()) - 1u))] = v;
  }
}

// ---- optimistic (histogram-free) layout helpers -----------------------------------------------
// cursors[q] = q * cap for q < count
__global__ void __launch_bounds__(256) rx_opt_init_kernel(u64* __restrict__ a, u32 na, u64 capa, u64* __restrict__ b,
                                                           u32 nb, u64 capb) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < na) a[i] = (u64)i * capa;
  if (i < nb) b[i] = (u64)i * capb;
}
// pass-1 output that arrived through the fused exchange: parent p = [off[p], off[p+1]) exactly
__global__ void __launch_bounds__(1024) rx_pre1_parents_kernel(const u64* __restrict__ off, u32 nparent, u32 tile,
                                                                u64* __restrict__ poff, u64* __restrict__ pend,
                                                                u64* __restrict__ ptiles) {
  for (u32 p = threadIdx.x; p <= nparent; p += blockDim.x) {
    poff[p] = off[p];
    if (p < nparent) pend[p] = off[p + 1];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    u64 t = 0;
    for (u32 p = 0; p < nparent; p++) {
      ptiles[p] = t;
      t += (pend[p] - poff[p] + tile - 1) / tile;
    }
    ptiles[nparent] = t;
  }
}
// after pass 1: parent p = [p*cap, cur[p]); tiles[p] = scatter tiles before parent p (one CTA)
__global__ void __launch_bounds__(1024) rx_opt_parents_kernel(const u64* __restrict__ cur, u32 nparent, u64 cap,
                                                               u32 tile, u64* __restrict__ poff,
                                                               u64* __restrict__ pend, u64* __restrict__ ptiles) {
  for (u32 p = threadIdx.x; p < nparent; p += blockDim.x) {
    poff[p] = (u64)p * cap;
    const u64 e = cur[p], lim = (u64)(p + 1) * cap;
    pend[p] = e < lim ? e : lim;  // an overflowed region (the host falls back anyway) must not be read past its end
  }
  if (threadIdx.x == 0) poff[nparent] = (u64)nparent * cap;
  __syncthreads();
  if (threadIdx.x == 0) {
    u64 t = 0;
    for (u32 p = 0; p < nparent; p++) {
      ptiles[p] = t;
      t += (pend[p] - poff[p] + tile - 1) / tile;
    }
    ptiles[nparent] = t;
  }
}


// ---- pipelined exchange over the copy engines (hj_xchg_*, hj_api.cu) --------------------------------
// Stage t (0 = R, 1.. = the chunks of S) of a rank's receive window: G source blocks of `stride` tuples,
// block src = F1 regions of cap tuples (region low = pass-1 partition `low` as scattered by rank src) + a
// scratch zone.  fill[src * F1 + low] = the sender's final cursor for that region, RELATIVE to its block.
// Parent p = src * F1 + low of pass 2 = [start, start + fill - low * cap), clamped to the region.
__global__ void __launch_bounds__(1024) rx_xchg_parents_kernel(const u64* __restrict__ fill, u32 nparent, u32 F1,
                                                                u64 cap, u64 stride, u32 tile, u64* __restrict__ poff,
                                                                u64* __restrict__ pend, u64* __restrict__ ptiles) {
  for (u32 p = threadIdx.x; p < nparent; p += blockDim.x) {
    const u32 src = p / F1, low = p - src * F1;
    const u64 start = (u64)src * stride + (u64)low * cap;
    u64 cnt = fill[p] - (u64)low * cap;
    if (fill[p] < (u64)low * cap || cnt > cap) cnt = cap;  // an overflowed region (flagged by its sender) is read up to its end
    poff[p] = start;
    pend[p] = start + cnt;
  }
  if (threadIdx.x == 0) poff[nparent] = (u64)(nparent / F1) * stride;
  __syncthreads();
  if (threadIdx.x == 0) {
    u64 t = 0;
    for (u32 p = 0; p < nparent; p++) {
      ptiles[p] = t;
      t += (pend[p] - poff[p] + tile - 1) / tile;
    }
    ptiles[nparent] = t;
  }
}
// cursors of every stage's local scatter: cur[t * gf + (dest << b1 | low)] = low * cap[t]; epoch word
__global__ void __launch_bounds__(256) rx_xchg_init_kernel(u64* __restrict__ cur, u32 nstages, u32 gf, u32 f1mask,
                                                            const u64* __restrict__ caps, u64* __restrict__ epoch_word,
                                                            u64 epoch) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nstages * gf) cur[i] = (u64)((i % gf) & f1mask) * caps[i / gf];
  if (i == 0) *epoch_word = epoch;
}
// Wait until every peer's flag for this stage has reached `epoch` (the peers' copy engines write it after the
// stage's data and fill levels).  One warp; gives up after timeout_ns and raises ctr[C_XCHG_TIMEOUT] so a lost
// peer can never hang the device.
__device__ __forceinline__ u64 ld_acquire_sys(const u64* p) {
  u64 r;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ u64 global_timer_ns() {
  u64 t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__global__ void __launch_bounds__(32) rx_xchg_wait_kernel(const u64* __restrict__ flags, u32 n, u32 self, u64 epoch,
                                                           u64* __restrict__ ctr, u64 timeout_ns) {
  const u32 i = threadIdx.x;
  if (i < n && i != self) {
    const u64 t0 = global_timer_ns();
    while (ld_acquire_sys(flags + i) < epoch) {
      if (global_timer_ns() - t0 > timeout_ns) {
        atomicExch(ctr + C_XCHG_TIMEOUT, 1ull);
        break;
      }
      __nanosleep(256);
    }
  }
  __syncwarp();
  __threadfence_system();
}

// ---- per-partition join --------------------------------------------------------
// Dynamic shared memory: slots * 8 bytes (slots = power of two).  R_p is inserted
// in chunks of slots/2 tuples so the load factor never exceeds 0.5; S_p is
// streamed once per chunk.  Hash = Fibonacci hash of the bits above the radix bits
// (the reference uses those bits directly: idx = (key & MASK) >> NUM_RADIX_BITS,
// parallel_radix_join.c:252,339).
// With S == nullptr this is the build-only PRO of the reference and the result
// is its "Results" checksum: sum of idx with N = NEXT_POW_2(|R_p|) (:239-257).
constexpr int kJoinThreads = 512;

// Fibonacci hashing of the key bits above the radix bits.  The reference indexes its
// per-partition table with those bits directly (parallel_radix_join.c:252); with dense
// keys that fills one contiguous run of slots, and a probe that must walk to the first
// empty slot (to count duplicates) would walk the whole run.  The golden-ratio multiply
// spreads consecutive keys evenly, so runs stay 1-2 slots long at load <= 0.5.
__device__ __forceinline__ u32 join_slot(u64 key, u32 radix_bits, u32 hshift) {
  // keys are 32-bit in every in-scope input; fold the (normally zero) upper half in so
  // wider keys still hash, then one 32-bit multiply
  const u32 k32 = (u32)(key >> radix_bits) ^ (u32)(key >> 32);
  return (k32 * 0x9E3779B1u) >> hshift;
}

template <int kJoinBatch, int MINB>
__global__ void __launch_bounds__(kJoinThreads, MINB)
rx_join_kernel(const u64* __restrict__ Rp, const u64* __restrict__ offR, const u64* __restrict__ Sp,
               const u64* __restrict__ offS, u32 nparts, u32 radix_bits, u32 slots, u64 key_mask,
               u64* __restrict__ ctr, int narrow_keys_elsewhere, u64 capR, u64 capS, u32 nsrc, u64 src_stride,
               u64 off_stride) {
  // nsrc > 1: S arrived in nsrc separately partitioned chunks (see rx_join32_kernel)
  // launched behind rx_join32_kernel: exactly one of the two does the work
  if (narrow_keys_elsewhere && ctr[C_KEY_HI] == 0ull) return;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  u64* tab = reinterpret_cast<u64*>(smem_raw);
  const u32 tid = threadIdx.x;
  const u32 smask = slots - 1;
  const u32 hshift = 32 - (31 - __clz(slots));  // slots is a power of two
  const u32 chunk_cap = slots / 2;
  u64 matches = 0, checksum = 0, maxpart = 0, chunked = 0;

  for (u32 p = blockIdx.x; p < nparts; p += gridDim.x) {
    // exact layout: partition p = [off[p], off[p+1]); histogram-free layout (cap != 0):
    // [p*cap, min(off[p], (p+1)*cap)) with off = the scatter's final cursors
    const u64 r0 = capR ? p * capR : offR[p];
    const u64 r1 = capR ? (offR[p] < r0 + capR ? offR[p] : r0 + capR) : offR[p + 1];
    const u64 nr = r1 - r0;
    maxpart = nr > maxpart ? nr : maxpart;
    if (Sp == nullptr) {
      // build-only: reference checksum (bucket_chaining_join)
      u32 N = 1;
      while (N < nr) N <<= 1;
      const u32 msk = (N - 1) << radix_bits;
      for (u64 i = r0 + tid; i < r1; i += kJoinThreads) checksum += ((u32)Rp[i] & msk) >> radix_bits;
      continue;
    }
    if (nr == 0) continue;
    if (nr > chunk_cap) chunked++;
    for (u64 c0 = r0; c0 < r1; c0 += chunk_cap) {
      const u64 c1 = c0 + chunk_cap < r1 ? c0 + chunk_cap : r1;
      __syncthreads();  // previous probe finished before the table is cleared
      for (u32 i = tid; i < slots; i += kJoinThreads) tab[i] = 0ull;
      __syncthreads();
      // build: kJoinBatch loads in flight per thread, then the CAS inserts
      for (u64 b = c0; b < c1; b += (u64)kJoinThreads * kJoinBatch) {
        u64 rv[kJoinBatch];
#pragma unroll
        for (int j = 0; j < kJoinBatch; j++) {
          const u64 i = b + (u64)j * kJoinThreads + tid;
          rv[j] = i < c1 ? ld_stream(Rp + i) : 0ull;
        }
#pragma unroll
        for (int j = 0; j < kJoinBatch; j++) {
          if (rv[j] == 0ull) continue;
          u32 slot = join_slot(rv[j] & key_mask, radix_bits, hshift);
          while (atomicCAS(tab + slot, 0ull, rv[j]) != 0ull) slot = (slot + 1) & smask;
        }
      }
      __syncthreads();
      // probe: a batch of S loads in flight, then the home slots of the batch, then the walks
      u32 matches32 = 0;  // < 2^32 per thread per chunk
      for (u32 src = 0; src < nsrc; src++) {
      const u64* __restrict__ Sp_ = Sp + (u64)src * src_stride;
      const u64* __restrict__ offS_ = offS + (u64)src * off_stride;
      const u64 s0 = capS ? p * capS : offS_[p];
      const u64 s1 = capS ? (offS_[p] < s0 + capS ? offS_[p] : s0 + capS) : offS_[p + 1];
      for (u64 b = s0; b < s1; b += (u64)kJoinThreads * kJoinBatch) {
        u64 sv[kJoinBatch], cur[kJoinBatch];
        u32 sl[kJoinBatch];
#pragma unroll
        for (int j = 0; j < kJoinBatch; j++) {
          const u64 i = b + (u64)j * kJoinThreads + tid;
          sv[j] = i < s1 ? ld_stream(Sp_ + i) : 0ull;
        }
#pragma unroll
        for (int j = 0; j < kJoinBatch; j++) {
          sl[j] = join_slot(sv[j] & key_mask, radix_bits, hshift);
          cur[j] = sv[j] ? tab[sl[j]] : 0ull;
        }
#pragma unroll
        for (int j = 0; j < kJoinBatch; j++) {
          const u64 key = sv[j] & key_mask;
          u64 c = cur[j];
          u32 slot = sl[j];
          while (c != 0ull) {
            if ((c & key_mask) == key) { ++matches32; checksum += c + sv[j]; }
            slot = (slot + 1) & smask;
            c = tab[slot];
          }
        }
      }
      }  // sources
      matches += matches32;
    }
  }
  u64 v[2] = {matches, checksum};
  u64* const d[2] = {ctr + C_MATCHES, ctr + C_CHECKSUM};
  block_sum_atomic<2, kJoinThreads>(v, d);
  // bookkeeping: largest partition, partitions that needed chunking
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

// ---- counter-based generators (hj_generate; CPU restatement in oracle/) ---------
// 4-round Feistel-style bijection on [0, 2^k): x -> perm(x).  For odd k the two
// halves differ by one bit; cycle-walking is avoided by using an unbalanced
// network whose round function folds into the narrower half.
__host__ __device__ __forceinline__ u64 mix64(u64 x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull;
  x ^= x >> 33;
  return x;
}
__host__ __device__ __forceinline__ u64 feistel_perm(u64 x, u32 k, u64 seed) {
  const u32 hb = k / 2, lb = k - hb;  // high half hb bits, low half lb bits (lb >= hb)
  u64 l = x & ((1ull << lb) - 1), h = x >> lb;
  if (k < 2) return x;
#pragma unroll
  for (u32 r = 0; r < 4; r++) {
    // alternate which half is rewritten; each step is invertible (xor with a function of the other half)
    h ^= mix64(l + seed + r * 0x9E3779B97F4A7C15ull) & ((1ull << hb) - 1);
    l ^= mix64(h + (seed ^ 0xD1B54A32D192ED03ull) + r * 0xBF58476D1CE4E5B9ull) & ((1ull << lb) - 1);
  }
  return (h << lb) | l;
}
__global__ void __launch_bounds__(256) gen_kernel(int kind, u64 first, u64 count, u32 kbits, u64 seed,
                                                   u64* __restrict__ out) {
  const u64 dmask = (1ull << kbits) - 1;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (u64)gridDim.x * blockDim.x) {
    const u64 g = first + i;
    out[i] = kind == 0 ? g + 1 : feistel_perm(g & dmask, kbits, seed + (g >> kbits)) + 1;
  }
}

}  // namespace hj
