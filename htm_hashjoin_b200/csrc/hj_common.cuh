// hj_common.cuh -- shared device helpers for the sm_100a join kernels.
//
// Everything on this path is 64-bit integer work bound by HBM / L2 traffic
// (SURVEY.md 8d); there is no dense contraction, so no tensor-core code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

typedef unsigned long long u64;
typedef unsigned int u32;

namespace hj {

// device-side counter block (one per handle, u64 each)
enum Counter {
  C_CONFLICTS = 0,   // tuples appended to the overflow list (AtomicHashBuild.hpp:62-63)
  C_CAS_RETRIES,     // lost CAS races
  C_PROBE_STEPS,     // occupied slots stepped over while inserting
  C_SECTORS,         // distinct 32 B table sectors per warp, summed (sample kernel)
  C_SAMPLED,         // tuples processed by the sample kernel
  C_MATCHES,         // true-join matches
  C_CHECKSUM,        // sum over pairs (r + s)
  C_WINDOW_MATCHES,  // matches found inside the probe window (reference totalMatches)
  C_SUM0, C_SUM1, C_SUM2,  // scratch for sum_u64
  C_MAX_PART,        // largest R partition (radix join)
  C_TOO_BIG,         // R partitions processed in >1 shared-memory chunk
  C_SPAN,            // sum over sampled CTA chunks of (max home - min home + 1), in slots
  C_CHUNKS,          // number of sampled CTA chunks
  C_KEY_HI,          // OR of the key bits above bit 31 over every tuple rx_hist saw (0: rx_join32 applies)
  C_OPT_OVF,         // a region of the histogram-free layout overflowed: redo on the exact path
  C_TILE0, C_TILE1,  // tile tickets of the L2 join kernels (rx_joinl2.cuh)
  C_NTASKS,          // join tasks = (partition, S slice) pairs built by rx_tasks_kernel (skew handling)
  C_MAT_TOTAL,       // pairs the materialising join will write (rx_joinmat.cuh, after the count pass)
  C_XCHG_TIMEOUT,    // a peer's stage flag did not arrive in time (pipelined exchange, rx_xchg_wait_kernel)
  C_ZERO_KEYS,       // input tuples equal to 0 (the empty-slot sentinel): ignored by the join, reported here
  C_PAY_HI,          // OR of the upper 32 bits of every tuple (0: no tuple carries a payload, rx_join32 skips them)
  C_COUNT = 24
};

constexpr u64 kOvfMul = 0x9E3779B97F4A7C15ull;  // Fibonacci hashing for the overflow table
constexpr u64 kOvfMinSlots = 1024;

__host__ __device__ __forceinline__ u64 pow2_ceil_u64(u64 v) {
  u64 p = 1;
  while (p < v) p <<= 1;
  return p;
}
// overflow-table geometry is a pure function of the overflow count so every
// kernel can recompute it from the device counter (no host round trip)
__host__ __device__ __forceinline__ u64 ovf_slots_for(u64 count) {
  u64 want = count * 2;
  return pow2_ceil_u64(want < kOvfMinSlots ? kOvfMinSlots : want);
}
__device__ __forceinline__ u64 ovf_hash(u64 key, u64 mask) { return ((key * kOvfMul) >> 20) & mask; }

// streaming 16-byte load of two tuples: read once, keep out of L1
__device__ __forceinline__ ulonglong2 ld_stream_v2(const u64* p) {
  ulonglong2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(r.x), "=l"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ u64 ld_stream(const u64* p) {
  u64 r;
  asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(r) : "l"(p));
  return r;
}
// table reads must observe other CTAs' CAS results: relaxed, gpu scope (served by L2)
__device__ __forceinline__ u64 ld_relaxed(const u64* p) {
  u64 r;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_stream(u64* p, u64 v) {
  asm volatile("st.global.L1::no_allocate.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_stream_v2(u64* p, u64 a, u64 b) {
  asm volatile("st.global.L1::no_allocate.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}

__device__ __forceinline__ u64 warp_sum(u64 v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ u32 lane_id() { return threadIdx.x & 31; }

// block-wide sum of up to 4 u64 values, one atomicAdd each from thread 0
template <int NV, int THREADS>
__device__ __forceinline__ void block_sum_atomic(u64 (&v)[NV], u64* const (&dst)[NV]) {
  __shared__ u64 red[NV][THREADS / 32];
#pragma unroll
  for (int i = 0; i < NV; i++) v[i] = warp_sum(v[i]);
  if (lane_id() == 0) {
#pragma unroll
    for (int i = 0; i < NV; i++) red[i][threadIdx.x >> 5] = v[i];
  }
  __syncthreads();
  if (threadIdx.x < 32) {
#pragma unroll
    for (int i = 0; i < NV; i++) {
      u64 x = threadIdx.x < THREADS / 32 ? red[i][threadIdx.x] : 0ull;
      x = warp_sum(x);
      if (threadIdx.x == 0 && x) atomicAdd(dst[i], x);
    }
  }
}

}  // namespace hj
