// hj_api.cu -- the C ABI of include/hj_b200.h on top of the sm_100a kernels.
//
// One handle owns one CUDA device: a compute stream, a copy stream, the global
// table, the overflow structures, the radix partition buffers and staging
// buffers for host inputs.  Buffers grow on demand and are reused, so -- like
// the reference, which allocates before gettimeofday (AtomicHashBuild.hpp:25-34)
// -- a warmed-up handle allocates nothing inside a timed call.
//
// There is deliberately no CPU path in this file: without a device every entry
// point fails with HJ_ECUDA.
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <new>

#include "../../include/hj_b200.h"
#include "gt_kernels.cuh"
#include "rx_kernels.cuh"
#include "rx_join32.cuh"
#include "rx_joinl2.cuh"
#include "rx_joinmat.cuh"
#include "rx_sortmerge.cuh"
#include "rx_wide.cuh"

using namespace hj;

namespace {

enum Ev { E_T0 = 0, E_CLEAR, E_SAMPLE, E_BUILD, E_PROBE0, E_PROBE, E_JOIN0, E_JOIN, E_END, E_CP0, E_CPR, E_CPS,
          E_RX,  // 5 events per relation: hist0, hist, scan, scatter1, scatter2
          E_COUNT = E_RX + 10 };

struct DevBuf {
  u64* p = nullptr;
  u64 cap = 0;  // in u64 elements
};

struct RelBufs {  // radix-side buffers for one relation
  DevBuf tmp, out, ghist, offsets, cursors, poff, pcur, ptiles, one, pend;
};

// A/B switches of the experiments under profiles/ (all default off), read when a handle is created:
// nothing on a timed path calls getenv.
struct Env {
  bool sc_no_bulk, sc_tmain, no_optimistic, force_l2, no_l2, no_join32, peer_no_bulk;
  int sc_ipt, sm_variant, sm_mode, sm_ctas, j32_variant, join_variant, peer_ipt, peer_ctas;
  u64 l2_min_need, join_slice;
};
long long env_num(const char* k, long long dflt) {
  const char* v = getenv(k);
  return v ? atoll(v) : dflt;
}
Env read_env() {
    auto num = [](const char* k, long long dflt) { const char* v = getenv(k); return v ? atoll(v) : dflt; };
    Env x;
    x.sc_no_bulk = getenv("HJ_SC_NO_BULK") != nullptr;
    x.sc_tmain = getenv("HJ_SC_TMAIN") != nullptr;
    x.no_optimistic = getenv("HJ_NO_OPTIMISTIC") != nullptr;
    x.force_l2 = getenv("HJ_FORCE_L2JOIN") != nullptr;
    x.no_l2 = getenv("HJ_NO_L2JOIN") != nullptr;
    x.no_join32 = getenv("HJ_NO_JOIN32") != nullptr;
    x.peer_no_bulk = getenv("HJ_PEER_NO_BULK") != nullptr;
    x.sc_ipt = (int)num("HJ_SC_IPT", 16);
    x.sm_variant = (int)num("HJ_SM_VARIANT", 1);
    x.sm_mode = (int)num("HJ_SM_MODE", -1);
    x.sm_ctas = (int)num("HJ_SM_CTAS", 2);
    x.j32_variant = (int)num("HJ_J32_VARIANT", -1);
    x.join_variant = (int)num("HJ_JOIN_VARIANT", 0);
    x.peer_ipt = (int)num("HJ_PEER_IPT", 32);
    x.peer_ctas = (int)num("HJ_PEER_CTAS", 0);
    x.l2_min_need = (u64)num("HJ_L2_MIN_NEED", 63000);
    x.join_slice = (u64)num("HJ_JOIN_SLICE", 0);
    return x;
}
Env g_env = read_env();  // refreshed by every hj_create
const Env& env() { return g_env; }

}  // namespace

struct Xchg;
struct hj_handle {
  hj_config cfg;
  int device = 0;
  int sm_count = 148;
  u64 l2_bytes = 126ull << 20;
  size_t smem_optin = 227 * 1024;
  cudaStream_t stream = nullptr, copy_stream = nullptr;
  cudaEvent_t ev[E_COUNT];
  bool ev_set[E_COUNT];
  u64* ctr = nullptr;    // device counters
  u64* h_ctr = nullptr;  // pinned mirror
  // global table
  DevBuf table, ovf_table, ovf_list;
  DevBuf claim;   // one bit per slot: the build's claims (gt_kernels.cuh)
  DevBuf ovfmap;  // one bit per home slot: some tuple of that home overflowed
  u64 slots = 0, mask = 0;
  bool built = false;
  bool no_optimistic = false;  // a histogram-free attempt overflowed on this handle: stay on the exact path
  // staging for host inputs
  DevBuf stageR, stageS;
  // radix
  RelBufs rb[2];
  DevBuf flush;
  DevBuf peer_ptrs;  // device copy of the peers' pass-1 buffer pointers (fused exchange)
  DevBuf tasks;      // join tasks (partition, S slice) of rx_join32 (skew handling)
  DevBuf mat_cnt;    // per-partition pair counts / offsets of the materialising join
  struct Xchg* xc = nullptr;  // pipelined exchange state (hj_xchg_*)
  u32 launches = 0;
  u64 allocs = 0;  // cudaMalloc calls made by ensure() over the handle's life (hj_alloc_count)
  char err[256];
};

namespace {

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      snprintf(h->err, sizeof(h->err), "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      return e_ == cudaErrorMemoryAllocation ? HJ_ENOMEM : HJ_ECUDA;                               \
    }                                                                                              \
  } while (0)

// cudaMalloc backs an allocation whose SIZE is not a multiple of 2 MiB with small pages throughout: a copy-engine
// write into such a buffer over NVLink runs at 530 GB/s instead of 760 (tools/py_copy_probe.py,
// profiles/r02_py_copy_probe.jsonl: a 9.78 GB window, 12 800 bytes past a 2 MiB multiple).  Every large allocation
// is therefore rounded up to whole 2 MiB pages.
inline u64 page_round_bytes(u64 bytes) {
  constexpr u64 kPage = 2ull << 20;
  return bytes >= kPage / 2 ? (bytes + kPage - 1) & ~(kPage - 1) : bytes;
}

int ensure(hj_handle* h, DevBuf& b, u64 elems) {
  if (b.cap >= elems) return HJ_OK;
  if (&b == &h->table) h->built = false;  // the old table goes away with its contents
  if (b.p) CK(cudaFree(b.p));
  b.p = nullptr;
  b.cap = 0;
  CK(cudaMalloc(&b.p, page_round_bytes(elems * sizeof(u64))));
  h->allocs++;
  b.cap = elems;
  return HJ_OK;
}

void mark(hj_handle* h, int e, cudaStream_t s = nullptr) {
  cudaEventRecord(h->ev[e], s ? s : h->stream);
  h->ev_set[e] = true;
}
float span_us(hj_handle* h, int a, int b) {
  if (!h->ev_set[a] || !h->ev_set[b]) return 0.f;
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, h->ev[a], h->ev[b]) != cudaSuccess) return 0.f;
  return ms * 1000.f;
}

bool is_device_ptr(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

int grid_for(hj_handle* h, const void* kernel, int threads, size_t smem) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem) != cudaSuccess || per_sm < 1) {
    cudaGetLastError();
    per_sm = 1;
  }
  return h->sm_count * per_sm;
}

u32 ipt_for(u32 tsize) {
  if (tsize == 0) return 4;
  if (tsize >= 8) return 8;
  if (tsize >= 4) return 4;
  if (tsize >= 2) return 2;
  return 1;
}

// ---- global table -------------------------------------------------------------

template <int IPT, bool STATS>
int launch_build_t(hj_handle* h, const u64* R, const SegMap& sm, u64 key_mask) {
  auto k = gt_build_kernel<IPT, STATS>;
  int grid = grid_for(h, (const void*)k, kGtThreads, 0);
  u64 chunks = ((sm.seg_len + kGtThreads * IPT - 1) / (kGtThreads * IPT)) * sm.nslices;
  if (chunks == 0) return HJ_OK;
  if ((u64)grid > chunks) grid = (int)chunks;
  k<<<grid, kGtThreads, 0, h->stream>>>(R, sm, h->table.p, reinterpret_cast<u32*>(h->claim.p), (u32)h->mask,
                                        (u32)(key_mask & h->mask), h->cfg.probe_length, h->ovf_list.p, h->ctr);
  h->launches++;
  CK(cudaGetLastError());
  return HJ_OK;
}
template <int IPT>
int launch_nocc_t(hj_handle* h, const u64* R, const SegMap& sm, u64 key_mask) {
  auto k = gt_nocc_build_kernel<IPT>;
  int grid = grid_for(h, (const void*)k, kGtThreads, 0);
  u64 chunks = ((sm.seg_len + kGtThreads * IPT - 1) / (kGtThreads * IPT)) * sm.nslices;
  if (chunks == 0) return HJ_OK;
  if ((u64)grid > chunks) grid = (int)chunks;
  k<<<grid, kGtThreads, 0, h->stream>>>(R, sm, h->table.p, h->mask, key_mask, h->cfg.probe_length, h->ovf_list.p,
                                        h->ctr);
  h->launches++;
  CK(cudaGetLastError());
  return HJ_OK;
}

// (8-byte loads: no alignment requirement beyond the tuple's own)
int launch_build(hj_handle* h, const u64* R, const SegMap& sm, bool nocc, bool stats, u32 ipt) {
  const u64 km = h->cfg.key_mask;
  if (stats) return launch_build_t<2, true>(h, R, sm, km);
  if (nocc) {
    switch (ipt) {
      case 1: return launch_nocc_t<1>(h, R, sm, km);
      case 2: return launch_nocc_t<2>(h, R, sm, km);
      case 4: return launch_nocc_t<4>(h, R, sm, km);
      default: return launch_nocc_t<8>(h, R, sm, km);
    }
  }
  switch (ipt) {
    case 1: return launch_build_t<1, false>(h, R, sm, km);
    case 2: return launch_build_t<2, false>(h, R, sm, km);
    case 4: return launch_build_t<4, false>(h, R, sm, km);
    default: return launch_build_t<8, false>(h, R, sm, km);
  }
}

template <int IPT>
int launch_probe_t(hj_handle* h, const u64* S, u64 m) {
  if (m == 0) return HJ_OK;
  {  // every tuple sits in its home slot (decided on the device from the build's counters): one read per S tuple
    constexpr int HIPT = IPT >= 4 ? 8 : IPT;
    auto k = gt_probe_home_kernel<HIPT>;
    int grid = grid_for(h, (const void*)k, kGtThreads, 0);
    const u64 chunks = (m + kGtThreads * HIPT - 1) / (kGtThreads * HIPT);
    if ((u64)grid > chunks) grid = (int)chunks;
    k<<<grid, kGtThreads, 0, h->stream>>>(S, m, h->table.p, h->mask, h->cfg.key_mask, h->ctr);
  }
  auto k = gt_probe_kernel<IPT>;
  int grid = grid_for(h, (const void*)k, kGtThreads, 0);
  u64 chunks = (m + kGtThreads * IPT - 1) / (kGtThreads * IPT);
  if ((u64)grid > chunks) grid = (int)chunks;
  k<<<grid, kGtThreads, 0, h->stream>>>(S, m, h->table.p, h->mask, h->cfg.key_mask, h->cfg.probe_length,
                                        h->ovf_table.p, h->ovf_table.cap, reinterpret_cast<const u32*>(h->ovfmap.p),
                                        h->ctr);
  h->launches += 2;
  CK(cudaGetLastError());
  return HJ_OK;
}
int launch_probe(hj_handle* h, const u64* S, u64 m, u32 ipt) {
  switch (ipt) {
    case 1: return launch_probe_t<1>(h, S, m);
    case 2: return launch_probe_t<2>(h, S, m);
    default: return launch_probe_t<4>(h, S, m);
  }
}

int zero_async(hj_handle* h, u64* p, u64 elems) {
  // 16 B stores from a full-chip grid; cudaMemsetAsync would do as well but a
  // kernel keeps the launch accounting honest
  if (elems == 0) return HJ_OK;
  u64 n2 = elems / 2;
  int grid = h->sm_count * 8;
  fill_zero_kernel<<<grid, 256, 0, h->stream>>>(reinterpret_cast<ulonglong2*>(p), n2);
  h->launches++;
  if (elems & 1) CK(cudaMemsetAsync(p + elems - 1, 0, 8, h->stream));
  CK(cudaGetLastError());
  return HJ_OK;
}

int prepare_table(hj_handle* h, u64 n) {
  u32 scale = h->cfg.scale_output ? h->cfg.scale_output : 2;
  u64 slots = pow2_ceil_u64(n * scale < 128 ? 128 : n * scale);  // >= 4 words of the slot bitmaps
  if (slots > (1ull << 32)) {  // 32-bit slot arithmetic (the reference's sizes are uint32_t, AtomicHashBuild.hpp:15,21)
    snprintf(h->err, sizeof(h->err), "global table: %llu slots exceed 2^32; use the radix path for |R| > 2^31 / scaleOutput", slots);
    return HJ_EINVAL;
  }
  int rc;
  if ((rc = ensure(h, h->table, slots))) return rc;
  if ((rc = ensure(h, h->ovf_table, slots))) return rc;  // worst case: everything overflows, load 0.5
  if ((rc = ensure(h, h->ovf_list, n ? n : 1))) return rc;
  if ((rc = ensure(h, h->claim, slots / 64))) return rc;   // one bit per slot (slots >= 64)
  if ((rc = ensure(h, h->ovfmap, slots / 64))) return rc;
  if (slots != h->slots) h->built = false;  // a table of another geometry: what was built is gone
  h->slots = slots;
  h->mask = slots - 1;
  return HJ_OK;
}

// table + claim bitmap to zero (the reference's `new ...{}`, AtomicHashBuild.hpp:25: untimed there)
int clear_table(hj_handle* h) {
  int rc;
  if ((rc = zero_async(h, h->table.p, h->slots))) return rc;
  CK(cudaMemsetAsync(h->claim.p, 0, h->slots / 8, h->stream));
  h->built = false;
  return HJ_OK;
}

// Stage a host relation on the copy stream (or pass a device pointer through).
// *pending is set when the compute stream still has to wait for ev_done.
int stage(hj_handle* h, const u64* p, u64 n, DevBuf& buf, const u64** dev, int ev_done, bool* pending,
          bool need16 = true) {
  *pending = false;
  if (n == 0 || p == nullptr) {
    *dev = nullptr;
    return HJ_OK;
  }
  int rc;
  if (is_device_ptr(p)) {
    if (!need16 || ((uintptr_t)p & 15) == 0) {  // (the global-table kernels read 8 bytes at a time)
      *dev = p;
      return HJ_OK;
    }
    // a device relation that is only 8-byte aligned (a slice t[1:], an odd tuple offset): the radix
    // kernels read tuples in 16-byte pairs, so it is moved to an aligned staging buffer first
    if ((rc = ensure(h, buf, n + (n & 1)))) return rc;
    CK(cudaMemcpyAsync(buf.p, p, n * sizeof(u64), cudaMemcpyDeviceToDevice, h->stream));
    *dev = buf.p;
    return HJ_OK;
  }
  if ((rc = ensure(h, buf, n + (n & 1)))) return rc;
  CK(cudaMemcpyAsync(buf.p, p, n * sizeof(u64), cudaMemcpyHostToDevice, h->copy_stream));
  mark(h, ev_done, h->copy_stream);
  *pending = true;
  *dev = buf.p;
  return HJ_OK;
}
int wait_staged(hj_handle* h, int ev_done, bool pending) {
  if (pending) CK(cudaStreamWaitEvent(h->stream, h->ev[ev_done], 0));
  return HJ_OK;
}

void reset_events(hj_handle* h) {
  for (int i = 0; i < E_COUNT; i++) h->ev_set[i] = false;
  h->launches = 0;
}

int fetch_counters(hj_handle* h) {
  CK(cudaMemcpyAsync(h->h_ctr, h->ctr, C_COUNT * sizeof(u64), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return HJ_OK;
}

void fill_common(hj_handle* h, hj_result* r) {
  const u64* c = h->h_ctr;
  r->matches = c[C_MATCHES];
  r->checksum = c[C_CHECKSUM];
  r->window_matches = c[C_WINDOW_MATCHES];
  r->conflicts = c[C_CONFLICTS];
  r->cas_retries = c[C_CAS_RETRIES];
  r->probe_steps = c[C_PROBE_STEPS];
  r->sampled = c[C_SAMPLED];
  r->sectors = c[C_SECTORS];
  r->max_partition = c[C_MAX_PART];
  r->zero_keys = c[C_ZERO_KEYS];
  r->kernel_launches = h->launches;
  r->us_total = span_us(h, E_T0, E_END);
  r->us_h2d = span_us(h, E_CP0, h->ev_set[E_CPS] ? E_CPS : E_CPR);
  r->us_clear = span_us(h, E_T0, E_CLEAR);
  r->us_sample = span_us(h, E_CLEAR, E_SAMPLE);
  r->us_build = span_us(h, h->ev_set[E_SAMPLE] ? E_SAMPLE : E_CLEAR, E_BUILD);
  r->us_probe = span_us(h, E_PROBE0, E_PROBE);
  for (int rel = 0; rel < 2; rel++) {
    const int b = E_RX + 5 * rel;
    r->us_hist += span_us(h, b, b + 1);
    r->us_scan += span_us(h, b + 1, b + 2);
    r->us_scatter1 += span_us(h, b + 2, b + 3);
    r->us_scatter2 += span_us(h, b + 3, b + 4);
  }
  r->us_join = span_us(h, E_JOIN0, E_JOIN);
}

// contention signal of the sampling round (see DESIGN.md "Adaptive switch")
double contention_of(hj_handle* h) {
  const u64* c = h->h_ctr;
  if (c[C_SAMPLED] == 0) return 0.0;
  double retry = (double)c[C_CAS_RETRIES] / (double)c[C_SAMPLED];
  double span_bytes = c[C_CHUNKS] ? 8.0 * (double)c[C_SPAN] / (double)c[C_CHUNKS] : 0.0;
  double locality = span_bytes / (double)h->l2_bytes;
  if (locality > 1.0) locality = 1.0;
  return locality + retry;
}

// Default switch point.  The span signal says how far apart the slots of one CTA chunk lie; what that costs depends on
// how much of the table the 126 MB L2 holds anyway.  Fitted to the crossovers of the round-2 build (R = local_shuffle W,
// experiments/results/w_sweep_n24.jsonl, w_sweep_n26.jsonl, w_sweep_n27.jsonl), with r = 2 L2 / table bytes:
//   |R| = 2^24 (r = 0.94): the bitmap-claimed table wins up to W = 2^15 (signal 0.004), ties at 2^16 (0.008);
//   |R| = 2^26 (r = 0.23): wins up to 2^14 (0.002), loses from 2^16 (0.008);
//   |R| = 2^27 (r = 0.12): ties from 2^12 to 2^14 (0.0005 .. 0.002), loses from 2^15 (0.004).
// Threshold 0.0068 r^0.55 (0.0066 / 0.0031 / 0.0021 at those sizes).  A table that fits L2 whole (r >= 2) keeps the
// old 0.25 -- only a span of a quarter of L2 sends it to the radix join -- with a log-linear blend in between.
double default_threshold(const hj_handle* h) {
  const double table_bytes = 8.0 * (double)h->slots;
  const double r = 2.0 * (double)h->l2_bytes / (table_bytes > 0 ? table_bytes : 1.0);
  constexpr double kAtOne = 0.0068, kInL2 = 0.25;
  if (r >= 2.0) return kInL2;
  if (r >= 1.0) return exp(log(kAtOne) + (log(kInL2) - log(kAtOne)) * (r - 1.0));
  return kAtOne * pow(r, 0.55);
}

SegMap whole(u64 n) {
  SegMap sm;
  sm.n = n;
  sm.seg_len = n;
  sm.slice_stride = 0;
  sm.slice_off = 0;
  sm.nslices = 1;
  return sm;
}

// Sampling geometry of HTMHashBuild.hpp:100-113: 64 slices, K*16384 tuples each.
void sample_geometry(const hj_handle* h, u64 n, u64* stride, u64* sample_len) {
  const u64 nsl = 64;
  u32 K = h->cfg.sample_rounds ? h->cfg.sample_rounds : 5;
  u64 st = n / nsl;
  u64 sl = (u64)K * 16384ull;
  if (st < 2) {  // tiny relation: one slice, sample an eighth
    *stride = 0;
    *sample_len = n / 8 ? n / 8 : n;
    return;
  }
  if (sl > st / 4) sl = st / 4 ? st / 4 : 1;  // never sample more than a quarter
  *stride = st;
  *sample_len = sl;
}

int run_sample(hj_handle* h, const u64* dR, u64 n, u64* stride_out, u64* len_out) {
  u64 stride, slen;
  sample_geometry(h, n, &stride, &slen);
  SegMap sm;
  sm.n = n;
  sm.seg_len = slen;
  sm.slice_stride = stride;
  sm.slice_off = 0;
  sm.nslices = stride ? 64 : 1;
  *stride_out = stride;
  *len_out = slen;
  return launch_build(h, dR, sm, false, true, 2);
}

// locality pre-check (gt_span_kernel): same sample geometry, no table
int run_span(hj_handle* h, const u64* dR, u64 n) {
  u64 stride, slen;
  sample_geometry(h, n, &stride, &slen);
  SegMap sm;
  sm.n = n; sm.seg_len = slen; sm.slice_stride = stride; sm.slice_off = 0; sm.nslices = stride ? 64 : 1;
  const u64 chunks = ((slen + kGtThreads * 2 - 1) / (kGtThreads * 2)) * sm.nslices;
  if (chunks == 0) return HJ_OK;
  int grid = grid_for(h, (const void*)gt_span_kernel, kGtThreads, 0);
  if ((u64)grid > chunks) grid = (int)chunks;
  gt_span_kernel<<<grid, kGtThreads, 0, h->stream>>>(dR, sm, h->mask, h->cfg.key_mask, h->ctr);
  h->launches++;
  CK(cudaGetLastError());
  return HJ_OK;
}

int run_build_rest(hj_handle* h, const u64* dR, u64 n, u64 stride, u64 slen, bool nocc, u32 ipt) {
  int rc;
  if (stride == 0) {  // single slice: the rest is [slen, n)
    SegMap sm;
    sm.n = n; sm.seg_len = n - slen; sm.slice_stride = 0; sm.slice_off = slen; sm.nslices = 1;
    return sm.seg_len ? launch_build(h, dR, sm, nocc, false, ipt) : HJ_OK;
  }
  SegMap sm;
  sm.n = n; sm.seg_len = stride - slen; sm.slice_stride = stride; sm.slice_off = slen; sm.nslices = 64;
  if ((rc = launch_build(h, dR, sm, nocc, false, ipt))) return rc;
  if (64 * stride < n) {  // tail not covered by the 64 slices
    SegMap t;
    t.n = n; t.seg_len = n - 64 * stride; t.slice_stride = 0; t.slice_off = 64 * stride; t.nslices = 1;
    if ((rc = launch_build(h, dR, t, nocc, false, ipt))) return rc;
  }
  return HJ_OK;
}

int run_overflow(hj_handle* h) {
  int grid = h->sm_count * 8;
  u32* ovfmap = reinterpret_cast<u32*>(h->ovfmap.p);
  gt_ovf_clear_kernel<<<grid, 256, 0, h->stream>>>(h->ovf_table.p, ovfmap, h->slots / 32, h->ctr, h->ovf_table.cap);
  gt_ovf_insert_kernel<<<grid, 256, 0, h->stream>>>(h->ovf_list.p, h->ovf_table.p, ovfmap, h->ctr, h->mask,
                                                    h->cfg.key_mask, h->ovf_table.cap);
  h->launches += 2;
  CK(cudaGetLastError());
  return HJ_OK;
}

int clear_probe_counters(hj_handle* h) {
  CK(cudaMemsetAsync(h->ctr + C_MATCHES, 0, 3 * sizeof(u64), h->stream));
  return HJ_OK;
}

// ---- radix --------------------------------------------------------------------

// The building-block entry points take device pointers in place: they must be 16-byte aligned
// (the kernels read tuples in pairs).  hj_join & co. re-stage a misaligned device input instead.
int need_aligned16(hj_handle* h, const void* p, const char* what) {
  if (((uintptr_t)p & 15) == 0) return HJ_OK;
  snprintf(h->err, sizeof(h->err), "%s: device pointer %p is not 16-byte aligned", what, p);
  return HJ_EINVAL;
}

// tile of the exact-layout passes (rx_hist + rx_scan offsets): 512 x 16 = 8192 tuples, as the histogram-free ones
constexpr int kExactIPT = 16;

struct RadixPlan {
  u32 bits, passes, b1, b2;
  u32 slots;        // join table slots
  u32 tile_ipt;     // scatter items per thread
};

constexpr u32 kMaxBitsPerPass = 12;
constexpr u32 kMaxTotalBits = 15;  // 2^15 bins * 4 B = 128 KiB privatised histogram

int plan_radix(hj_handle* h, u64 n, RadixPlan* pl, u64 target_per = 16384) {
  u32 bits = 0;
  if (h->cfg.num_partitions) {
    u32 f = h->cfg.num_partitions;
    if (f & (f - 1)) return HJ_EINVAL;
    while ((1u << bits) < f) bits++;
  } else {
    // default: R partitions of <= 16384 tuples -- the largest rx_join32 joins in one piece (one
    // 1024-thread CTA per SM); measured at C3: 13 bits (6+7) 11.9 ms, 14 bits 12.2, 15 bits 12.7
    // (cheaper scatter passes outweigh the slower 1-CTA join).  Mid-size inputs get at least 2048
    // partitions so every SM has several.
    while (bits < kMaxTotalBits && (n >> bits) > target_per) bits++;
    while (bits < 11 && (n >> bits) > 2048) bits++;
    if (bits < 4) bits = 4;
  }
  if (bits < 1 || bits > kMaxTotalBits) return HJ_EINVAL;
  u32 passes = h->cfg.passes;
  if (passes == 0) passes = bits > 8 ? 2 : 1;
  if (passes > 2) return HJ_EINVAL;
  // 1 pass beyond 2^12 (BASELINE configs[2] sweeps 1-pass up to 2^14, prj_params.h:15-22 builds it): no
  // shared-memory tile holds a run per digit any more -- the direct scatter takes those (rx_scatter_direct_kernel)
  if (passes == 2 && bits < 2) passes = 1;
  pl->bits = bits;
  pl->passes = passes;
  pl->b1 = passes == 2 ? bits / 2 : bits;  // prj_params.h:76-78: pass 1 gets floor(b/2)
  if (passes == 2 && h->cfg.pass1_bits) {
    const u32 want = (u32)h->cfg.pass1_bits;
    if (want < 1 || want >= bits || want > kMaxBitsPerPass || bits - want > kMaxBitsPerPass) return HJ_EINVAL;
    pl->b1 = want;
  }
  pl->b2 = bits - pl->b1;
  // join table: expected partition * 2, power of two, capped by shared memory
  u64 per = (n >> bits) + 1;
  u64 slots = pow2_ceil_u64(per * 2 < 1024 ? 1024 : per * 2);
  u64 cap = 1;
  while (cap * 2 * sizeof(u64) <= h->smem_optin - 1024) cap *= 2;  // 16384 slots at 227 KB
  if (slots > cap) slots = cap;
  pl->slots = (u32)slots;
  pl->tile_ipt = kExactIPT;
  return HJ_OK;
}

// region capacity of the histogram-free layout for a relation of n tuples in 2^bits partitions:
// mean + 8 sigma of a random split + 64, even
u64 opt_region_cap(u64 n, u32 bits) {
  const u64 avg = (n >> bits) + 1;
  u64 r = 1;
  while (r * r < avg * 64) r++;  // 8 * sqrt(avg)
  return (avg + r + 64 + 1) & ~1ull;
}

int ensure_rel(hj_handle* h, RelBufs& rb, u64 n, const RadixPlan& pl, bool pre1 = false, bool opt = false,
               u32 n_src = 1) {
  int rc;
  u64 nb = 1ull << pl.bits;
  u64 np = (1ull << pl.b1) * n_src;  // parents of pass 2 (one per source rank after a region exchange)
  // histogram-free layout: 2^bits regions + one scatter tile of scratch for diverted claims
  const u64 elems = opt ? nb * opt_region_cap(n, pl.bits) + kScThreads * 32 + 2 : n + 2;
  if ((rc = ensure(h, rb.out, elems))) return rc;
  if (pl.passes == 2 && !pre1 && (rc = ensure(h, rb.tmp, elems))) return rc;
  if (opt && (rc = ensure(h, rb.pend, np + 2))) return rc;
  if ((rc = ensure(h, rb.ghist, nb))) return rc;
  if ((rc = ensure(h, rb.offsets, nb + 1))) return rc;
  if ((rc = ensure(h, rb.cursors, nb))) return rc;
  if ((rc = ensure(h, rb.poff, np + 2))) return rc;
  if ((rc = ensure(h, rb.pcur, np + 2))) return rc;
  if ((rc = ensure(h, rb.ptiles, np + 2))) return rc;
  if ((rc = ensure(h, rb.one, 4))) return rc;
  return HJ_OK;
}

// Histogram launch: 512-thread CTAs, or 1024 when the privatised histogram (> 64 KiB)
// leaves room for one CTA per SM only.
int launch_hist(hj_handle* h, const u64* in, u64 n, const Digit& dg, u32 nbins, u64* ghist) {
  u64* ctr = h->ctr;
  const size_t smem = (size_t)nbins * 4;
  const bool big = smem > 64 * 1024;
  const int threads = big ? 1024 : 512;
  const void* k = big ? (const void*)rx_hist_kernel<true, 1024> : (const void*)rx_hist_kernel<true, 512>;
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int grid = grid_for(h, k, threads, smem);
  u64 tiles = (n + (u64)threads * kHistIPT - 1) / ((u64)threads * kHistIPT);
  if ((u64)grid > tiles) grid = (int)(tiles ? tiles : 1);
  if (big) rx_hist_kernel<true, 1024><<<grid, 1024, smem, h->stream>>>(in, n, dg, nbins, ghist, ctr);
  else rx_hist_kernel<true, 512><<<grid, 512, smem, h->stream>>>(in, n, dg, nbins, ghist, ctr);
  h->launches++;
  CK(cudaGetLastError());
  return HJ_OK;
}

bool sc_bulk() { return !env().sc_no_bulk; }
// Input tiles staged by TMA: measured SLOWER than register loads at C3 (scatter 1: 4.49 vs 3.80 ms,
// scatter 2: 4.86 vs 4.16 ms, 3 CTAs/SM in both): the staging pass adds a shared-memory write and a
// read per tuple to a kernel whose shared-memory pipe is already the busiest unit.  Opt-in only.
bool sc_tmain() { return env().sc_tmain; }
// Bulk (TMA) flush per pass: one cp.async.bulk per run pays off while a run is long -- 4096-tuple tiles
// over F <= 128 digits = 32+ tuples (scatter 2 at C3: 4.68 -> 4.09 ms); at 256 / 512 digits (16 / 8 tuples
// per run, the pass 2 after a 4- / 8-GPU exchange) it LOSES to 8-byte stores from consecutive lanes:
// 2.48 vs 2.27 ms and 3.33 vs 2.60 ms per 2^30-tuple launch.
bool sc_bulk_for(u32 F, u32 tile) { return sc_bulk() && F % 2 == 0 && F <= tile / 32; }

// Tile of the histogram-free passes: 512 threads x IPT tuples.  8192-tuple tiles (IPT 16, two CTAs per SM,
// 64 registers) beat 4096-tuple ones (IPT 8, three CTAs, 40 registers, a few spills): runs are twice as long
// and every tuple pays for half as many barriers -- C3: scatter 1 3.77 -> 3.37 ms (87 % of the HBM roofline),
// scatter 2 4.14 -> 3.57 ms, 116 k -> 128 k Mtuples/s.  HJ_SC_IPT = 8 | 16 | 32 overrides.
u32 sc_opt_ipt() {
  const int x = env().sc_ipt;
  return (u32)(x == 8 || x == 32 ? x : 16);
}
using ScKernel = void (*)(const u64*, u64*, Digit, u32, u32, const u64*, const u64*, u64*, u64* const*, ScOpt);
template <int IPT>
ScKernel sc_opt_kernel(int opt, bool bulk) {
  if (opt == 1) return bulk ? rx_scatter_kernel<IPT, 0, false, 1, true> : rx_scatter_kernel<IPT, 0, false, 1, false>;
  return bulk ? rx_scatter_kernel<IPT, 0, false, 2, true> : rx_scatter_kernel<IPT, 0, false, 2, false>;
}
// BULK shared memory on top of tile + F*16: one spare tile slot per digit (+2) and the run lengths
size_t sc_bulk_extra(u32 F) { return (size_t)(F + 2) * 8 + (size_t)F * 4; }

template <int IPT, int MODE>
int launch_scatter_tm(hj_handle* h, const u64* in, u64* out, const Digit& dg, u32 F, u32 nparent, const u64* poff,
                      const u64* ptiles, u64* cursors, u64 max_tiles) {
  // runs leave through bulk async copies (TMA) when the fan-out is even; HJ_SC_NO_BULK: 8-byte stores
  const bool bulk = sc_bulk_for(F, kScThreads * IPT);
  auto k = bulk ? rx_scatter_kernel<IPT, MODE, false, 0, true> : rx_scatter_kernel<IPT, MODE, false, 0, false>;
  size_t smem = (size_t)kScThreads * IPT * 8 + (size_t)F * 16 + (bulk ? sc_bulk_extra(F) : 0);
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int grid = grid_for(h, (const void*)k, kScThreads, smem);
  if ((u64)grid > max_tiles) grid = (int)(max_tiles ? max_tiles : 1);
  k<<<grid, kScThreads, smem, h->stream>>>(in, out, dg, F, nparent, poff, ptiles, cursors, nullptr, ScOpt{});
  h->launches++;
  CK(cudaGetLastError());
  return HJ_OK;
}
// mode-0 digits on bits below 32 take the specialised kernel; everything else the generic one
template <int IPT>
int launch_scatter_t(hj_handle* h, const u64* in, u64* out, const Digit& dg, u32 F, u32 nparent, const u64* poff,
                     const u64* ptiles, u64* cursors, u64 max_tiles) {
  const bool low32 = (h->cfg.key_mask & 0xffffffffull) == 0xffffffffull;
  if (dg.mode == 0 && dg.shift + dg.bits <= 32 && low32 && (F & (F - 1)) == 0)
    return launch_scatter_tm<IPT, 0>(h, in, out, dg, F, nparent, poff, ptiles, cursors, max_tiles);
  return launch_scatter_tm<IPT, 2>(h, in, out, dg, F, nparent, poff, ptiles, cursors, max_tiles);
}

// Partition `in` (device, n tuples) by pl into rb.out; rb.offsets gets 2^bits+1 offsets.
// evb = first of this relation's five phase events (or -1: untimed).
// pre1: `in` already IS the pass-1 output (partitioned by the low b1 bits, partition-major, no
// gaps -- written by the peers' fused exchange scatter): histogram it, skip pass 1, run pass 2 on it.
int radix_partition_rel(hj_handle* h, const u64* in, u64 n, RelBufs& rb, const RadixPlan& pl, int evb,
                        u64* max_bin = nullptr, bool pre1 = false) {
  const u32 nb = 1u << pl.bits;
  const u32 F1 = 1u << pl.b1, F2 = 1u << pl.b2;
  const u32 tile = kScThreads * pl.tile_ipt;
  Digit full;
  memset(&full, 0, sizeof(full));
  full.key_mask = h->cfg.key_mask;
  if (pl.passes == 2) { full.mode = 1; full.b1 = pl.b1; full.b2 = pl.b2; }
  else { full.mode = 0; full.shift = 0; full.bits = pl.bits; }

  if (evb >= 0) mark(h, evb);
  CK(cudaMemsetAsync(rb.ghist.p, 0, nb * sizeof(u64), h->stream));
  {
    int rc0 = launch_hist(h, in, n, full, nb, rb.ghist.p);
    if (rc0) return rc0;
  }
  if (evb >= 0) mark(h, evb + 1);
  // scan: final offsets, the pass-1 ("parent") view and the one-parent view {0,n}.
  // Single pass: f2 = nb makes parent 0 span every bin.
  rx_scan_kernel<<<1, kScanThreads, 0, h->stream>>>(rb.ghist.p, nb, rb.offsets.p, rb.cursors.p,
                                                    pl.passes == 2 ? F2 : nb, rb.poff.p, rb.pcur.p, rb.ptiles.p,
                                                    rb.one.p, tile, max_bin, n);
  h->launches++;
  CK(cudaGetLastError());
  if (evb >= 0) mark(h, evb + 2);

  const u64 max_tiles1 = (n + tile - 1) / tile;
  int rc;
  if (pl.passes == 1 && pl.bits > kMaxBitsPerPass) {
    // 2^13 .. 2^15 partitions in ONE pass: a tile of 8192 tuples holds less than one tuple per digit, so
    // there is nothing to write-combine; every tuple claims its slot from the partition's cursor
    rx_scatter_direct_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>(in, n, rb.out.p, full, rb.cursors.p);
    h->launches++;
    CK(cudaGetLastError());
    if (evb >= 0) { mark(h, evb + 3); mark(h, evb + 4); }
  } else if (pl.passes == 1) {
    if ((rc = launch_scatter_t<kExactIPT>(h, in, rb.out.p, full, nb, 1, rb.poff.p, rb.ptiles.p, rb.cursors.p, max_tiles1)))
      return rc;
    if (evb >= 0) { mark(h, evb + 3); mark(h, evb + 4); }
  } else {
    // pass 1: the whole relation by the low b1 bits into tmp, claiming from the parent cursors
    Digit d1;
    memset(&d1, 0, sizeof(d1));
    d1.key_mask = h->cfg.key_mask; d1.mode = 0; d1.shift = 0; d1.bits = pl.b1;
    if (!pre1 &&
        (rc = launch_scatter_t<kExactIPT>(h, in, rb.tmp.p, d1, F1, 1, rb.one.p, rb.one.p + 2, rb.pcur.p, max_tiles1)))
      return rc;
    if (evb >= 0) mark(h, evb + 3);
    // pass 2: every pass-1 partition by the next b2 bits into out (radix_cluster, :402-440)
    Digit d2;
    memset(&d2, 0, sizeof(d2));
    d2.key_mask = h->cfg.key_mask; d2.mode = 0; d2.shift = pl.b1; d2.bits = pl.b2;
    if ((rc = launch_scatter_t<kExactIPT>(h, pre1 ? in : rb.tmp.p, rb.out.p, d2, F2, F1, rb.poff.p, rb.ptiles.p,
                                  rb.cursors.p, max_tiles1 + F1)))
      return rc;
    if (evb >= 0) mark(h, evb + 4);
  }
  return HJ_OK;
}

// Histogram-free 2-pass partitioning (ScOpt in rx_kernels.cuh): both passes write into
// fixed-capacity regions, rb.cursors ends up holding the fill level of every final region.
// pre1_off != nullptr: `in` is already the pass-1 output (fused exchange) with exact parent offsets
// pre1_off[0..F1]: only pass 2 runs, and it gathers the key-width flag.
// What a histogram-free exchange leaves in a receive buffer: pass-1 partition `low` as n_src regions of
// `cap` tuples, region low * n_src + src = [start, cur[low * n_src + src]) (hj_join_regions).
struct Pre1Regions {
  const u64* cur;
  u64 cap;
  u32 n_src;
};

int radix_partition_rel_opt(hj_handle* h, const u64* in, u64 n, RelBufs& rb, const RadixPlan& pl, int evb,
                            const u64* pre1_off = nullptr, const Pre1Regions* reg = nullptr) {
  const u32 nb = 1u << pl.bits, F1 = 1u << pl.b1, F2 = 1u << pl.b2;
  const u32 ipt = sc_opt_ipt();
  const u32 tile = kScThreads * ipt;
  const u64 C2 = opt_region_cap(n, pl.bits), C1 = C2 * F2;
  const u64 ntiles = (n + tile - 1) / tile;
  if (evb >= 0) { mark(h, evb); mark(h, evb + 1); }
  rx_opt_init_kernel<<<(nb + 255) / 256, 256, 0, h->stream>>>(rb.pcur.p, F1, C1, rb.cursors.p, nb, C2);
  const u64 one[4] = {0, n, 0, ntiles};  // pass 1: one parent [0, n) of ntiles tiles
  CK(cudaMemcpyAsync(rb.one.p, one, sizeof(one), cudaMemcpyHostToDevice, h->stream));
  h->launches++;
  if (evb >= 0) mark(h, evb + 2);
  // per pass: runs leave through bulk copies (TMA) while they are long enough (sc_bulk_for)
  const bool bulk1 = sc_bulk_for(F1, tile), bulk2 = sc_bulk_for(F2, tile);
  // input tiles staged through shared memory by one bulk copy each (TMAIN): 4096-tuple tiles + bulk flush only
  const bool tmain = ipt == 8 && bulk1 && bulk2 && sc_tmain();
  ScKernel k, k1b, k2;  // pass 1; pass 2 after an exchange (key width still unknown); pass 2
  if (tmain) {
    k = k1b = rx_scatter_kernel<8, 0, false, 1, true, true>;
    k2 = rx_scatter_kernel<8, 0, false, 2, true, true>;
  } else if (ipt == 32) {
    k = sc_opt_kernel<32>(1, bulk1); k1b = sc_opt_kernel<32>(1, bulk2); k2 = sc_opt_kernel<32>(2, bulk2);
  } else if (ipt == 16) {
    k = sc_opt_kernel<16>(1, bulk1); k1b = sc_opt_kernel<16>(1, bulk2); k2 = sc_opt_kernel<16>(2, bulk2);
  } else {
    k = sc_opt_kernel<8>(1, bulk1); k1b = sc_opt_kernel<8>(1, bulk2); k2 = sc_opt_kernel<8>(2, bulk2);
  }
  const size_t stage = tmain ? (size_t)tile * 8 : 0;
  Digit d1, d2;
  memset(&d1, 0, sizeof(d1));
  d1.key_mask = h->cfg.key_mask; d1.mode = 0; d1.shift = 0; d1.bits = pl.b1;
  d2 = d1; d2.shift = pl.b1; d2.bits = pl.b2;
  u32 nparent2 = F1, pshift = 0;
  if (reg) {
    nparent2 = F1 * reg->n_src;
    while ((1u << pshift) < reg->n_src) pshift++;
    if ((1u << pshift) != reg->n_src || nparent2 > (u32)kScParentCache) return HJ_EINVAL;
  }
  if (!pre1_off && !reg) {
    size_t smem = stage + (size_t)tile * 8 + (size_t)F1 * 16 + (bulk1 ? sc_bulk_extra(F1) : 0);
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = grid_for(h, (const void*)k, kScThreads, smem);
    if ((u64)grid > ntiles) grid = (int)(ntiles ? ntiles : 1);
    ScOpt o{rb.one.p + 1, C1, (u64)F1 * C1, h->ctr};
    k<<<grid, kScThreads, smem, h->stream>>>(in, rb.tmp.p, d1, F1, 1, rb.one.p, rb.one.p + 2, rb.pcur.p, nullptr, o);
    h->launches++;
    CK(cudaGetLastError());
  }
  if (evb >= 0) mark(h, evb + 3);
  if (reg) rx_opt_parents_kernel<<<1, 1024, 0, h->stream>>>(reg->cur, nparent2, reg->cap, tile, rb.poff.p, rb.pend.p, rb.ptiles.p);
  else if (pre1_off) rx_pre1_parents_kernel<<<1, 1024, 0, h->stream>>>(pre1_off, F1, tile, rb.poff.p, rb.pend.p, rb.ptiles.p);
  else rx_opt_parents_kernel<<<1, 1024, 0, h->stream>>>(rb.pcur.p, F1, C1, tile, rb.poff.p, rb.pend.p, rb.ptiles.p);
  h->launches++;
  {
    auto kp2 = (pre1_off || reg) ? k1b : k2;  // after the fused exchange nobody has looked at the key width yet
    size_t smem = stage + (size_t)tile * 8 + (size_t)F2 * 16 + (bulk2 ? sc_bulk_extra(F2) : 0);
    CK(cudaFuncSetAttribute(kp2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = grid_for(h, (const void*)kp2, kScThreads, smem);
    if ((u64)grid > ntiles + nparent2) grid = (int)(ntiles + nparent2);
    ScOpt o{rb.pend.p, C2, (u64)nb * C2, h->ctr, pshift, 0, 0, 0};
    kp2<<<grid, kScThreads, smem, h->stream>>>((pre1_off || reg) ? in : rb.tmp.p, rb.out.p, d2, F2, nparent2, rb.poff.p,
                                               rb.ptiles.p, rb.cursors.p, nullptr, o);
    h->launches++;
    CK(cudaGetLastError());
  }
  if (evb >= 0) mark(h, evb + 4);
  return HJ_OK;
}

// Which layout and which join a radix call takes -- one function, so that hj_reserve sizes exactly
// what hj_join will use.
struct RadixChoice {
  RadixPlan pl;
  u64 need;     // R tuples per partition incl. 8 sigma of a random split
  bool use_l2;  // partitions too large for shared memory: the L2-resident partition-local table
  bool opt;     // histogram-free layout (fixed-capacity regions)
};
int choose_radix(hj_handle* h, u64 n, bool has_s, bool pre1, bool pre1_known, bool force_exact, bool sortmerge,
                 RadixChoice* ch) {
  int rc;
  // sort-merge: R partitions of <= 4096 tuples sort in one 8192-element shared-memory network with slack
  if ((rc = plan_radix(h, n, &ch->pl, sortmerge ? 4096 : 16384))) return rc;
  const RadixPlan& pl = ch->pl;
  // Which join: shared memory while an R partition (plus 8 sigma of a random split) fits the
  // largest table one CTA can hold, else the L2-resident partition-local table (rx_joinl2.cuh).
  const u64 per = (n >> pl.bits) + 1;
  u64 need = per + 64;
  for (u64 r = 1; r * r < per * 64; r++) need = per + 64 + r;  // + 8*sqrt(per)
  ch->need = need;
  // up to three shared-memory chunks per R partition (S_p is probed once per chunk) still beat the L2 table:
  // 2^12 partitions at C3 (2 chunks) 17.6 ms through L2; HJ_L2_MIN_NEED overrides the switch point
  ch->use_l2 = !sortmerge && has_s && (env().force_l2 || need > env().l2_min_need) && !env().no_l2;
  // Histogram-free partitioning (ScOpt, rx_kernels.cuh) for the 2-pass shared-memory join; the
  // exact, histogram-based layout everywhere else and after an overflow on this handle.
  ch->opt = has_s && (!pre1 || pre1_known) && !ch->use_l2 && pl.passes == 2 && (1u << pl.b1) <= (u32)kScParentCache &&
            !h->no_optimistic && !force_exact && !env().no_optimistic;
  return HJ_OK;
}

// The per-partition shared-memory join over h->rb[0].out / h->rb[1].out (exact layout: offsets; histogram-free
// layout: regions of capR / capS tuples whose fill levels are the scatter's cursors).
// nsrc > 1: S was partitioned in nsrc separate chunks (hj_xchg): chunk c's partitions at rb[1].out + c * src_stride,
// its fill levels at rb[1].cursors + c * off_stride; R_p's table is built once and probed with every chunk's S_p.
int launch_partition_join(hj_handle* h, const RadixPlan& pl, u64 need, bool opt, u64 n, bool has_s, u64 m, u64 capR,
                          u64 capS, u32 nsrc = 1, u64 src_stride = 0, u64 off_stride = 0, u32 src_first = 0) {
  int rc;
  const u32 nparts = 1u << pl.bits;
  // 32-bit-key fast path first (returns at once if rx_hist saw wider keys), then the generic
  // kernel (returns at once if the fast path took the join).  Build-only PRO is generic.
  const bool try32 = has_s && pl.bits >= 1 && (h->cfg.key_mask & 0xffffffffull) == 0xffffffffull &&
                     !env().no_join32;
  if (try32) {
    const int forced = env().j32_variant;
    // CTAs/SM by table size (chunk capacity = 0.75 * slots): 4 -> 5.3k, 2 -> 10.8k, 1 -> 21.7k tuples
    const int variant = forced >= 0 ? forced : need <= 5300 ? 7 : need <= 10800 ? 3 : 10;
    const void* k;
    int threads, ctas;
    int W = 4;
#define J32V(T, B, M, WW) k = (const void*)rx_join32_kernel<T, B, M, WW>; threads = T; ctas = M; W = WW; break
    switch (variant) {  // measured at C3 (profiles/r01_join32_variants.jsonl): 7 is the fastest
      case 0: J32V(512, 8, 2, 4);
      case 2: J32V(512, 4, 3, 2);
      case 3: J32V(512, 8, 2, 2);
      case 5: J32V(256, 8, 4, 2);
      case 10: J32V(1024, 4, 1, 2);
      case 11: J32V(1024, 6, 1, 2);
      case 12: J32V(1024, 8, 1, 2);
      default: J32V(384, 4, 4, 2);
    }
#undef J32V
    // table: 4 slots per expected R tuple if shared memory allows `ctas` CTAs per SM with it
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    const size_t per_sm = 228 * 1024, reserved = 1024;
    size_t avail = per_sm / ctas - reserved - fa.sharedSizeBytes;
    if (avail > h->smem_optin - fa.sharedSizeBytes) avail = h->smem_optin - fa.sharedSizeBytes;
    u64 want = (n >> pl.bits) + 1;
    if (want < 64) want = 64;
    // want = expected R tuples per partition -> 4 slots each; buckets hold W slots in 8*W bytes
    u32 nbuckets = (u32)(want * 32 <= avail ? want * 4 / W : avail / (8 * W));
    nbuckets &= ~1u;  // key array = whole 16-byte words for either W
    const size_t smem = (size_t)nbuckets * 8 * W;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = grid_for(h, k, threads, smem);
    if ((u32)grid > nparts) grid = (int)nparts;
    // histogram-free layout: the partitions' ends are the scatter's final cursors
    const u64 *aR = h->rb[0].out.p, *aoR = opt ? h->rb[0].cursors.p : h->rb[0].offsets.p,
              *aS = h->rb[1].out.p + (u64)src_first * src_stride,
              *aoS = (opt ? h->rb[1].cursors.p : h->rb[1].offsets.p) + (u64)src_first * off_stride;
    u32 a_np = nparts, a_bits = pl.bits;
    u64* a_ctr = h->ctr;
    u64 a_capR = capR, a_capS = capS;
    // tasks = (partition, S slice): S partitions beyond 4x the mean are probed by several CTAs
    u64 slice = 4 * (m >> pl.bits);
    if (slice < 32768) slice = 32768;
    if (env().join_slice) slice = env().join_slice;
    if (nsrc > 1) slice = 1ull << 40;  // one task per partition: every source is probed by the CTA that built R_p
    slice &= ~1ull;
    const u64 max_tasks = (u64)nparts + m / slice + 1;
    if ((rc = ensure(h, h->tasks, max_tasks))) return rc;
    rx_tasks_kernel<<<1, 1024, 0, h->stream>>>(aoS, nparts, capS, slice, h->tasks.p, max_tasks, h->ctr);
    const u64* a_tasks = h->tasks.p;
    void* args[] = {&aR, &aoR, &aS, &aoS, &a_np, &a_bits, &nbuckets, &a_ctr, &a_capR, &a_capS, &a_tasks, &slice,
                    &nsrc, &src_stride, &off_stride};
    CK(cudaLaunchKernel(k, dim3(grid), dim3(threads), args, smem, h->stream));
    h->launches += 2;
  }
  {
    size_t smem = (size_t)pl.slots * 8;
    // <loads in flight per thread, CTAs/SM the register budget is set for>
    const int variant = env().join_variant;
    auto k = variant == 1 ? rx_join_kernel<4, 3> : variant == 2 ? rx_join_kernel<6, 3> : rx_join_kernel<8, 2>;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = grid_for(h, (const void*)k, kJoinThreads, smem);
    if ((u32)grid > nparts) grid = (int)nparts;
    k<<<grid, kJoinThreads, smem, h->stream>>>(h->rb[0].out.p, opt ? h->rb[0].cursors.p : h->rb[0].offsets.p,
                                               has_s ? h->rb[1].out.p + (u64)src_first * src_stride : nullptr,
                                               has_s ? (opt ? h->rb[1].cursors.p : h->rb[1].offsets.p) + (u64)src_first * off_stride
                                                     : nullptr, nparts,
                                               pl.bits, pl.slots, h->cfg.key_mask, h->ctr, try32 ? 1 : 0, capR, capS,
                                               nsrc, src_stride, off_stride);
    h->launches++;
    CK(cudaGetLastError());
  }
  return HJ_OK;
}

// materialisation request (hj_join_materialize): out may be null (count only)
struct MatReq {
  ulonglong2* out;
  u64 capacity;  // pairs
  u64 needed;    // pairs the join produces (set after the count pass)
  bool written;
};

int run_radix_join(hj_handle* h, const u64* dR, u64 n, bool pendR, const u64* dS, u64 m, bool pendS, hj_result* res,
                   bool pre1 = false, bool force_exact = false, const u64* offR1 = nullptr,
                   const u64* offS1 = nullptr, MatReq* mat = nullptr, const Pre1Regions* regR = nullptr,
                   const Pre1Regions* regS = nullptr, bool sortmerge = false) {
  RadixChoice ch;
  int rc;
  h->built = false;  // the counter block is about to be zeroed and (L2 join) the table reused: no probe after this
  if ((rc = choose_radix(h, n, dS != nullptr, pre1, (offR1 && offS1) || (regR && regS), force_exact, sortmerge, &ch))) {
    snprintf(h->err, sizeof(h->err), "radix plan: numPartitions must be a power of two <= 2^%u, passes 1|2, each of two passes <= 2^%u",
             kMaxTotalBits, kMaxBitsPerPass);
    return rc;
  }
  const RadixPlan& pl = ch.pl;
  if (pre1 && pl.passes != 2) return HJ_EINVAL;
  const u32 nparts = 1u << pl.bits;
  const u64 need = ch.need;
  const bool use_l2 = ch.use_l2, opt = ch.opt;
  if (sortmerge) {
    if (!dS || mat || pre1) return HJ_EINVAL;
  }
  const u64 capR = opt ? opt_region_cap(n, pl.bits) : 0, capS = opt ? opt_region_cap(m, pl.bits) : 0;
  if ((regR || regS) && !opt) return HJ_EINVAL;  // regions need the histogram-free 2-pass shared-memory plan
  if ((rc = ensure_rel(h, h->rb[0], n, pl, pre1, opt, regR ? regR->n_src : 1))) return rc;
  if (dS && (rc = ensure_rel(h, h->rb[1], m, pl, pre1, opt, regS ? regS->n_src : 1))) return rc;
  CK(cudaMemsetAsync(h->ctr, 0, C_COUNT * sizeof(u64), h->stream));
  if ((rc = wait_staged(h, E_CPR, pendR))) return rc;
  if (opt) rc = radix_partition_rel_opt(h, dR, n, h->rb[0], pl, E_RX, pre1 ? offR1 : nullptr, regR);
  else rc = radix_partition_rel(h, dR, n, h->rb[0], pl, E_RX, h->ctr + C_MAX_PART, pre1);
  if (rc) return rc;
  if (dS) {
    if ((rc = wait_staged(h, E_CPS, pendS))) return rc;
    if (opt) rc = radix_partition_rel_opt(h, dS, m, h->rb[1], pl, E_RX + 5, pre1 ? offS1 : nullptr, regS);
    else rc = radix_partition_rel(h, dS, m, h->rb[1], pl, E_RX + 5, nullptr, pre1);
    if (rc) return rc;
  }
  mark(h, E_JOIN0);
  if (mat) {
    // count / scan / write over the partitions (rx_joinmat.cuh); shared-memory partitions of 32-bit keys only
    if (use_l2 || !dS || pl.bits < 1 || (h->cfg.key_mask & 0xffffffffull) != 0xffffffffull) {
      snprintf(h->err, sizeof(h->err), "materialise: needs shared-memory partitions (fan-out 2^%u too low for |R|) "
               "and 32-bit keys", pl.bits);
      return HJ_EINVAL;
    }
    if ((rc = ensure(h, h->mat_cnt, (u64)nparts + 1))) return rc;
    const void* kc = (const void*)rx_join32_mat_kernel<1024, 2, false>;
    const void* kw = (const void*)rx_join32_mat_kernel<1024, 2, true>;
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, kw));
    size_t avail = h->smem_optin - fa.sharedSizeBytes - 1024;
    u64 want = (n >> pl.bits) + 1;
    if (want < 64) want = 64;
    u32 nbuckets = (u32)(want * 32 <= avail ? want * 2 : avail / 16);  // 4 slots per expected R tuple, W = 2
    nbuckets &= ~1u;
    const size_t smem = (size_t)nbuckets * 16;
    CK(cudaFuncSetAttribute(kc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(kw, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = grid_for(h, kw, 1024, smem);
    if ((u32)grid > nparts) grid = (int)nparts;
    const u64 *aR = h->rb[0].out.p, *aoR = opt ? h->rb[0].cursors.p : h->rb[0].offsets.p, *aS = h->rb[1].out.p,
              *aoS = opt ? h->rb[1].cursors.p : h->rb[1].offsets.p;
    u32 a_np = nparts, a_bits = pl.bits;
    u64* a_ctr = h->ctr;
    u64 a_capR = capR, a_capS = capS;
    u64* a_cnt = h->mat_cnt.p;
    ulonglong2* a_out = mat->out;
    void* args[] = {&aR, &aoR, &aS, &aoS, &a_np, &a_bits, &nbuckets, &a_ctr, &a_capR, &a_capS, &a_cnt, &a_out};
    CK(cudaLaunchKernel(kc, dim3(grid), dim3(1024), args, smem, h->stream));
    rx_mat_scan_kernel<<<1, 1024, 0, h->stream>>>(h->mat_cnt.p, nparts, h->ctr + C_MAT_TOTAL);
    h->launches += 2;
    if ((rc = fetch_counters(h))) return rc;
    if (opt && h->h_ctr[C_OPT_OVF]) {
      h->no_optimistic = true;
      return run_radix_join(h, dR, n, false, dS, m, false, res, pre1, true, nullptr, nullptr, mat);
    }
    if (h->h_ctr[C_KEY_HI]) {
      snprintf(h->err, sizeof(h->err), "materialise: keys wider than 32 bits");
      return HJ_EINVAL;
    }
    mat->needed = h->h_ctr[C_MAT_TOTAL];
    mat->written = false;
    if (mat->out && mat->needed <= mat->capacity) {
      CK(cudaLaunchKernel(kw, dim3(grid), dim3(1024), args, smem, h->stream));
      h->launches++;
      mat->written = true;
    }
    mark(h, E_JOIN);
    mark(h, E_END);
    if ((rc = fetch_counters(h))) return rc;
    if (res) {
      fill_common(h, res);
      if (!mat->written) res->matches = mat->needed;
      res->path_taken = HJ_PATH_RADIX;
      res->radix_bits = pl.bits;
      res->passes = pl.passes;
      // two join passes read both relations, the write pass emits 16 B per pair
      res->hbm_bytes = 8ull * (2 * pl.passes + 2) * (n + m) + (mat->written ? 16ull * mat->needed : 0ull);
    }
    if (mat->out && !mat->written) {
      snprintf(h->err, sizeof(h->err), "materialise: %llu pairs do not fit the buffer of %llu", mat->needed, mat->capacity);
      return HJ_EOVERFLOW;
    }
    return HJ_OK;
  }
  if (sortmerge) {
    // per-partition sort + merge (rx_sortmerge.cuh) on either layout
    // register sort (default) or the shared-memory-only network (HJ_SM_VARIANT=0, the A/B record)
    const int variant = env().sm_variant;
    const u32 tileB = 4096;
    u32 capA = variant ? 4096 : 1024;
    while (capA < 8192 && capA < need) capA <<= 1;  // larger R partitions are merged chunk by chunk
    // encoding of the register sort by key mask (rx_sortmerge.cuh); HJ_SM_MODE=0 forces the generic order
    int mode = h->cfg.key_mask == 0xffffffffull ? 1 : h->cfg.key_mask == ~0ull ? 2 : 0;
    if (env().sm_mode == 0) mode = 0;
    // HJ_SM_CTAS=3 (opt-in, the A/B record): the scan of R's histogram has left the largest R partition in the
    // counters; when none exceeds 4096 tuples (dense keys split evenly by their low bits) A needs 32 KB only and
    // three 40-register CTAs fit an SM.  Measured at C3: 27.61 ms vs 27.73 ms with two -- the kernel is bound by
    // the ALU pipe, not by latency -- so the default stays at two and saves the counter read-back.
    const int want_ctas = env().sm_ctas;
    int ctas = 2;
    if (variant && want_ctas == 3 && !opt) {
      if ((rc = fetch_counters(h))) return rc;
      if (h->h_ctr[C_MAX_PART] <= 4096) { capA = 4096; ctas = 3; }
    }
    const size_t smem = (size_t)(capA + tileB) * 8;
    const void* k2 = mode == 1 ? (const void*)rx_sortmerge_regs_kernel<1, 2>
                     : mode == 2 ? (const void*)rx_sortmerge_regs_kernel<2, 2> : (const void*)rx_sortmerge_regs_kernel<0, 2>;
    const void* k3 = mode == 1 ? (const void*)rx_sortmerge_regs_kernel<1, 3>
                     : mode == 2 ? (const void*)rx_sortmerge_regs_kernel<2, 3> : (const void*)rx_sortmerge_regs_kernel<0, 3>;
    const void* k = !variant ? (const void*)rx_sortmerge_kernel : ctas == 3 ? k3 : k2;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = grid_for(h, k, kSmThreads, smem);
    if ((u32)grid > nparts) grid = (int)nparts;
    if (variant) {
      // histogram-free layout: the partitions' ends are the scatter's final cursors
      const u64 *aR = h->rb[0].out.p, *aoR = opt ? h->rb[0].cursors.p : h->rb[0].offsets.p, *aS = h->rb[1].out.p,
                *aoS = opt ? h->rb[1].cursors.p : h->rb[1].offsets.p;
      u32 a_np = nparts;
      u64 a_km = h->cfg.key_mask, a_capR = capR, a_capS = capS;
      u64* a_ctr = h->ctr;
      void* args[] = {&aR, &aoR, &aS, &aoS, &a_np, &capA, &a_km, &a_ctr, &a_capR, &a_capS};
      CK(cudaLaunchKernel(k, dim3(grid), dim3(kSmThreads), args, smem, h->stream));
    } else
      rx_sortmerge_kernel<<<grid, kSmThreads, smem, h->stream>>>(
          h->rb[0].out.p, opt ? h->rb[0].cursors.p : h->rb[0].offsets.p, h->rb[1].out.p,
          opt ? h->rb[1].cursors.p : h->rb[1].offsets.p, nparts, capA, tileB, h->cfg.key_mask, h->ctr, capR, capS);
    h->launches++;
    CK(cudaGetLastError());
    mark(h, E_JOIN);
    mark(h, E_END);
    if ((rc = fetch_counters(h))) return rc;
    if (opt && h->h_ctr[C_OPT_OVF]) {
      // a region overflowed (skewed partition sizes): nothing was lost -- redo on the exact layout
      h->no_optimistic = true;
      return run_radix_join(h, dR, n, false, dS, m, false, res, false, true, nullptr, nullptr, nullptr, nullptr, nullptr,
                            true);
    }
    if (res) {
      fill_common(h, res);
      res->path_taken = HJ_PATH_SORTMERGE;
      res->radix_bits = pl.bits;
      res->passes = pl.passes;
      res->hbm_bytes = 8ull * (2 * pl.passes + 1) * (n + m);  // as the radix join: the sort is on chip
    }
    return HJ_OK;
  }
  if (use_l2) {
    Digit full;
    memset(&full, 0, sizeof(full));
    full.key_mask = h->cfg.key_mask;
    if (pl.passes == 2) { full.mode = 1; full.b1 = pl.b1; full.b2 = pl.b2; }
    else { full.mode = 0; full.shift = 0; full.bits = pl.bits; }
    const u64 tslots = l2_table_slots(n, nparts);
    if ((rc = ensure(h, h->table, tslots))) return rc;
    h->built = false;
    CK(cudaMemsetAsync(h->table.p, 0, tslots * sizeof(u64), h->stream));
    constexpr int IPT = 4;
    int gb = grid_for(h, (const void*)rx_l2_build_kernel<IPT>, kL2Threads, 0);
    const int narrow_possible = (h->cfg.key_mask & 0xffffffffull) == 0xffffffffull ? 1 : 0;
    rx_l2_build_kernel<IPT><<<gb, kL2Threads, 0, h->stream>>>(h->rb[0].out.p, n, h->rb[0].offsets.p, full, pl.bits,
                                                              h->table.p, h->ctr + C_TILE0);
    int gp = grid_for(h, (const void*)rx_l2_probe_kernel<IPT, true>, kL2Threads, 0);
    rx_l2_probe_kernel<IPT, true><<<gp, kL2Threads, 0, h->stream>>>(h->rb[1].out.p, m, h->rb[0].offsets.p, full,
                                                                    pl.bits, h->table.p, h->ctr, narrow_possible);
    gp = grid_for(h, (const void*)rx_l2_probe_kernel<IPT, false>, kL2Threads, 0);
    rx_l2_probe_kernel<IPT, false><<<gp, kL2Threads, 0, h->stream>>>(h->rb[1].out.p, m, h->rb[0].offsets.p, full,
                                                                     pl.bits, h->table.p, h->ctr, narrow_possible);
    h->launches += 3;
    CK(cudaGetLastError());
  }
  if (!use_l2 && (rc = launch_partition_join(h, pl, need, opt, n, dS != nullptr, m, capR, capS))) return rc;
  mark(h, E_JOIN);
  mark(h, E_END);
  if ((rc = fetch_counters(h))) return rc;
  if (opt && h->h_ctr[C_OPT_OVF]) {
    // a region overflowed (skewed partition sizes): nothing was lost -- redo on the exact path
    if (regR) {  // the exact path needs contiguous pass-1 partitions: the caller redoes the exchange
      snprintf(h->err, sizeof(h->err), "join_regions: a pass-2 region overflowed; redo with hj_peer_hist + hj_join_pass1");
      return HJ_EOVERFLOW;
    }
    h->no_optimistic = true;
    return run_radix_join(h, dR, n, false, dS, m, false, res, pre1, true);
  }
  if (res) {
    fill_common(h, res);
    res->path_taken = HJ_PATH_RADIX;
    res->radix_bits = pl.bits;
    res->passes = pl.passes;
    res->hbm_bytes = 8ull * (2 * pl.passes + 1) * (n + (dS ? m : 0));  // 8*(2P+1) per tuple
  }
  return HJ_OK;
}

int run_global_join(hj_handle* h, const u64* dR, u64 n, const u64* dS, u64 m, bool pendS, bool nocc,
                    bool adaptive_prefix, u64 stride, u64 slen, hj_result* res) {
  int rc;
  const u32 ipt = ipt_for(h->cfg.transaction_size);
  if (!adaptive_prefix) {
    if ((rc = launch_build(h, dR, whole(n), nocc, false, ipt))) return rc;
  } else {
    if ((rc = run_build_rest(h, dR, n, stride, slen, nocc, ipt))) return rc;
  }
  if ((rc = run_overflow(h))) return rc;
  mark(h, E_BUILD);
  h->built = true;
  if (dS && m) {
    if ((rc = wait_staged(h, E_CPS, pendS))) return rc;
    mark(h, E_PROBE0);
    if ((rc = launch_probe(h, dS, m, ipt))) return rc;
    mark(h, E_PROBE);
  }
  mark(h, E_END);
  if ((rc = fetch_counters(h))) return rc;
  if (res) {
    fill_common(h, res);
    res->path_taken = HJ_PATH_GLOBAL;
    res->hbm_bytes = 16ull * n + 16ull * (dS ? m : 0);  // 16 B per R tuple + 16 B per S tuple
  }
  return HJ_OK;
}

}  // namespace

// ---- 16-byte tuples (cfg.tuple_bytes == 16; mc's KEY_8B tuple, mc/src/types.h:22-37) -----------------------
namespace {

bool wide(const hj_handle* h) { return h->cfg.tuple_bytes == 16; }

// partition a relation of n 16-byte tuples by pl (exact layout): rb.out = partitions, rb.offsets = 2^bits + 1 offsets
int wide_partition_rel(hj_handle* h, const tup16* in, u64 n, RelBufs& rb, const RadixPlan& pl, int evb) {
  const u32 nb = 1u << pl.bits, F1 = 1u << pl.b1, F2 = 1u << pl.b2;
  constexpr u32 tile = kWScThreads * kWScIPT;
  Digit full;
  memset(&full, 0, sizeof(full));
  full.key_mask = h->cfg.key_mask;
  if (pl.passes == 2) { full.mode = 1; full.b1 = pl.b1; full.b2 = pl.b2; }
  else { full.mode = 0; full.shift = 0; full.bits = pl.bits; }
  if (evb >= 0) mark(h, evb);
  CK(cudaMemsetAsync(rb.ghist.p, 0, nb * sizeof(u64), h->stream));
  {
    const size_t smem = (size_t)nb * 4;
    CK(cudaFuncSetAttribute(w_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = grid_for(h, (const void*)w_hist_kernel, kWHistThreads, smem);
    const u64 tiles = (n + (u64)kWHistThreads * kWHistIPT - 1) / ((u64)kWHistThreads * kWHistIPT);
    if ((u64)grid > tiles) grid = (int)(tiles ? tiles : 1);
    w_hist_kernel<<<grid, kWHistThreads, smem, h->stream>>>(in, n, full, nb, rb.ghist.p, h->ctr);
    h->launches++;
  }
  if (evb >= 0) mark(h, evb + 1);
  rx_scan_kernel<<<1, kScanThreads, 0, h->stream>>>(rb.ghist.p, nb, rb.offsets.p, rb.cursors.p, pl.passes == 2 ? F2 : nb,
                                                    rb.poff.p, rb.pcur.p, rb.ptiles.p, rb.one.p, tile, h->ctr + C_MAX_PART, n);
  h->launches++;
  CK(cudaGetLastError());
  if (evb >= 0) mark(h, evb + 2);
  auto scatter = [&](const tup16* src, tup16* dst, const Digit& dg, u32 F, u32 nparent, const u64* poff, const u64* ptiles,
                     u64* cursors, u64 max_tiles) {
    const size_t smem = (size_t)tile * 16 + (size_t)F * 16;
    if (cudaFuncSetAttribute(w_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return HJ_ECUDA;
    int grid = grid_for(h, (const void*)w_scatter_kernel, kWScThreads, smem);
    if ((u64)grid > max_tiles) grid = (int)(max_tiles ? max_tiles : 1);
    w_scatter_kernel<<<grid, kWScThreads, smem, h->stream>>>(src, dst, dg, F, nparent, poff, ptiles, cursors);
    h->launches++;
    return cudaGetLastError() == cudaSuccess ? HJ_OK : HJ_ECUDA;
  };
  const u64 max_tiles = (n + tile - 1) / tile;
  tup16* out = reinterpret_cast<tup16*>(rb.out.p);
  tup16* tmp = reinterpret_cast<tup16*>(rb.tmp.p);
  int rc;
  if (pl.passes == 1) {
    if ((rc = scatter(in, out, full, nb, 1, rb.poff.p, rb.ptiles.p, rb.cursors.p, max_tiles))) return rc;
    if (evb >= 0) { mark(h, evb + 3); mark(h, evb + 4); }
  } else {
    Digit d1, d2;
    memset(&d1, 0, sizeof(d1));
    d1.key_mask = h->cfg.key_mask; d1.mode = 0; d1.shift = 0; d1.bits = pl.b1;
    d2 = d1; d2.shift = pl.b1; d2.bits = pl.b2;
    if ((rc = scatter(in, tmp, d1, F1, 1, rb.one.p, rb.one.p + 2, rb.pcur.p, max_tiles))) return rc;
    if (evb >= 0) mark(h, evb + 3);
    if ((rc = scatter(tmp, out, d2, F2, F1, rb.poff.p, rb.ptiles.p, rb.cursors.p, max_tiles + F1))) return rc;
    if (evb >= 0) mark(h, evb + 4);
  }
  return HJ_OK;
}

int run_wide_radix_join(hj_handle* h, const u64* dR, u64 n, bool pendR, const u64* dS, u64 m, bool pendS, hj_result* res) {
  if (!dS) {
    snprintf(h->err, sizeof(h->err), "16-byte tuples: the radix join needs both relations");
    return HJ_EINVAL;
  }
  RadixPlan pl;
  int rc;
  h->built = false;
  if ((rc = plan_radix(h, n, &pl, 4096))) return rc;  // R partitions of <= 4096 tuples: one chunk of an 8192-slot table
  if (pl.passes == 1 && pl.bits > kMaxBitsPerPass) return HJ_EINVAL;
  if ((rc = ensure_rel(h, h->rb[0], 2 * n, pl))) return rc;
  if ((rc = ensure_rel(h, h->rb[1], 2 * m, pl))) return rc;
  CK(cudaMemsetAsync(h->ctr, 0, C_COUNT * sizeof(u64), h->stream));
  if ((rc = wait_staged(h, E_CPR, pendR))) return rc;
  if ((rc = wide_partition_rel(h, reinterpret_cast<const tup16*>(dR), n, h->rb[0], pl, E_RX))) return rc;
  if ((rc = wait_staged(h, E_CPS, pendS))) return rc;
  if ((rc = wide_partition_rel(h, reinterpret_cast<const tup16*>(dS), m, h->rb[1], pl, E_RX + 5))) return rc;
  mark(h, E_JOIN0);
  {
    const u32 nparts = 1u << pl.bits;
    u64 per = (n >> pl.bits) + 1;
    u32 slots = 1024;
    while (slots < 8192 && slots < 4 * per) slots <<= 1;
    const size_t smem = (size_t)slots * 16;
    CK(cudaFuncSetAttribute(w_join_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = grid_for(h, (const void*)w_join_kernel, kWJoinThreads, smem);
    if ((u32)grid > nparts) grid = (int)nparts;
    w_join_kernel<<<grid, kWJoinThreads, smem, h->stream>>>(reinterpret_cast<const tup16*>(h->rb[0].out.p), h->rb[0].offsets.p,
                                                            reinterpret_cast<const tup16*>(h->rb[1].out.p), h->rb[1].offsets.p,
                                                            nparts, pl.bits, slots, h->cfg.key_mask, h->ctr);
    h->launches++;
    CK(cudaGetLastError());
  }
  mark(h, E_JOIN);
  mark(h, E_END);
  if ((rc = fetch_counters(h))) return rc;
  if (res) {
    fill_common(h, res);
    res->path_taken = HJ_PATH_RADIX;
    res->radix_bits = pl.bits;
    res->passes = pl.passes;
    res->hbm_bytes = 16ull * (3 * pl.passes + 1) * (n + m);  // per pass: histogram read (exact layout) + read + write; join read
  }
  return HJ_OK;
}

int wide_prepare_table(hj_handle* h, u64 n) {
  u32 scale = h->cfg.scale_output ? h->cfg.scale_output : 2;
  u64 slots = pow2_ceil_u64(n * scale < 128 ? 128 : n * scale);
  if (slots > (1ull << 32)) return HJ_EINVAL;
  int rc;
  if ((rc = ensure(h, h->table, 2 * slots))) return rc;
  if ((rc = ensure(h, h->ovf_table, 2 * slots))) return rc;
  if ((rc = ensure(h, h->ovf_list, 2 * (n ? n : 1)))) return rc;
  if ((rc = ensure(h, h->claim, slots / 64))) return rc;
  if (slots != h->slots) h->built = false;
  h->slots = slots;
  h->mask = slots - 1;
  return HJ_OK;
}

int wide_build(hj_handle* h, const u64* dR, u64 n) {
  int rc;
  if ((rc = zero_async(h, h->table.p, 2 * h->slots))) return rc;
  CK(cudaMemsetAsync(h->claim.p, 0, h->slots / 8, h->stream));
  mark(h, E_CLEAR);
  const u64 chunks = (n + kGtThreads * kWGtIPT - 1) / (kGtThreads * kWGtIPT);
  int grid = grid_for(h, (const void*)gw_build_kernel, kGtThreads, 0);
  if ((u64)grid > chunks) grid = (int)(chunks ? chunks : 1);
  gw_build_kernel<<<grid, kGtThreads, 0, h->stream>>>(reinterpret_cast<const tup16*>(dR), n, reinterpret_cast<tup16*>(h->table.p),
                                                      reinterpret_cast<u32*>(h->claim.p), (u32)h->mask, h->cfg.key_mask,
                                                      h->cfg.probe_length, reinterpret_cast<tup16*>(h->ovf_list.p), h->ctr);
  gw_ovf_clear_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>(reinterpret_cast<tup16*>(h->ovf_table.p), h->ctr, h->ovf_table.cap / 2);
  gw_ovf_insert_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>(reinterpret_cast<const tup16*>(h->ovf_list.p),
                                                               reinterpret_cast<tup16*>(h->ovf_table.p), h->ctr, h->cfg.key_mask,
                                                               h->ovf_table.cap / 2);
  h->launches += 3;
  CK(cudaGetLastError());
  mark(h, E_BUILD);
  h->built = true;
  return HJ_OK;
}

int wide_probe(hj_handle* h, const u64* dS, u64 m) {
  mark(h, E_PROBE0);
  if (m) {
    int grid = grid_for(h, (const void*)gw_probe_kernel, kGtThreads, 0);
    gw_probe_kernel<<<grid, kGtThreads, 0, h->stream>>>(reinterpret_cast<const tup16*>(dS), m, reinterpret_cast<const tup16*>(h->table.p),
                                                        (u32)h->mask, h->cfg.key_mask, h->cfg.probe_length,
                                                        reinterpret_cast<const tup16*>(h->ovf_table.p), h->ovf_table.cap / 2, h->ctr);
    h->launches++;
    CK(cudaGetLastError());
  }
  mark(h, E_PROBE);
  return HJ_OK;
}

// hj_join / hj_build / hj_probe on 16-byte tuples
int wide_join(hj_handle* h, const u64* R, u64 n, const u64* S, u64 m, hj_result* res, bool build, bool probe) {
  int rc;
  const u64 *dR = nullptr, *dS = nullptr;
  bool pendR = false, pendS = false;
  const int algo = h->cfg.algo;
  if (algo != HJ_ALGO_RADIX && algo != HJ_ALGO_ATOMIC) {
    snprintf(h->err, sizeof(h->err), "16-byte tuples: HJ_ALGO_RADIX and HJ_ALGO_ATOMIC only");
    return HJ_EINVAL;
  }
  mark(h, E_T0);
  mark(h, E_CP0, h->copy_stream);
  if (build && (rc = stage(h, R, 2 * n, h->stageR, &dR, E_CPR, &pendR))) return rc;
  if (probe && (rc = stage(h, S, 2 * m, h->stageS, &dS, E_CPS, &pendS))) return rc;
  if (algo == HJ_ALGO_RADIX) {
    if (!build || !probe) return HJ_ESTATE;
    return run_wide_radix_join(h, dR, n, pendR, dS, m, pendS, res);
  }
  if (build) {
    if ((rc = wide_prepare_table(h, n))) return rc;
    CK(cudaMemsetAsync(h->ctr, 0, C_COUNT * sizeof(u64), h->stream));
    if ((rc = wait_staged(h, E_CPR, pendR))) return rc;
    if ((rc = wide_build(h, dR, n))) return rc;
  } else {
    if (!h->built) return HJ_ESTATE;
    if ((rc = clear_probe_counters(h))) return rc;
  }
  if (probe) {
    if ((rc = wait_staged(h, E_CPS, pendS))) return rc;
    if ((rc = wide_probe(h, dS, m))) return rc;
  }
  mark(h, E_END);
  if ((rc = fetch_counters(h))) return rc;
  if (res) {
    fill_common(h, res);
    res->path_taken = HJ_PATH_GLOBAL;
    res->hbm_bytes = 32ull * (build ? n : 0) + 32ull * (probe ? m : 0);  // 16 B read + 16 B slot per tuple
  }
  return HJ_OK;
}

}  // namespace

void xchg_free(hj_handle* h);  // defined with the pipelined exchange at the end of this file

// ===========================================================================
// C ABI
// ===========================================================================
extern "C" {

int hj_abi_version(void) { return HJ_ABI_VERSION; }

int hj_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

const char* hj_strerror(int code) {
  switch (code) {
    case HJ_OK: return "ok";
    case HJ_EINVAL: return "invalid argument";
    case HJ_ENOMEM: return "out of memory";
    case HJ_ECUDA: return "CUDA error (no usable device, or a runtime failure; see hj_last_error)";
    case HJ_EOVERFLOW: return "partition overflow";
    case HJ_ESTATE: return "call order violated";
    default: return "unknown error";
  }
}

const char* hj_last_error(const hj_handle* h) { return h ? h->err : "null handle"; }

void hj_config_default(hj_config* cfg) {
  if (!cfg) return;
  memset(cfg, 0, sizeof(*cfg));
  cfg->struct_size = sizeof(hj_config);
  cfg->algo = HJ_ALGO_ATOMIC;
  cfg->device = 0;
  cfg->probe_length = 4;      // main.cpp:81
  cfg->scale_output = 2;      // main.cpp:84
  cfg->num_partitions = 0;
  cfg->passes = 0;
  cfg->transaction_size = 16; // main.cpp:83
  cfg->enable_probe = 1;
  cfg->sample_rounds = 5;     // HTMHashBuild.hpp:48-52 (HTM_SWITCH)
  cfg->key_mask = ~0ull;
  cfg->adaptive_threshold = 0.0;
}

int hj_create(const hj_config* cfg, hj_handle** out) {
  if (!cfg || !out || cfg->struct_size != sizeof(hj_config)) return HJ_EINVAL;
  if (cfg->algo < HJ_ALGO_ATOMIC || cfg->algo > HJ_ALGO_SORTMERGE) return HJ_EINVAL;
  if (cfg->probe_length == 0 || cfg->probe_length > 1024) return HJ_EINVAL;
  if (cfg->tuple_bytes != 0 && cfg->tuple_bytes != 8 && cfg->tuple_bytes != 16) return HJ_EINVAL;
  if (cfg->tuple_bytes == 16 && cfg->algo != HJ_ALGO_RADIX && cfg->algo != HJ_ALGO_ATOMIC) return HJ_EINVAL;
  int ndev = hj_device_count();
  if (ndev <= 0 || cfg->device < 0 || cfg->device >= ndev) return HJ_ECUDA;
  hj_handle* h = new (std::nothrow) hj_handle();
  if (!h) return HJ_ENOMEM;
  g_env = read_env();
  h->cfg = *cfg;
  if (h->cfg.key_mask == 0) h->cfg.key_mask = ~0ull;
  h->device = cfg->device;
  h->err[0] = 0;
  for (int i = 0; i < E_COUNT; i++) h->ev_set[i] = false;
  auto fail = [&](int code) { hj_destroy(h); return code; };
  if (cudaSetDevice(h->device) != cudaSuccess) return fail(HJ_ECUDA);
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, h->device) != cudaSuccess) return fail(HJ_ECUDA);
  h->sm_count = prop.multiProcessorCount;
  h->l2_bytes = prop.l2CacheSize ? (u64)prop.l2CacheSize : (126ull << 20);
  h->smem_optin = prop.sharedMemPerBlockOptin;
  if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) return fail(HJ_ECUDA);
  if (cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking) != cudaSuccess) return fail(HJ_ECUDA);
  for (int i = 0; i < E_COUNT; i++)
    if (cudaEventCreate(&h->ev[i]) != cudaSuccess) return fail(HJ_ECUDA);
  if (cudaMalloc(&h->ctr, C_COUNT * sizeof(u64)) != cudaSuccess) return fail(HJ_ENOMEM);
  if (cudaMallocHost(&h->h_ctr, C_COUNT * sizeof(u64)) != cudaSuccess) return fail(HJ_ENOMEM);
  cudaMemset(h->ctr, 0, C_COUNT * sizeof(u64));
  *out = h;
  return HJ_OK;
}

void hj_destroy(hj_handle* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  xchg_free(h);
  DevBuf* bufs[] = {&h->table, &h->ovf_table, &h->ovf_list, &h->claim, &h->ovfmap, &h->stageR, &h->stageS, &h->flush,
                    &h->peer_ptrs, &h->tasks, &h->mat_cnt};
  for (DevBuf* b : bufs)
    if (b->p) cudaFree(b->p);
  for (RelBufs& rb : h->rb) {
    DevBuf* rbufs[] = {&rb.tmp, &rb.out, &rb.ghist, &rb.offsets, &rb.cursors, &rb.poff, &rb.pcur, &rb.ptiles, &rb.one, &rb.pend};
    for (DevBuf* b : rbufs)
      if (b->p) cudaFree(b->p);
  }
  if (h->ctr) cudaFree(h->ctr);
  if (h->h_ctr) cudaFreeHost(h->h_ctr);
  for (int i = 0; i < E_COUNT; i++)
    if (h->ev[i]) cudaEventDestroy(h->ev[i]);
  if (h->stream) cudaStreamDestroy(h->stream);
  if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
  delete h;
}

int hj_reserve(hj_handle* h, uint64_t max_r, uint64_t max_s) {
  if (!h) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  int rc;
  const bool sortmerge = h->cfg.algo == HJ_ALGO_SORTMERGE;
  if (h->cfg.algo != HJ_ALGO_RADIX && !sortmerge) {
    if ((rc = prepare_table(h, max_r))) return rc;
  }
  if (h->cfg.algo == HJ_ALGO_RADIX || h->cfg.algo == HJ_ALGO_ADAPTIVE || sortmerge) {
    // exactly what run_radix_join will pick for these sizes: layout (histogram-free regions or exact),
    // join tasks, the L2 table when the partitions are too large for shared memory
    const bool has_s = max_s && h->cfg.enable_probe;
    RadixChoice ch;
    if ((rc = choose_radix(h, max_r, has_s, false, false, false, sortmerge, &ch))) return rc;
    if ((rc = ensure_rel(h, h->rb[0], max_r, ch.pl, false, ch.opt))) return rc;
    if (has_s && (rc = ensure_rel(h, h->rb[1], max_s, ch.pl, false, ch.opt))) return rc;
    if (has_s) {
      u64 slice = 4 * (max_s >> ch.pl.bits);
      if (slice < 32768) slice = 32768;
      if ((rc = ensure(h, h->tasks, (1ull << ch.pl.bits) + max_s / slice + 1))) return rc;
      if ((rc = ensure(h, h->mat_cnt, (1ull << ch.pl.bits) + 1))) return rc;
    }
    if (ch.use_l2 && (rc = ensure(h, h->table, l2_table_slots(max_r, 1ull << ch.pl.bits)))) return rc;
  }
  return HJ_OK;
}

uint64_t hj_alloc_count(const hj_handle* h) { return h ? h->allocs : 0; }

int hj_reserve_staging(hj_handle* h, uint64_t max_r, uint64_t max_s) {
  if (!h) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  int rc;
  if (max_r && (rc = ensure(h, h->stageR, max_r + (max_r & 1)))) return rc;
  if (max_s && (rc = ensure(h, h->stageS, max_s + (max_s & 1)))) return rc;
  return HJ_OK;
}

int hj_build(hj_handle* h, const uint64_t* R, uint64_t n, hj_result* res) {
  if (!h || !R || !n) return HJ_EINVAL;
  if (h->cfg.algo == HJ_ALGO_RADIX || h->cfg.algo == HJ_ALGO_SORTMERGE) return HJ_ESTATE;  // no standalone table
  CK(cudaSetDevice(h->device));
  if (res) memset(res, 0, sizeof(*res));
  reset_events(h);
  if (wide(h)) return wide_join(h, (const u64*)R, n, nullptr, 0, res, true, false);
  int rc;
  if ((rc = prepare_table(h, n))) return rc;
  mark(h, E_T0);
  CK(cudaMemsetAsync(h->ctr, 0, C_COUNT * sizeof(u64), h->stream));
  if ((rc = clear_table(h))) return rc;
  mark(h, E_CLEAR);
  const u64* dR = nullptr;
  bool pendR = false;
  mark(h, E_CP0, h->copy_stream);
  if ((rc = stage(h, (const u64*)R, n, h->stageR, &dR, E_CPR, &pendR, false))) return rc;
  if ((rc = wait_staged(h, E_CPR, pendR))) return rc;
  return run_global_join(h, dR, n, nullptr, 0, false, h->cfg.algo == HJ_ALGO_NOCC, false, 0, 0, res);
}

int hj_probe(hj_handle* h, const uint64_t* S, uint64_t m, hj_result* res) {
  if (!h || (!S && m)) return HJ_EINVAL;
  if (!h->built) return HJ_ESTATE;
  CK(cudaSetDevice(h->device));
  if (res) memset(res, 0, sizeof(*res));
  reset_events(h);
  if (wide(h)) return wide_join(h, nullptr, 0, (const u64*)S, m, res, false, true);
  int rc;
  const u64* dS = nullptr;
  bool pendS = false;
  mark(h, E_CP0, h->copy_stream);
  if ((rc = stage(h, (const u64*)S, m, h->stageS, &dS, E_CPS, &pendS, false))) return rc;
  if ((rc = wait_staged(h, E_CPS, pendS))) return rc;
  mark(h, E_T0);
  if ((rc = clear_probe_counters(h))) return rc;
  mark(h, E_PROBE0);
  if (m && (rc = launch_probe(h, dS, m, ipt_for(h->cfg.transaction_size)))) return rc;
  mark(h, E_PROBE);
  mark(h, E_END);
  if ((rc = fetch_counters(h))) return rc;
  if (res) {
    fill_common(h, res);
    res->path_taken = HJ_PATH_GLOBAL;
    res->hbm_bytes = 16ull * m;
  }
  return HJ_OK;
}

int hj_sample(hj_handle* h, const uint64_t* R, uint64_t n, hj_result* res) {
  if (!h || !R || !n) return HJ_EINVAL;
  if (h->cfg.algo == HJ_ALGO_RADIX || h->cfg.algo == HJ_ALGO_SORTMERGE) return HJ_ESTATE;
  CK(cudaSetDevice(h->device));
  if (res) memset(res, 0, sizeof(*res));
  reset_events(h);
  int rc;
  if ((rc = prepare_table(h, n))) return rc;
  mark(h, E_T0);
  CK(cudaMemsetAsync(h->ctr, 0, C_COUNT * sizeof(u64), h->stream));
  if ((rc = clear_table(h))) return rc;
  mark(h, E_CLEAR);
  const u64* dR = nullptr;
  bool pendR = false;
  mark(h, E_CP0, h->copy_stream);
  if ((rc = stage(h, (const u64*)R, n, h->stageR, &dR, E_CPR, &pendR, false))) return rc;
  if ((rc = wait_staged(h, E_CPR, pendR))) return rc;
  u64 stride, slen;
  if ((rc = run_sample(h, dR, n, &stride, &slen))) return rc;
  mark(h, E_SAMPLE);
  mark(h, E_END);
  if ((rc = fetch_counters(h))) return rc;
  h->built = false;
  if (res) {
    fill_common(h, res);
    res->us_build = 0.f;
    res->contention = contention_of(h);
    res->path_taken = HJ_PATH_NONE;
  }
  return HJ_OK;
}

int hj_join(hj_handle* h, const uint64_t* R, uint64_t n, const uint64_t* S, uint64_t m, hj_result* res) {
  if (!h || !R || !n) return HJ_EINVAL;
  if (!S) m = 0;
  if (!h->cfg.enable_probe || m == 0) { S = nullptr; m = 0; }
  CK(cudaSetDevice(h->device));
  if (res) memset(res, 0, sizeof(*res));
  reset_events(h);
  if (wide(h)) return wide_join(h, (const u64*)R, n, (const u64*)S, m, res, true, S != nullptr);
  int rc;
  const int algo = h->cfg.algo;
  const u64 *dR = nullptr, *dS = nullptr;
  bool pendR = false, pendS = false;
  if (algo == HJ_ALGO_RADIX || algo == HJ_ALGO_SORTMERGE) {
    if (algo == HJ_ALGO_SORTMERGE && !S) return HJ_EINVAL;  // a merge needs both sides
    mark(h, E_T0);
    mark(h, E_CP0, h->copy_stream);
    if ((rc = stage(h, (const u64*)R, n, h->stageR, &dR, E_CPR, &pendR))) return rc;
    if ((rc = stage(h, (const u64*)S, m, h->stageS, &dS, E_CPS, &pendS))) return rc;
    return run_radix_join(h, dR, n, pendR, dS, m, pendS, res, false, false, nullptr, nullptr, nullptr, nullptr, nullptr,
                          algo == HJ_ALGO_SORTMERGE);
  }
  // the table clear needs no input: enqueue it first so it runs under the R copy
  if ((rc = prepare_table(h, n))) return rc;
  mark(h, E_T0);
  CK(cudaMemsetAsync(h->ctr, 0, C_COUNT * sizeof(u64), h->stream));
  if (algo == HJ_ALGO_ADAPTIVE && is_device_ptr(R)) {
    // Locality pre-check on a device-resident R: the span signal needs no table, so an input that is
    // going to the radix join anyway never pays for clearing one.  (Host inputs keep the order below:
    // their clear hides under the R copy.)
    if ((rc = stage(h, (const u64*)R, n, h->stageR, &dR, E_CPR, &pendR))) return rc;  // re-staged if misaligned
    R = reinterpret_cast<const uint64_t*>(dR);
    if ((rc = run_span(h, (const u64*)R, n))) return rc;
    if ((rc = fetch_counters(h))) return rc;
    const double locality = contention_of(h);
    const double thr0 = h->cfg.adaptive_threshold > 0 ? h->cfg.adaptive_threshold : default_threshold(h);
    if (locality > thr0) {
      const u64 sampled = h->h_ctr[C_SAMPLED];
      mark(h, E_CLEAR);
      mark(h, E_SAMPLE);
      h->built = false;
      if ((rc = stage(h, (const u64*)S, m, h->stageS, &dS, E_CPS, &pendS))) return rc;
      rc = run_radix_join(h, (const u64*)R, n, false, dS, m, pendS, res);
      if (res && rc == HJ_OK) {
        res->contention = locality;
        res->sampled = sampled;
        res->us_sample = span_us(h, E_T0, E_SAMPLE);
      }
      return rc;
    }
    CK(cudaMemsetAsync(h->ctr, 0, C_COUNT * sizeof(u64), h->stream));
  }
  if ((rc = clear_table(h))) return rc;
  mark(h, E_CLEAR);
  mark(h, E_CP0, h->copy_stream);
  const bool global_only = algo == HJ_ALGO_ATOMIC || algo == HJ_ALGO_NOCC;
  if ((rc = stage(h, (const u64*)R, n, h->stageR, &dR, E_CPR, &pendR, !global_only))) return rc;
  if ((rc = wait_staged(h, E_CPR, pendR))) return rc;
  if (global_only) {
    // launch the build before staging S so the S copy runs under it
    // (run_global_join waits for S right before the probe)
    const bool host_s = S && !is_device_ptr(S);
    if (!host_s) {
      if ((rc = stage(h, (const u64*)S, m, h->stageS, &dS, E_CPS, &pendS, false))) return rc;
      return run_global_join(h, dR, n, dS, m, pendS, algo == HJ_ALGO_NOCC, false, 0, 0, res);
    }
    const u32 ipt = ipt_for(h->cfg.transaction_size);
    if ((rc = launch_build(h, dR, whole(n), algo == HJ_ALGO_NOCC, false, ipt))) return rc;
    if ((rc = run_overflow(h))) return rc;
    mark(h, E_BUILD);
    h->built = true;
    if ((rc = stage(h, (const u64*)S, m, h->stageS, &dS, E_CPS, &pendS))) return rc;
    if ((rc = wait_staged(h, E_CPS, pendS))) return rc;
    mark(h, E_PROBE0);
    if ((rc = launch_probe(h, dS, m, ipt))) return rc;
    mark(h, E_PROBE);
    mark(h, E_END);
    if ((rc = fetch_counters(h))) return rc;
    if (res) {
      fill_common(h, res);
      res->path_taken = HJ_PATH_GLOBAL;
      res->hbm_bytes = 16ull * n + 16ull * m;
    }
    return HJ_OK;
  }

  // HJ_ALGO_ADAPTIVE: HTMHashBuild.hpp:100-156 -- sample, measure, decide.
  u64 stride, slen;
  if ((rc = run_sample(h, dR, n, &stride, &slen))) return rc;
  mark(h, E_SAMPLE);
  if ((rc = stage(h, (const u64*)S, m, h->stageS, &dS, E_CPS, &pendS))) return rc;
  if ((rc = fetch_counters(h))) return rc;
  const double contention = contention_of(h);
  const u64 sampled = h->h_ctr[C_SAMPLED], retries = h->h_ctr[C_CAS_RETRIES], sectors = h->h_ctr[C_SECTORS];
  const float us_sample = span_us(h, E_CLEAR, E_SAMPLE), us_clear = span_us(h, E_T0, E_CLEAR);
  const double thr = h->cfg.adaptive_threshold > 0 ? h->cfg.adaptive_threshold : default_threshold(h);
  if (contention > thr) {
    // radix path; the sampled prefix is abandoned with the table
    h->built = false;
    rc = run_radix_join(h, dR, n, false, dS, m, pendS, res);
    if (res && rc == HJ_OK) {
      res->contention = contention;
      res->sampled = sampled;
      res->cas_retries = retries;
      res->sectors = sectors;
      res->us_sample = us_sample;
      res->us_clear = us_clear;
    }
    return rc;
  }
  // global path: the sampled prefix stays inserted (HTMHashBuild.hpp:170)
  rc = run_global_join(h, dR, n, dS, m, pendS, false, true, stride, slen, res);
  if (res && rc == HJ_OK) {
    res->contention = contention;
    res->us_sample = us_sample;
  }
  return rc;
}

int hj_join_materialize(hj_handle* h, const uint64_t* R, uint64_t n, const uint64_t* S, uint64_t m, uint64_t* out_pairs,
                        uint64_t capacity_pairs, uint64_t* n_pairs, hj_result* res) {
  if (!h || !R || !n || !S || !m) return HJ_EINVAL;
  if (out_pairs && (!is_device_ptr(out_pairs) || ((uintptr_t)out_pairs & 15))) {
    snprintf(h->err, sizeof(h->err), "materialise: out_pairs must be a 16-byte aligned device pointer");
    return HJ_EINVAL;
  }
  CK(cudaSetDevice(h->device));
  if (res) memset(res, 0, sizeof(*res));
  reset_events(h);
  int rc;
  const u64 *dR = nullptr, *dS = nullptr;
  bool pendR = false, pendS = false;
  mark(h, E_T0);
  mark(h, E_CP0, h->copy_stream);
  if ((rc = stage(h, (const u64*)R, n, h->stageR, &dR, E_CPR, &pendR))) return rc;
  if ((rc = stage(h, (const u64*)S, m, h->stageS, &dS, E_CPS, &pendS))) return rc;
  MatReq mat{reinterpret_cast<ulonglong2*>(out_pairs), capacity_pairs, 0, false};
  rc = run_radix_join(h, dR, n, pendR, dS, m, pendS, res, false, false, nullptr, nullptr, &mat);
  if (n_pairs) *n_pairs = mat.needed;
  return rc;
}

int hj_checksums(hj_handle* h, const uint64_t* R, uint64_t n, hj_result* res) {
  if (!h || !res) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  int rc;
  const u64* dR = nullptr;
  if (R && n) {
    bool pend = false;
    if ((rc = stage(h, (const u64*)R, (wide(h) ? 2 : 1) * n, h->stageR, &dR, E_CPR, &pend, wide(h)))) return rc;
    if ((rc = wait_staged(h, E_CPR, pend))) return rc;
  }
  CK(cudaMemsetAsync(h->ctr + C_SUM0, 0, 3 * sizeof(u64), h->stream));
  int grid = h->sm_count * 8;
  const u64 words = wide(h) ? 2 : 1;  // 16-byte tuples: sums of key + payload words
  if (dR) sum_u64_kernel<<<grid, 256, 0, h->stream>>>(dR, words * n, h->ctr + C_SUM0);
  if (h->built) {
    sum_u64_kernel<<<grid, 256, 0, h->stream>>>(h->table.p, words * h->slots, h->ctr + C_SUM1);
    if (wide(h)) sum_ovf16_kernel<<<grid, 256, 0, h->stream>>>(reinterpret_cast<const tup16*>(h->ovf_list.p), h->ctr, h->ctr + C_SUM2);
    else sum_ovf_kernel<<<grid, 256, 0, h->stream>>>(h->ovf_list.p, h->ctr, h->ctr + C_SUM2);
  }
  CK(cudaGetLastError());
  if ((rc = fetch_counters(h))) return rc;
  res->input_sum = h->h_ctr[C_SUM0];
  res->table_sum = h->h_ctr[C_SUM1];
  res->overflow_sum = h->h_ctr[C_SUM2];
  res->conflicts = h->h_ctr[C_CONFLICTS];
  return HJ_OK;
}

int hj_radix_partition(hj_handle* h, const uint64_t* in, uint64_t n, uint32_t shift, uint32_t bits, uint64_t* out,
                       uint64_t* offsets) {
  if (!h || !in || !out || !offsets || bits < 1 || bits > kMaxBitsPerPass) return HJ_EINVAL;
  if (!is_device_ptr(in) || !is_device_ptr(out) || !is_device_ptr(offsets)) return HJ_EINVAL;
  if (need_aligned16(h, in, "hj_radix_partition(in)") || need_aligned16(h, out, "hj_radix_partition(out)")) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  RelBufs& rb = h->rb[0];
  const u32 F = 1u << bits;
  int rc;
  if ((rc = ensure(h, rb.ghist, F < 8 ? 8 : F))) return rc;
  if ((rc = ensure(h, rb.cursors, F))) return rc;
  if ((rc = ensure(h, rb.poff, 4))) return rc;
  if ((rc = ensure(h, rb.pcur, 4))) return rc;
  if ((rc = ensure(h, rb.ptiles, 4))) return rc;
  if ((rc = ensure(h, rb.one, 4))) return rc;
  Digit dg;
  memset(&dg, 0, sizeof(dg));
  dg.key_mask = h->cfg.key_mask; dg.mode = 0; dg.shift = shift; dg.bits = bits;
  const u32 tile = kScThreads * kExactIPT;
  CK(cudaMemsetAsync(rb.ghist.p, 0, F * sizeof(u64), h->stream));
  if ((rc = launch_hist(h, (const u64*)in, n, dg, F, rb.ghist.p))) return rc;
  rx_scan_kernel<<<1, kScanThreads, 0, h->stream>>>(rb.ghist.p, F, (u64*)offsets, rb.cursors.p, F, rb.poff.p,
                                                    rb.pcur.p, rb.ptiles.p, rb.one.p, tile, nullptr, n);
  CK(cudaGetLastError());
  if ((rc = launch_scatter_t<kExactIPT>(h, (const u64*)in, (u64*)out, dg, F, 1, rb.poff.p, rb.ptiles.p, rb.cursors.p,
                                (n + tile - 1) / tile)))
    return rc;
  CK(cudaStreamSynchronize(h->stream));
  return HJ_OK;
}

int hj_exchange_partition(hj_handle* h, const uint64_t* in, uint64_t n, uint32_t n_dest, uint32_t domain_bits,
                          uint64_t* out, uint64_t* offsets) {
  if (!h || !in || !out || !offsets || n_dest < 1 || n_dest > 4096 || domain_bits > 62) return HJ_EINVAL;
  if (!is_device_ptr(in) || !is_device_ptr(out) || !is_device_ptr(offsets)) return HJ_EINVAL;
  if (need_aligned16(h, in, "hj_exchange_partition(in)") || need_aligned16(h, out, "hj_exchange_partition(out)")) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  RelBufs& rb = h->rb[0];
  const u32 F = n_dest;
  int rc;
  if ((rc = ensure(h, rb.ghist, F < 8 ? 8 : F))) return rc;
  if ((rc = ensure(h, rb.cursors, F))) return rc;
  if ((rc = ensure(h, rb.poff, 4))) return rc;
  if ((rc = ensure(h, rb.pcur, 4))) return rc;
  if ((rc = ensure(h, rb.ptiles, 4))) return rc;
  if ((rc = ensure(h, rb.one, 4))) return rc;
  Digit dg;
  memset(&dg, 0, sizeof(dg));
  dg.key_mask = h->cfg.key_mask; dg.mode = 2; dg.ndest = n_dest; dg.domain_bits = domain_bits;
  const u32 tile = kScThreads * kExactIPT;
  CK(cudaMemsetAsync(rb.ghist.p, 0, F * sizeof(u64), h->stream));
  if ((rc = launch_hist(h, (const u64*)in, n, dg, F, rb.ghist.p))) return rc;
  rx_scan_kernel<<<1, kScanThreads, 0, h->stream>>>(rb.ghist.p, F, (u64*)offsets, rb.cursors.p, F, rb.poff.p,
                                                    rb.pcur.p, rb.ptiles.p, rb.one.p, tile, nullptr, n);
  CK(cudaGetLastError());
  if ((rc = launch_scatter_t<kExactIPT>(h, (const u64*)in, (u64*)out, dg, F, 1, rb.poff.p, rb.ptiles.p, rb.cursors.p,
                                (n + tile - 1) / tile)))
    return rc;
  CK(cudaStreamSynchronize(h->stream));
  return HJ_OK;
}

// ---- exchange fused with radix pass 1 (multi-GPU) ------------------------------
// Each rank scatters its slice by (owner GPU of the key range, low b1 bits) straight into the
// owners' pass-1 buffers over NVLink (peer memory mapped through CUDA IPC), so the all-to-all IS
// radix pass 1: no send buffer, no separate collective, no extra pass over the data.

int hj_radix_plan(hj_handle* h, uint64_t n_build, uint32_t* bits, uint32_t* b1, uint32_t* b2, uint32_t* passes) {
  if (!h) return HJ_EINVAL;
  RadixPlan pl;
  int rc = plan_radix(h, n_build, &pl);
  if (rc) return rc;
  if (bits) *bits = pl.bits;
  if (b1) *b1 = pl.b1;
  if (b2) *b2 = pl.b2;
  if (passes) *passes = pl.passes;
  return HJ_OK;
}

int hj_ipc_export(hj_handle* h, const void* dptr, void* handle64) {
  if (!h || !dptr || !handle64) return HJ_EINVAL;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle travels as 64 bytes");
  CK(cudaSetDevice(h->device));
  CK(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(handle64), const_cast<void*>(dptr)));
  return HJ_OK;
}
int hj_ipc_open(hj_handle* h, const void* handle64, void** dptr) {
  if (!h || !handle64 || !dptr) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  cudaIpcMemHandle_t hd;
  memcpy(&hd, handle64, sizeof(hd));
  CK(cudaIpcOpenMemHandle(dptr, hd, cudaIpcMemLazyEnablePeerAccess));
  return HJ_OK;
}
int hj_ipc_close(hj_handle* h, void* dptr) {
  if (!h || !dptr) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  CK(cudaIpcCloseMemHandle(dptr));
  return HJ_OK;
}

static int peer_digit(hj_handle* h, uint32_t n_dest, uint32_t domain_bits, uint32_t b1, Digit* dg) {
  if (n_dest < 1 || n_dest > (uint32_t)kScMaxPeers || domain_bits > 62 || b1 > 8 || (n_dest << b1) > 4096) return HJ_EINVAL;
  memset(dg, 0, sizeof(*dg));
  dg->key_mask = h->cfg.key_mask;
  dg->mode = 3; dg->ndest = n_dest; dg->domain_bits = domain_bits; dg->b1 = b1;
  return HJ_OK;
}

// counts[(dest << b1) | low bits] += tuples of `in`; counts is a DEVICE array of n_dest << b1 u64, zeroed here.
int hj_peer_hist(hj_handle* h, const uint64_t* in, uint64_t n, uint32_t n_dest, uint32_t domain_bits, uint32_t b1,
                 uint64_t* counts) {
  if (!h || !in || !counts || !is_device_ptr(in) || !is_device_ptr(counts)) return HJ_EINVAL;
  if (need_aligned16(h, in, "hj_peer_hist(in)")) return HJ_EINVAL;
  Digit dg;
  int rc = peer_digit(h, n_dest, domain_bits, b1, &dg);
  if (rc) return rc;
  CK(cudaSetDevice(h->device));
  const u32 F = n_dest << b1;
  CK(cudaMemsetAsync(counts, 0, F * sizeof(u64), h->stream));
  if ((rc = launch_hist(h, (const u64*)in, n, dg, F, (u64*)counts))) return rc;
  CK(cudaStreamSynchronize(h->stream));
  return HJ_OK;
}

// peer_bufs: HOST array of n_dest device pointers (entry `self` = the local buffer, the others
// opened with hj_ipc_open); cursors: DEVICE array of n_dest << b1 u64, on entry the first element
// of this rank's exclusive range for every (dest, low bits), advanced by the claims.
struct PeerRegions {  // histogram-free exchange: this rank's regions in every destination buffer
  u32 rank;
  u64 region_cap, trash_off;
};
static int peer_scatter_impl(hj_handle* h, const uint64_t* in, uint64_t n, uint32_t n_dest, uint32_t domain_bits,
                             uint32_t b1, uint64_t* const* peer_bufs, uint64_t* cursors, uint32_t max_ctas, bool sync,
                             const PeerRegions* reg = nullptr);

int hj_peer_scatter(hj_handle* h, const uint64_t* in, uint64_t n, uint32_t n_dest, uint32_t domain_bits, uint32_t b1,
                    uint64_t* const* peer_bufs, uint64_t* cursors) {
  return peer_scatter_impl(h, in, n, n_dest, domain_bits, b1, peer_bufs, cursors, 0, true);
}
// Same, but only enqueued (hj_sync waits for it), on at most max_ctas CTAs (0: all SMs).  The scatter is
// NVLink-bound, so a third of the SMs keeps the links busy while another handle's pass 2 + join of the
// previous chunk runs on the rest (bench.py --chunks).
int hj_peer_scatter_async(hj_handle* h, const uint64_t* in, uint64_t n, uint32_t n_dest, uint32_t domain_bits,
                          uint32_t b1, uint64_t* const* peer_bufs, uint64_t* cursors, uint32_t max_ctas) {
  return peer_scatter_impl(h, in, n, n_dest, domain_bits, b1, peer_bufs, cursors, max_ctas, false);
}

// Histogram-free exchange, sending side: no hj_peer_hist, no counts before the scatter.  Region
// low * n_dest + rank of every destination buffer (region_cap tuples each, the same on all ranks) belongs
// to this rank; cursors[(dest << b1) | low] must hold that region's first element on entry and holds
// its fill level on return -- the receivers need those values (all-gather) for hj_join_regions.
// *overflowed = 1: a region was too small (skewed keys); its surplus went to the scratch zone at
// trash_off of the destination buffer (one 16384-tuple tile + 2) and the exchange has to be redone
// with the exact protocol (hj_peer_hist / hj_peer_scatter / hj_join_pass1).
int hj_peer_scatter_regions(hj_handle* h, const uint64_t* in, uint64_t n, uint32_t n_dest, uint32_t domain_bits,
                            uint32_t b1, uint64_t* const* peer_bufs, uint64_t* cursors, uint32_t rank,
                            uint64_t region_cap, uint64_t trash_off, uint32_t* overflowed) {
  if (!h || rank >= n_dest || !region_cap || (region_cap & 1) || (trash_off & 1)) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  CK(cudaMemsetAsync(h->ctr + C_OPT_OVF, 0, sizeof(u64), h->stream));
  PeerRegions reg{rank, region_cap, trash_off};
  int rc = peer_scatter_impl(h, in, n, n_dest, domain_bits, b1, peer_bufs, cursors, 0, false, &reg);
  if (rc) return rc;
  CK(cudaMemcpyAsync(h->h_ctr + C_OPT_OVF, h->ctr + C_OPT_OVF, sizeof(u64), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  if (overflowed) *overflowed = h->h_ctr[C_OPT_OVF] ? 1u : 0u;
  return HJ_OK;
}

static int peer_scatter_impl(hj_handle* h, const uint64_t* in, uint64_t n, uint32_t n_dest, uint32_t domain_bits,
                             uint32_t b1, uint64_t* const* peer_bufs, uint64_t* cursors, uint32_t max_ctas, bool sync,
                             const PeerRegions* reg) {
  if (!h || !in || !peer_bufs || !cursors || !is_device_ptr(in) || !is_device_ptr(cursors)) return HJ_EINVAL;
  if (need_aligned16(h, in, "hj_peer_scatter(in)")) return HJ_EINVAL;
  Digit dg;
  int rc = peer_digit(h, n_dest, domain_bits, b1, &dg);
  if (rc) return rc;
  CK(cudaSetDevice(h->device));
  RelBufs& rb = h->rb[0];
  if ((rc = ensure(h, rb.one, 4))) return rc;
  if ((rc = ensure(h, h->peer_ptrs, kScMaxPeers))) return rc;
  const u32 F = n_dest << b1;
  // tile = 512 x IPT tuples: 32 -> one CTA per SM, the longest runs; 16 (HJ_PEER_IPT=16) -> two CTAs per SM
  const int peer_ipt = env().peer_ipt;
  const u32 tile = kScThreads * (peer_ipt == 16 ? 16 : 32);
  const u64 ntiles = (n + tile - 1) / tile;
  const u64 one[4] = {0, n, 0, ntiles};  // the one-parent view: offsets {0, n}, tiles {0, ntiles}
  CK(cudaMemcpyAsync(rb.one.p, one, sizeof(one), cudaMemcpyHostToDevice, h->stream));
  CK(cudaMemcpyAsync(h->peer_ptrs.p, peer_bufs, n_dest * sizeof(u64*), cudaMemcpyHostToDevice, h->stream));
  // runs leave through bulk async copies (TMA) unless HJ_PEER_NO_BULK is set (then: 8-byte stores)
  const bool bulk = !env().peer_no_bulk && (F % 2 == 0);
  if (reg && !bulk) return HJ_EINVAL;
  auto k = reg              ? rx_scatter_kernel<32, 2, true, 1, true>
           : peer_ipt == 16 ? rx_scatter_kernel<16, 2, true, 0, true>
           : bulk           ? rx_scatter_kernel<32, 2, true, 0, true>
                            : rx_scatter_kernel<32, 2, true, 0, false>;
  ScOpt so{};
  if (reg) so = ScOpt{rb.one.p + 1, reg->region_cap, reg->trash_off, h->ctr, 0, (1u << b1) - 1u, n_dest, reg->rank};
  size_t smem = (size_t)tile * 8 + (size_t)F * 16 + (bulk ? sc_bulk_extra(F) : 0);
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int grid = grid_for(h, (const void*)k, kScThreads, smem);
  if (env().peer_ctas > 0) grid = h->sm_count * env().peer_ctas;
  if (max_ctas && (u32)grid > max_ctas) grid = (int)max_ctas;
  if ((u64)grid > ntiles) grid = (int)(ntiles ? ntiles : 1);
  k<<<grid, kScThreads, smem, h->stream>>>((const u64*)in, nullptr, dg, F, 1, rb.one.p, rb.one.p + 2, (u64*)cursors,
                                           reinterpret_cast<u64* const*>(h->peer_ptrs.p), so);
  h->launches++;
  CK(cudaGetLastError());
  if (sync) CK(cudaStreamSynchronize(h->stream));
  return HJ_OK;
}

// Join of relations that are already pass-1 partitioned (by the low b1 bits of hj_radix_plan,
// partition-major, contiguous): pass 2 + per-partition join.  Device pointers only.  offR1/offS1:
// the 2^b1 + 1 pass-1 partition offsets (known from the exchange protocol) -- with them pass 2
// runs histogram-free; without them (nullptr) the arrived data is histogrammed first.
int hj_join_pass1(hj_handle* h, const uint64_t* R1, uint64_t n, const uint64_t* offR1, const uint64_t* S1,
                  uint64_t m, const uint64_t* offS1, hj_result* res) {
  if (!h || !R1 || !S1 || !is_device_ptr(R1) || !is_device_ptr(S1)) return HJ_EINVAL;
  if (need_aligned16(h, R1, "hj_join_pass1(R1)") || need_aligned16(h, S1, "hj_join_pass1(S1)")) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  if (res) memset(res, 0, sizeof(*res));
  reset_events(h);
  mark(h, E_T0);
  if ((offR1 && !is_device_ptr(offR1)) || (offS1 && !is_device_ptr(offS1))) return HJ_EINVAL;
  return run_radix_join(h, (const u64*)R1, n, false, (const u64*)S1, m, false, res, true, false, (const u64*)offR1,
                        (const u64*)offS1);
}

// hj_join_pass1 for what a histogram-free exchange (hj_peer_scatter_regions) delivered: pass-1 partition
// `low` sits in n_src regions of cap tuples, region low * n_src + src = [(low * n_src + src) * cap,
// cur[low * n_src + src]); curR / curS: DEVICE arrays of 2^b1 * n_src fill levels (the senders' final
// cursors), n / m: tuples that arrived.  n_src must be a power of two, 2^b1 * n_src <= 256.
int hj_join_regions(hj_handle* h, const uint64_t* R1, uint64_t n, const uint64_t* curR, uint64_t capR,
                    const uint64_t* S1, uint64_t m, const uint64_t* curS, uint64_t capS, uint32_t n_src,
                    hj_result* res) {
  if (!h || !R1 || !S1 || !curR || !curS || !n_src || !is_device_ptr(R1) || !is_device_ptr(S1) ||
      !is_device_ptr(curR) || !is_device_ptr(curS))
    return HJ_EINVAL;
  if (need_aligned16(h, R1, "hj_join_regions(R1)") || need_aligned16(h, S1, "hj_join_regions(S1)")) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  if (res) memset(res, 0, sizeof(*res));
  reset_events(h);
  mark(h, E_T0);
  Pre1Regions rr{(const u64*)curR, capR, n_src}, rs{(const u64*)curS, capS, n_src};
  return run_radix_join(h, (const u64*)R1, n, false, (const u64*)S1, m, false, res, true, false, nullptr, nullptr, nullptr,
                        &rr, &rs);
}

int hj_generate(hj_handle* h, int kind, uint64_t first, uint64_t count, uint64_t domain, uint64_t seed,
                uint64_t* out) {
  if (!h || !out || kind < 0 || kind > 1) return HJ_EINVAL;
  if (kind == 1 && (domain == 0 || (domain & (domain - 1)))) return HJ_EINVAL;
  if (!is_device_ptr(out)) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  u32 k = 0;
  while ((1ull << k) < domain) k++;
  gen_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>(kind, first, count, k, seed, (u64*)out);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(h->stream));
  return HJ_OK;
}

int hj_device_alloc(hj_handle* h, uint64_t bytes, void** out) {
  if (!h || !out) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  CK(cudaMalloc(out, bytes ? page_round_bytes(bytes) : 8));
  return HJ_OK;
}
int hj_device_free(hj_handle* h, void* p) {
  if (!h) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  CK(cudaFree(p));
  return HJ_OK;
}
int hj_host_alloc(uint64_t bytes, void** out) {
  if (!out) return HJ_EINVAL;
  if (cudaMallocHost(out, bytes ? bytes : 8) != cudaSuccess) {
    cudaGetLastError();
    return HJ_ENOMEM;
  }
  return HJ_OK;
}
int hj_host_free(void* p) {
  if (cudaFreeHost(p) != cudaSuccess) {
    cudaGetLastError();
    return HJ_ECUDA;
  }
  return HJ_OK;
}
int hj_copy(hj_handle* h, void* dst, const void* src, uint64_t bytes) {
  if (!h || !dst || !src) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return HJ_OK;
}
int hj_flush_l2(hj_handle* h) {
  if (!h) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  const u64 elems = (256ull << 20) / 8;  // 256 MiB > 126 MB L2
  int rc;
  if ((rc = ensure(h, h->flush, elems))) return rc;
  fill_zero_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>(reinterpret_cast<ulonglong2*>(h->flush.p), elems / 2);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(h->stream));
  return HJ_OK;
}
int hj_sync(hj_handle* h) {
  if (!h) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  CK(cudaStreamSynchronize(h->stream));
  CK(cudaStreamSynchronize(h->copy_stream));
  return HJ_OK;
}

}  // extern "C"

// ===========================================================================
// Pipelined exchange over the copy engines (multi-GPU)
// ===========================================================================
// The fused exchange of round 1 stores over NVLink from the SMs: the whole chip scatters for 13.4 ms at
// N=8 and only then joins for 7.9 ms.  Here the SMs never touch the link:
//   * a stage (stage 0 = this rank's slice of R, stages 1..C = C chunks of its slice of S) is scattered
//     LOCALLY, at HBM speed, by (owner GPU, low b1 key bits) into per-destination blocks -- the same
//     tile-sorting kernel, same fixed-capacity regions, every destination pointer local;
//   * one stream per peer then ships block `dest` of the stage to rank dest's receive window with a plain
//     device-to-device copy (copy engines over NVLink through the CUDA-IPC mapping), followed by the
//     block's fill levels (the scatter's final cursors) and a flag word carrying the step's epoch;
//   * the compute stream scatters stage t+1, then waits (rx_xchg_wait_kernel, bounded) for the flags of
//     stage t from every peer, runs pass 2 on what arrived (one parent per (source, low bits) region) and
//     joins it -- so the copies of stage t+1 run under pass 2 + join of stage t, and nothing but the final
//     counter read-back ever synchronises with the host.
// Contract: all ranks call hj_xchg_start with the same, strictly increasing epoch, and nobody starts epoch
// e+1 before every rank finished epoch e (the all-reduce of the result is that barrier).
typedef CUresult (*StreamWaitValue64)(CUstream, CUdeviceptr, cuuint64_t, unsigned int);
static StreamWaitValue64 stream_wait_value() {
  static StreamWaitValue64 fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue64", &p, cudaEnableDefault, &qr) != cudaSuccess ||
        qr != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      p = nullptr;
    }
    return (StreamWaitValue64)p;
  }();
  return fn;
}

struct XStage {
  u64 n_local, cap, stride, recv_off, fill_off, flag_off, tok_off;
  u64* send;
  bool ev_joined;  // a join was launched right after this stage's pass 2 (ev[5] .. ev[6])
  cudaEvent_t scattered;
  cudaEvent_t ev[8];  // scatter0/1, wait0/1, pass2 0/1, join end, spare
};
constexpr u32 kXMaxStages = 17;
constexpr u64 kXPage = (2ull << 20) / sizeof(u64);  // 2 MiB in tuples
constexpr u64 kXScratch = (u64)kScThreads * 32 + 2;  // diverted claims of one scatter tile (+ alignment)
struct Xchg {
  u32 G = 0, rank = 0, b1 = 0, F1 = 0, nstages = 0, domain_bits = 0;
  bool connected = false, started = false;
  bool pow2_domain = false;  // keys lie in [1, 2^domain_bits] (caller's promise, hj_xchg_create): the shift form applies
  u64 window_elems = 0, epoch = 0;
  u64* window = nullptr;
  u64* peer_window[kScMaxPeers];
  XStage st[kXMaxStages];
  u64 *cur = nullptr, *caps_dev = nullptr, *epoch_word = nullptr, *ones = nullptr;
  u64** dest_ptrs = nullptr;  // device [nstages][kScMaxPeers]: where the local scatter puts block `dest`
  cudaStream_t cs[kScMaxPeers];
  // a block travels as `nsub` slices on nsub streams (one copy engine serves one stream at a time: with a single
  // peer one stream would leave the other engines idle); the fill levels + flag follow on cs[p] after all slices
  u32 nsub = 1;
  cudaStream_t sub[kScMaxPeers][4];
  cudaEvent_t sub_done[kScMaxPeers][4];
  RadixPlan pl;
  u64 C2R = 0, C2S = 0, nR = 0, mS = 0;
  u64 s_out_stride = 0;  // pass-2 output of S chunk c: rb[1].out + c * s_out_stride, fill levels rb[1].cursors + c * 2^bits
  cudaEvent_t ev_begin, ev_end;
  // HJ_XCHG_TRACE=1: a timeline of the step (copies per stage and peer, compute-stream phases) on stderr
  bool one_stream = true, peer_copy = false;
  // Ring kept in lock step (one_stream && G > 2): before this rank's copy stream ships round k+1 of a stage (target
  // rank+k+1) it waits -- cuStreamWaitValue64, no SM involved -- for a token from its successor rank+1, the previous
  // sender to that target, saying that its round k is done.  Without it the ranks drift apart, two of them write into
  // one GPU while another receives nothing, and a block takes 1.0-1.15 ms instead of 0.75 (tools/ce_probe_ipc.cu:
  // 715 GB/s in step, 475-550 GB/s with the ranks 0.4 ms apart, 600-620 GB/s with tokens, the skew included).
  bool tokens = false;
  u64 tok_first = 0, tok_words = 0;  // the token area of the window
  cudaStream_t cs_all = nullptr;
  int peer_dev[kScMaxPeers];
  bool trace = false;
  cudaEvent_t cev[kXMaxStages][kScMaxPeers][2];
};

// CUDA loads kernels lazily, at their first launch, and that load synchronises with running kernels: a step
// whose first pass-2 / join launch came AFTER its wait kernel was enqueued would block the host inside
// hj_xchg_start until the peers' stage arrived (and, with several logical ranks in one process, for ever).
// Touch every kernel the pipeline can launch once, at create time.
static void xchg_preload_kernels() {
  cudaFuncAttributes fa;
  const void* ks[] = {
      (const void*)rx_xchg_init_kernel, (const void*)rx_xchg_wait_kernel, (const void*)rx_xchg_parents_kernel,
      (const void*)rx_opt_init_kernel, (const void*)rx_tasks_kernel,
      (const void*)rx_scatter_kernel<16, 2, true, 1, true>, (const void*)rx_scatter_kernel<16, 3, true, 1, true>,
      (const void*)sc_opt_kernel<8>(1, true), (const void*)sc_opt_kernel<8>(1, false),
      (const void*)sc_opt_kernel<16>(1, true), (const void*)sc_opt_kernel<16>(1, false),
      (const void*)sc_opt_kernel<32>(1, true), (const void*)sc_opt_kernel<32>(1, false),
      (const void*)rx_join32_kernel<512, 8, 2, 4>, (const void*)rx_join32_kernel<512, 4, 3, 2>,
      (const void*)rx_join32_kernel<512, 8, 2, 2>, (const void*)rx_join32_kernel<256, 8, 4, 2>,
      (const void*)rx_join32_kernel<1024, 4, 1, 2>, (const void*)rx_join32_kernel<1024, 6, 1, 2>,
      (const void*)rx_join32_kernel<1024, 8, 1, 2>, (const void*)rx_join32_kernel<384, 4, 4, 2>,
      (const void*)rx_join_kernel<4, 3>, (const void*)rx_join_kernel<6, 3>, (const void*)rx_join_kernel<8, 2>};
  for (const void* k : ks) cudaFuncGetAttributes(&fa, k);
  cudaGetLastError();
}

static u64 xchg_region_cap(u64 n_stage, u32 G, u32 F1) {
  const u64 mean = n_stage / ((u64)G * F1) + 1;
  u64 r = 1;
  while (r * r < mean) r++;
  return (mean + 8 * r + 64 + 1) & ~1ull;
}

void xchg_free(hj_handle* h) {
  Xchg* x = h->xc;
  if (!x) return;
  if (x->tokens && x->window) {  // release token waits still pending (a step that was started and never finished)
    cudaMemset(x->window + x->tok_first, 0xFF, x->tok_words * sizeof(u64));
  }
  if (x->cs_all) { cudaStreamSynchronize(x->cs_all); cudaStreamDestroy(x->cs_all); }
  for (u32 p = 0; p < x->G; p++) {
    if (x->cs[p] && x->cs[p] != x->cs_all) { cudaStreamSynchronize(x->cs[p]); cudaStreamDestroy(x->cs[p]); }
    for (u32 k = 0; k < 4; k++) {
      if (x->sub[p][k]) { cudaStreamSynchronize(x->sub[p][k]); cudaStreamDestroy(x->sub[p][k]); }
      if (x->sub_done[p][k]) cudaEventDestroy(x->sub_done[p][k]);
    }
  }
  for (u32 t = 0; t < x->nstages; t++) {
    if (x->st[t].send) cudaFree(x->st[t].send);
    if (x->st[t].scattered) cudaEventDestroy(x->st[t].scattered);
    for (cudaEvent_t e : x->st[t].ev)
      if (e) cudaEventDestroy(e);
    for (u32 p = 0; p < x->G; p++)
      for (cudaEvent_t e : x->cev[t][p])
        if (e) cudaEventDestroy(e);
  }
  void* bufs[] = {x->window, x->cur, x->caps_dev, x->epoch_word, x->ones, x->dest_ptrs};
  for (void* b : bufs)
    if (b) cudaFree(b);
  if (x->ev_begin) cudaEventDestroy(x->ev_begin);
  if (x->ev_end) cudaEventDestroy(x->ev_end);
  delete x;
  h->xc = nullptr;
}

extern "C" {

int hj_xchg_create(hj_handle* h, uint32_t n_ranks, uint32_t rank, uint64_t n_local, uint64_t m_local,
                   uint32_t s_chunks, uint32_t domain_bits, void** window, uint64_t* window_bytes) {
  if (!h || !window || !window_bytes || n_ranks < 2 || n_ranks > (uint32_t)kScMaxPeers || (n_ranks & (n_ranks - 1)) ||
      rank >= n_ranks || !n_local || !m_local || s_chunks < 1 || s_chunks + 1 > kXMaxStages || domain_bits > 62 ||
      m_local % (2ull * s_chunks))
    return HJ_EINVAL;
  if (h->cfg.algo != HJ_ALGO_RADIX) return HJ_ESTATE;
  CK(cudaSetDevice(h->device));
  xchg_free(h);
  xchg_preload_kernels();
  int rc;
  RadixChoice ch;
  if ((rc = choose_radix(h, n_local, true, true, true, false, false, &ch))) return rc;
  if (!ch.opt || ch.pl.passes != 2 || (n_ranks << ch.pl.b1) > (uint32_t)kScParentCache) {
    snprintf(h->err, sizeof(h->err), "hj_xchg_create: needs the 2-pass shared-memory plan with n_ranks << pass1_bits <= %d "
             "(got bits %u, pass 1 %u)", kScParentCache, ch.pl.bits, ch.pl.b1);
    return HJ_EINVAL;
  }
  Xchg* x = new (std::nothrow) Xchg();
  if (!x) return HJ_ENOMEM;
  memset(x->peer_window, 0, sizeof(x->peer_window));
  memset(x->st, 0, sizeof(x->st));
  memset(x->cs, 0, sizeof(x->cs));
  memset(x->sub, 0, sizeof(x->sub));
  memset(x->sub_done, 0, sizeof(x->sub_done));
  memset(x->cev, 0, sizeof(x->cev));
  x->trace = env_num("HJ_XCHG_TRACE", 0) != 0;
  x->ev_begin = x->ev_end = nullptr;
  h->xc = x;
  x->G = n_ranks; x->rank = rank; x->pl = ch.pl; x->b1 = ch.pl.b1; x->F1 = 1u << ch.pl.b1;
  x->nstages = 1 + s_chunks; x->domain_bits = domain_bits;
  x->pow2_domain = true;  // dest = ((key-1) * G) >> domain_bits == (key-1) >> (domain_bits - log2 G) for G = 2^k
  x->nR = n_local; x->mS = m_local / s_chunks;
  x->C2R = opt_region_cap(x->nR, ch.pl.bits);
  x->C2S = opt_region_cap(x->mS, ch.pl.bits);
  const u32 G = x->G, F1 = x->F1, gf = G * F1;
  // window layout (identical on every rank): the stages' G source blocks, then fill levels, then flags
  u64 off = 0;
  for (u32 t = 0; t < x->nstages; t++) {
    XStage& st = x->st[t];
    st.n_local = t == 0 ? n_local : x->mS;
    st.cap = xchg_region_cap(st.n_local, G, F1);
    // Blocks start on 2 MiB boundaries on both sides and travel as whole 2 MiB pages: the copy engines move a block
    // that is only 256-byte aligned at 475-540 GB/s, a page-aligned one at 720-740 GB/s (tools/ce_probe_ipc.cu with
    // PROBE_BLOCK_BYTES, profiles/r02_ce_probe_ipc_align.jsonl)
    st.stride = ((u64)F1 * st.cap + kXScratch + kXPage - 1) & ~(kXPage - 1);
    st.recv_off = off;
    off += (u64)G * st.stride;
  }
  for (u32 t = 0; t < x->nstages; t++) { x->st[t].fill_off = off; off += (gf + 31) & ~31u; }
  for (u32 t = 0; t < x->nstages; t++) { x->st[t].flag_off = off; off += 32; }
  x->tok_first = off;
  for (u32 t = 0; t < x->nstages; t++) { x->st[t].tok_off = off; off += 32; }
  x->tok_words = off - x->tok_first;
  x->window_elems = off;
  auto fail = [&](int code) { xchg_free(h); return code; };
#define XCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    snprintf(h->err, sizeof(h->err), "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
    return fail(e_ == cudaErrorMemoryAllocation ? HJ_ENOMEM : HJ_ECUDA); } } while (0)
  XCK(cudaMalloc(&x->window, page_round_bytes(off * sizeof(u64))));
  h->allocs++;
  XCK(cudaMemset(x->window + x->st[0].fill_off, 0, (off - x->st[0].fill_off) * sizeof(u64)));  // fills + flags
  for (u32 t = 0; t < x->nstages; t++) {
    XCK(cudaMalloc(&x->st[t].send, page_round_bytes((u64)G * x->st[t].stride * sizeof(u64))));
    h->allocs++;
    XCK(cudaEventCreateWithFlags(&x->st[t].scattered, cudaEventDisableTiming));
    for (cudaEvent_t& e : x->st[t].ev) XCK(cudaEventCreate(&e));
  }
  XCK(cudaEventCreate(&x->ev_begin));
  XCK(cudaEventCreate(&x->ev_end));
  XCK(cudaMalloc(&x->cur, (u64)x->nstages * gf * sizeof(u64)));
  XCK(cudaMalloc(&x->caps_dev, x->nstages * sizeof(u64)));
  XCK(cudaMalloc(&x->epoch_word, 32));
  XCK(cudaMalloc(&x->ones, (u64)x->nstages * 4 * sizeof(u64)));
  XCK(cudaMalloc(&x->dest_ptrs, (u64)x->nstages * kScMaxPeers * sizeof(u64*)));
  {
    u64 caps[kXMaxStages], ones[kXMaxStages * 4];
    const u64 tile = (u64)kScThreads * 16;
    for (u32 t = 0; t < x->nstages; t++) {
      caps[t] = x->st[t].cap;
      const u64 n = x->st[t].n_local;
      ones[4 * t] = 0; ones[4 * t + 1] = n; ones[4 * t + 2] = 0; ones[4 * t + 3] = (n + tile - 1) / tile;
    }
    XCK(cudaMemcpy(x->caps_dev, caps, x->nstages * sizeof(u64), cudaMemcpyHostToDevice));
    XCK(cudaMemcpy(x->ones, ones, x->nstages * 4 * sizeof(u64), cudaMemcpyHostToDevice));
  }
  {
    const long long want = env_num("HJ_XCHG_SUBCOPIES", 0);
    // measured at N=2 (profiles/r02_n2_xchg_subcopies.jsonl): 1 slice 13.85 ms, 2 slices 14.57, 4 slices 14.73 --
    // one copy saturates what the concurrently running scatter / pass-2 kernels leave of HBM; default 1
    x->nsub = want > 0 ? (u32)(want > 4 ? 4 : want) : 1u;
  }
  if (x->trace)
    for (u32 t = 0; t < x->nstages; t++)
      for (u32 p = 0; p < G; p++)
        for (cudaEvent_t& e : x->cev[t][p]) XCK(cudaEventCreate(&e));
  // One copy stream for all peers (default): the copy engines serve ONE of a GPU's streams until it has nothing
  // ready and only then the next (HJ_XCHG_TRACE at N=4, per-peer streams: the first S chunk was complete after
  // 11.8 ms although it was scattered after 1.6 ms, because the stream of peer r+1 shipped ITS blocks of every stage
  // first); a single stream keeps the copies in stage order, ring order inside a stage.  HJ_XCHG_PEER_STREAMS=1: one
  // stream per peer.
  x->one_stream = env_num("HJ_XCHG_PEER_STREAMS", 0) == 0;
  if (x->one_stream) XCK(cudaStreamCreateWithFlags(&x->cs_all, cudaStreamNonBlocking));
  x->peer_copy = env_num("HJ_XCHG_PEERCOPY", 0) != 0;
  x->tokens = x->one_stream && G > 2 && env_num("HJ_XCHG_RING", 1) != 0 && env_num("HJ_XCHG_TOKENS", 1) != 0 &&
              stream_wait_value() != nullptr;
  for (u32 p = 0; p < G; p++) {
    if (p == rank) continue;
    if (x->one_stream) x->cs[p] = x->cs_all;
    else XCK(cudaStreamCreateWithFlags(&x->cs[p], cudaStreamNonBlocking));
    for (u32 k = 1; k < x->nsub; k++) {  // slice 0 travels on cs[p] itself
      XCK(cudaStreamCreateWithFlags(&x->sub[p][k], cudaStreamNonBlocking));
      XCK(cudaEventCreateWithFlags(&x->sub_done[p][k], cudaEventDisableTiming));
    }
  }
#undef XCK
  // pass 2 + join buffers for what arrives (expected: as much as was sent)
  if ((rc = ensure_rel(h, h->rb[0], x->nR, x->pl, true, true, G))) return fail(rc);
  if ((rc = ensure_rel(h, h->rb[1], x->mS, x->pl, true, true, G))) return fail(rc);
  // every S chunk keeps its own pass-2 output and fill levels: ONE join at the end builds each R_p once
  x->s_out_stride = ((1ull << x->pl.bits) * x->C2S + kScThreads * 32 + 2 + 1) & ~1ull;
  if ((rc = ensure(h, h->rb[1].out, (u64)s_chunks * x->s_out_stride))) return fail(rc);
  if ((rc = ensure(h, h->rb[1].cursors, (u64)s_chunks << x->pl.bits))) return fail(rc);
  {
    u64 slice = 4 * (x->mS >> x->pl.bits);
    if (slice < 32768) slice = 32768;
    if ((rc = ensure(h, h->tasks, (1ull << x->pl.bits) + 2 * x->mS / slice + 1))) return fail(rc);
  }
  *window = x->window;
  *window_bytes = off * sizeof(u64);
  return HJ_OK;
}

int hj_xchg_connect(hj_handle* h, void* const* peer_windows) {
  if (!h || !h->xc || !peer_windows) return HJ_EINVAL;
  Xchg* x = h->xc;
  CK(cudaSetDevice(h->device));
  u64* dest[kXMaxStages * kScMaxPeers];
  memset(dest, 0, sizeof(dest));
  for (u32 p = 0; p < x->G; p++) {
    x->peer_window[p] = p == x->rank ? x->window : (u64*)peer_windows[p];
    if (!x->peer_window[p] || ((uintptr_t)x->peer_window[p] & 15)) return HJ_EINVAL;
    cudaPointerAttributes pa;
    x->peer_dev[p] = cudaPointerGetAttributes(&pa, x->peer_window[p]) == cudaSuccess ? pa.device : -1;
    cudaGetLastError();
  }
  for (u32 t = 0; t < x->nstages; t++)
    for (u32 d = 0; d < x->G; d++)  // the own block is written in place, the others into the send buffer
      dest[t * kScMaxPeers + d] = d == x->rank ? x->window + x->st[t].recv_off + (u64)x->rank * x->st[t].stride
                                               : x->st[t].send + (u64)d * x->st[t].stride;
  CK(cudaMemcpy(x->dest_ptrs, dest, (u64)x->nstages * kScMaxPeers * sizeof(u64*), cudaMemcpyHostToDevice));
  x->connected = true;
  return HJ_OK;
}

}  // extern "C"

namespace {

// local scatter of stage t + the copies that ship it
int xchg_scatter_stage(hj_handle* h, Xchg* x, u32 t, const u64* in) {
  XStage& st = x->st[t];
  const u32 G = x->G, F1 = x->F1, gf = G * F1, F = gf;
  Digit dg;
  memset(&dg, 0, sizeof(dg));
  dg.key_mask = h->cfg.key_mask; dg.mode = 3; dg.ndest = G; dg.domain_bits = x->domain_bits; dg.b1 = x->b1;
  constexpr int IPT = 16;  // 8192-tuple tiles, two CTAs per SM: the HBM-speed shape of the local passes
  const u32 tile = kScThreads * IPT;
  // 32-bit keys and a power-of-two domain: the shift form of the owner function (sc_digit<3>)
  u32 lg = 0;
  while ((1u << lg) < G) lg++;
  const bool fast = (h->cfg.key_mask & 0xffffffffull) == 0xffffffffull && x->domain_bits >= lg && x->domain_bits <= 32 &&
                    x->pow2_domain;
  if (fast) dg.shift = x->domain_bits - lg;
  auto k = fast ? rx_scatter_kernel<IPT, 3, true, 1, true> : rx_scatter_kernel<IPT, 2, true, 1, true>;
  const size_t smem = (size_t)tile * 8 + (size_t)F * 16 + sc_bulk_extra(F);
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int grid = grid_for(h, (const void*)k, kScThreads, smem);
  const u64 ntiles = (st.n_local + tile - 1) / tile;
  if ((u64)grid > ntiles) grid = (int)(ntiles ? ntiles : 1);
  const u64* one = x->ones + 4 * t;
  ScOpt so{one + 1, st.cap, (u64)F1 * st.cap, h->ctr, 0, F1 - 1u, 1u, 0u};
  cudaEventRecord(st.ev[0], h->stream);
  k<<<grid, kScThreads, smem, h->stream>>>(in, nullptr, dg, F, 1, one, one + 2, x->cur + (u64)t * gf,
                                           reinterpret_cast<u64* const*>(x->dest_ptrs + (u64)t * kScMaxPeers), so);
  h->launches++;
  CK(cudaGetLastError());
  cudaEventRecord(st.ev[1], h->stream);
  CK(cudaEventRecord(st.scattered, h->stream));
  // own block: written in place by the scatter; its fill levels go next to the others'
  CK(cudaMemcpyAsync(x->window + st.fill_off + (u64)x->rank * F1, x->cur + (u64)t * gf + (u64)x->rank * F1,
                     F1 * sizeof(u64), cudaMemcpyDeviceToDevice, h->stream));
  // Peers in ring order, rank+1 first: the copy engines take the streams' copies roughly in the order they were
  // enqueued, and if every rank started with peer 0 all of them would write into ONE GPU at a time (N=8, peers in
  // index order: 175 GB/s per GPU, 35 ms of a 48 ms step spent waiting; profiles/r02_bench_n8_xchg_v1*.json).
  const bool ring = env_num("HJ_XCHG_RING", 1) != 0;
  for (u32 k = 1; k < G; k++) {
    const u32 p = ring ? (x->rank + k) % G : (k <= x->rank ? k - 1 : k);
    cudaStream_t cs = x->cs[p];
    u64* pw = x->peer_window[p];
    CK(cudaStreamWaitEvent(cs, st.scattered, 0));
    if (x->tokens && k >= 2) {  // the successor's round k-1 went to the same target: let it finish first
      const CUresult cr = stream_wait_value()((CUstream)cs, (CUdeviceptr)(x->window + st.tok_off + (k - 1)), x->epoch,
                                              CU_STREAM_WAIT_VALUE_GEQ);
      if (cr != CUDA_SUCCESS) {
        snprintf(h->err, sizeof(h->err), "hj_xchg: cuStreamWaitValue64 failed (%d)", (int)cr);
        return HJ_ECUDA;
      }
    }
    if (x->trace) cudaEventRecord(x->cev[t][p][0], cs);
    {
      u64* const dst = pw + st.recv_off + (u64)x->rank * st.stride;
      const u64* const src = st.send + (u64)p * st.stride;
      const u64 total = ((u64)F1 * st.cap + kXPage - 1) & ~(kXPage - 1);  // whole pages (<= stride: the tail is scratch)
      const u64 per = ((total + x->nsub - 1) / x->nsub + 31) & ~31ull;  // 256-byte aligned slices
      for (u32 k = 1; k < x->nsub; k++) {
        const u64 b = (u64)k * per, e = b + per < total ? b + per : total;
        if (b >= total) break;
        CK(cudaStreamWaitEvent(x->sub[p][k], st.scattered, 0));
        CK(cudaMemcpyAsync(dst + b, src + b, (e - b) * sizeof(u64), cudaMemcpyDeviceToDevice, x->sub[p][k]));
        CK(cudaEventRecord(x->sub_done[p][k], x->sub[p][k]));
      }
      if (x->peer_copy && x->peer_dev[p] >= 0 && x->peer_dev[p] != h->device)
        CK(cudaMemcpyPeerAsync(dst, x->peer_dev[p], src, h->device, (per < total ? per : total) * sizeof(u64), cs));
      else
        CK(cudaMemcpyAsync(dst, src, (per < total ? per : total) * sizeof(u64), cudaMemcpyDeviceToDevice, cs));
      for (u32 k = 1; k < x->nsub; k++)
        if ((u64)k * per < total) CK(cudaStreamWaitEvent(cs, x->sub_done[p][k], 0));
    }
    if (x->tokens && k + 2 <= G)  // round k done: the predecessor's round k+1 has the target to itself
      CK(cudaMemcpyAsync(x->peer_window[(x->rank + G - 1) % G] + st.tok_off + k, x->epoch_word, sizeof(u64),
                         cudaMemcpyDeviceToDevice, cs));
    CK(cudaMemcpyAsync(pw + st.fill_off + (u64)x->rank * F1, x->cur + (u64)t * gf + (u64)p * F1, F1 * sizeof(u64),
                       cudaMemcpyDeviceToDevice, cs));
    CK(cudaMemcpyAsync(pw + st.flag_off + x->rank, x->epoch_word, sizeof(u64), cudaMemcpyDeviceToDevice, cs));
    if (x->trace) cudaEventRecord(x->cev[t][p][1], cs);
  }
  return HJ_OK;
}

int xchg_wait_stage(hj_handle* h, Xchg* x, u32 t) {
  XStage& st = x->st[t];
  cudaEventRecord(st.ev[2], h->stream);
  rx_xchg_wait_kernel<<<1, 32, 0, h->stream>>>(x->window + st.flag_off, x->G, x->rank, x->epoch, h->ctr,
                                               10ull * 1000 * 1000 * 1000);
  h->launches++;
  CK(cudaGetLastError());
  cudaEventRecord(st.ev[3], h->stream);
  return HJ_OK;
}

// pass 2 of what stage t delivered: one parent per (source, low bits) region of the window
int xchg_pass2_stage(hj_handle* h, Xchg* x, u32 t) {
  XStage& st = x->st[t];
  RelBufs& rb = h->rb[t == 0 ? 0 : 1];
  const RadixPlan& pl = x->pl;
  const u32 nb = 1u << pl.bits, F1 = x->F1, F2 = 1u << pl.b2, nparent = x->G * F1;
  const u64 C2 = t == 0 ? x->C2R : x->C2S;
  u64* const out = t == 0 ? rb.out.p : rb.out.p + (u64)(t - 1) * x->s_out_stride;
  u64* const cursors = t == 0 ? rb.cursors.p : rb.cursors.p + ((u64)(t - 1) << pl.bits);
  const u32 ipt = sc_opt_ipt();
  const u32 tile = kScThreads * ipt;
  cudaEventRecord(st.ev[4], h->stream);
  rx_opt_init_kernel<<<(nb + 255) / 256, 256, 0, h->stream>>>(rb.pcur.p, 0, 0, cursors, nb, C2);
  rx_xchg_parents_kernel<<<1, 1024, 0, h->stream>>>(x->window + st.fill_off, nparent, F1, st.cap, st.stride, tile,
                                                    rb.poff.p, rb.pend.p, rb.ptiles.p);
  h->launches += 2;
  const bool bulk2 = sc_bulk_for(F2, tile);
  ScKernel k = ipt == 32 ? sc_opt_kernel<32>(1, bulk2) : ipt == 16 ? sc_opt_kernel<16>(1, bulk2) : sc_opt_kernel<8>(1, bulk2);
  const size_t smem = (size_t)tile * 8 + (size_t)F2 * 16 + (bulk2 ? sc_bulk_extra(F2) : 0);
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int grid = grid_for(h, (const void*)k, kScThreads, smem);
  Digit d2;
  memset(&d2, 0, sizeof(d2));
  d2.key_mask = h->cfg.key_mask; d2.mode = 0; d2.shift = pl.b1; d2.bits = pl.b2;
  ScOpt o{rb.pend.p, C2, (u64)nb * C2, h->ctr, 0, F1 - 1u, 0, 0};
  k<<<grid, kScThreads, smem, h->stream>>>(x->window + st.recv_off, out, d2, F2, nparent, rb.poff.p, rb.ptiles.p,
                                           cursors, nullptr, o);
  h->launches++;
  CK(cudaGetLastError());
  cudaEventRecord(st.ev[5], h->stream);
  return HJ_OK;
}

}  // namespace

extern "C" {

int hj_xchg_start(hj_handle* h, const uint64_t* R, const uint64_t* S, uint64_t epoch) {
  if (!h || !h->xc || !R || !S || !is_device_ptr(R) || !is_device_ptr(S)) return HJ_EINVAL;
  Xchg* x = h->xc;
  if (!x->connected || x->started || epoch <= x->epoch) return HJ_ESTATE;
  if (need_aligned16(h, R, "hj_xchg_start(R)") || need_aligned16(h, S, "hj_xchg_start(S)")) return HJ_EINVAL;
  CK(cudaSetDevice(h->device));
  reset_events(h);
  h->built = false;
  int rc;
  const u32 gf = x->G * x->F1, C = x->nstages - 1;
  x->epoch = epoch;
  CK(cudaEventRecord(x->ev_begin, h->stream));
  CK(cudaMemsetAsync(h->ctr, 0, C_COUNT * sizeof(u64), h->stream));
  rx_xchg_init_kernel<<<(x->nstages * gf + 255) / 256, 256, 0, h->stream>>>(x->cur, x->nstages, gf, x->F1 - 1u,
                                                                            x->caps_dev, x->epoch_word, epoch);
  h->launches++;
  auto stage_in = [&](u32 t) { return t == 0 ? (const u64*)R : (const u64*)S + (u64)(t - 1) * x->mS; };
  // Every stage is scattered up front -- the copy engines then have work from the first scatter on and never
  // wait for the SMs (scattering only one stage ahead left the link idle between stages: N=2, 4 chunks,
  // 2.6 ms of waiting per step) -- then stage after stage is waited for and put through pass 2.  The chunks
  // are joined in groups: a group's join builds every R_p once and probes it with S_p of all the group's
  // chunks; with more than one group the join of the earlier chunks runs while the last ones still travel.
  const u32 ahead = (u32)env_num("HJ_XCHG_AHEAD", 1000);
  u32 scattered = 0;
  auto scatter_upto = [&](u32 t_last) {
    int r = HJ_OK;
    while (scattered <= t_last && scattered < x->nstages && r == HJ_OK) {
      r = xchg_scatter_stage(h, x, scattered, stage_in(scattered));
      scattered++;
    }
    return r;
  };
  const u64 per = (x->nR >> x->pl.bits) + 1;
  u64 need = per + 64;
  for (u64 r = 1; r * r < per * 64; r++) need = per + 64 + r;
  // measured at N=2 (profiles/r02_n2_xchg_order_groups.jsonl): 1 group 13.6 ms (2.2 ms waiting for the last chunks
  // before the one join), 2 groups 12.8 ms (0.1-0.8 ms waiting; the second build of the R tables costs 0.75 ms)
  // With the page-aligned window (copies at 740 GB/s) N=2 never waits for the link: one join, every R table built
  // once (profiles/r02_n2_pg_*.json: 4 chunks, 1 group 11.5 ms, 2 groups 12.0 ms).  From N=4 on the last chunks
  // arrive late enough that joining the first half meanwhile pays.
  // At N=8 the link is the longer pole (7/8 of the tuples leave): every chunk is joined as it arrives, the extra
  // builds of the R tables run while the next chunk travels and only one chunk's join is left after the last copy
  // (profiles/r02_n8_pg_c4*.json: 4 groups 15.0 ms, 2 groups 16.1 ms; at N=4 the same trade loses, r02_n4_pg_c8_g4).
  u32 ngroups = (u32)env_num("HJ_XCHG_JOIN_GROUPS", x->G <= 2 ? 1 : x->G <= 4 ? 2 : (long long)C);
  if (ngroups < 1) ngroups = 1;
  if (ngroups > C) ngroups = C;
  const u32 gsize = (C + ngroups - 1) / ngroups;
  u32 joined = 0;  // chunks already joined
  if ((rc = scatter_upto(ahead))) return rc;
  for (u32 c = 1; c <= C; c++) {
    if ((rc = scatter_upto(c + ahead))) return rc;
    if (c == 1) {
      if ((rc = xchg_wait_stage(h, x, 0))) return rc;
      if ((rc = xchg_pass2_stage(h, x, 0))) return rc;
    }
    if ((rc = xchg_wait_stage(h, x, c))) return rc;
    if ((rc = xchg_pass2_stage(h, x, c))) return rc;
    if (c - joined == gsize || c == C) {
      if ((rc = launch_partition_join(h, x->pl, need, true, x->nR, true, x->mS, x->C2R, x->C2S, c - joined,
                                      x->s_out_stride, 1ull << x->pl.bits, joined)))
        return rc;
      cudaEventRecord(x->st[c].ev[6], h->stream);
      x->st[c].ev_joined = true;
      joined = c;
    } else {
      x->st[c].ev_joined = false;
    }
  }
  CK(cudaEventRecord(x->ev_end, h->stream));
  x->started = true;
  return HJ_OK;
}

int hj_xchg_finish(hj_handle* h, hj_result* res) {
  if (!h || !h->xc) return HJ_EINVAL;
  Xchg* x = h->xc;
  if (!x->started) return HJ_ESTATE;
  CK(cudaSetDevice(h->device));
  x->started = false;
  int rc;
  if ((rc = fetch_counters(h))) return rc;  // waits for the compute stream
  if (h->h_ctr[C_XCHG_TIMEOUT] && x->tokens) {  // a peer is gone: nothing may keep this rank's copy stream waiting
    CK(cudaMemsetAsync(x->window + x->tok_first, 0xFF, x->tok_words * sizeof(u64), h->stream));
    CK(cudaStreamSynchronize(h->stream));
  }
  for (u32 p = 0; p < x->G; p++)
    if (x->cs[p]) CK(cudaStreamSynchronize(x->cs[p]));  // (waits for the slices' streams too) the send buffers are free again
  if (res) {
    memset(res, 0, sizeof(*res));
    fill_common(h, res);
    auto span = [&](cudaEvent_t a, cudaEvent_t b) { float ms = 0.f; return cudaEventElapsedTime(&ms, a, b) == cudaSuccess ? ms * 1000.f : 0.f; };
    res->us_total = span(x->ev_begin, x->ev_end);
    for (u32 t = 0; t < x->nstages; t++) {
      res->us_scatter1 += span(x->st[t].ev[0], x->st[t].ev[1]);
      res->us_reserved[0] += span(x->st[t].ev[2], x->st[t].ev[3]);  // waiting for the peers' copies
      res->us_scatter2 += span(x->st[t].ev[4], x->st[t].ev[5]);
      if (t && x->st[t].ev_joined) res->us_join += span(x->st[t].ev[5], x->st[t].ev[6]);
    }
    res->path_taken = HJ_PATH_RADIX;
    res->radix_bits = x->pl.bits;
    res->passes = 2;
    res->hbm_bytes = 40ull * (x->nR + x->mS * (x->nstages - 1));  // 8*(2P+1) per tuple, P = 2
  }
  if (x->trace) {
    auto at = [&](cudaEvent_t e) { float ms = 0.f; return cudaEventElapsedTime(&ms, x->ev_begin, e) == cudaSuccess ? ms : -1.f; };
    for (u32 t = 0; t < x->nstages; t++) {
      const XStage& st = x->st[t];
      char line[1024];
      int n = snprintf(line, sizeof(line), "[xchg r%u e%llu] stage %u: scatter %.2f-%.2f wait %.2f-%.2f pass2 %.2f-%.2f join-end %.2f | copies",
                       x->rank, (unsigned long long)x->epoch, t, at(st.ev[0]), at(st.ev[1]), at(st.ev[2]), at(st.ev[3]),
                       at(st.ev[4]), at(st.ev[5]), t && st.ev_joined ? at(st.ev[6]) : -1.f);
      for (u32 k = 1; k < x->G && n < 900; k++) {
        const u32 p = (x->rank + k) % x->G;
        n += snprintf(line + n, sizeof(line) - n, " ->%u(d%d) %.2f-%.2f", p, x->peer_dev[p], at(x->cev[t][p][0]), at(x->cev[t][p][1]));
      }
      snprintf(line + n, sizeof(line) - n, "\n");
      fputs(line, stderr);
    }
    fprintf(stderr, "[xchg r%u] end %.2f\n", x->rank, at(x->ev_end));
  }
  if (h->h_ctr[C_XCHG_TIMEOUT]) {
    snprintf(h->err, sizeof(h->err), "hj_xchg: a peer's stage did not arrive within 10 s (epoch %llu)", x->epoch);
    return HJ_ECUDA;
  }
  if (h->h_ctr[C_OPT_OVF]) {
    snprintf(h->err, sizeof(h->err), "hj_xchg: a fixed-capacity region overflowed (skewed keys): redo the step with the exact "
             "protocol (hj_peer_hist / hj_peer_scatter / hj_join_pass1)");
    return HJ_EOVERFLOW;
  }
  return HJ_OK;
}

}  // extern "C"

// ===========================================================================
// hj_join_noalloc: the global-table join with every byte of scratch in a static device array
// ===========================================================================
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it).  The global-table kernels of
// gt_kernels.cuh -- build, overflow insert, probe -- run here on a table, claim bitmap, overflow table / list /
// map and counters carved out of ONE file-scope __device__ array instead of the handle's cudaMalloc'ed buffers.
// The array costs kNaWords * 8 bytes (2.53 MiB) on every GPU that loads the library, until the process exits.
// The name g_na_scratch is a host-side stand-in in host code, not a device address: the address the kernels get
// comes from cudaGetSymbolAddress, once per device, with that device current.
// Every call shares the array and finds the last call's contents in it: the counters, the table and the claim
// bitmap are cleared at the start of each call (the overflow table and map by gt_ovf_clear_kernel when something
// overflowed).  Calls that overlap on one GPU (two streams, two threads) race on it.
namespace {

constexpr u64 kNaMaxR = HJ_NOALLOC_MAX_R;      // tuples of R
constexpr u64 kNaSlotsMax = 2 * kNaMaxR;       // scaleOutput 2 (the reference's default, main.cpp:80)
constexpr u64 kNaCtrOff = 0;                   // C_COUNT counters, padded to 32 words (256 bytes)
constexpr u64 kNaTableOff = 32;
constexpr u64 kNaOvfTableOff = kNaTableOff + kNaSlotsMax;
constexpr u64 kNaOvfListOff = kNaOvfTableOff + kNaSlotsMax;
constexpr u64 kNaClaimOff = kNaOvfListOff + kNaMaxR;
constexpr u64 kNaOvfMapOff = kNaClaimOff + kNaSlotsMax / 64;
constexpr u64 kNaWords = kNaOvfMapOff + kNaSlotsMax / 64;
static_assert(C_COUNT <= 32, "counter block of the no-alloc scratch");
static_assert(C_CHECKSUM == C_MATCHES + 1, "hj_join_noalloc copies {matches, checksum} as one 16-byte block");

static __device__ __align__(256) unsigned long long g_na_scratch[kNaWords];

constexpr int kNaMaxDevices = 64;
struct NaDevice {
  u64* base = nullptr;  // device address of g_na_scratch on this device
  int sm_count = 0;
};
NaDevice g_na_dev[kNaMaxDevices];
std::mutex g_na_mutex;

int na_fail(char* err, uint32_t cap, int code, const char* fmt, ...) {
  if (err && cap) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err, cap, fmt, ap);
    va_end(ap);
  }
  return code;
}

}  // namespace

extern "C" int hj_join_noalloc(int device, void* stream, const uint64_t* R, uint64_t n, const uint64_t* S, uint64_t m,
                               uint32_t probe_length, uint64_t* out, char* err, uint32_t err_cap) {
  if (err && err_cap) err[0] = 0;
  if (device < 0 || device >= kNaMaxDevices) return na_fail(err, err_cap, HJ_EINVAL, "hj_join_noalloc: device %d out of range", device);
  if (!R || !n || !out || (m && !S)) return na_fail(err, err_cap, HJ_EINVAL, "hj_join_noalloc: null input or output, or |R| = 0");
  if (n > kNaMaxR)
    return na_fail(err, err_cap, HJ_EINVAL,
                   "hj_join_noalloc: |R| = %llu exceeds HJ_NOALLOC_MAX_R = %llu (the static scratch holds a table of %llu "
                   "slots); use hj_join on a handle for larger builds",
                   (unsigned long long)n, (unsigned long long)kNaMaxR, (unsigned long long)kNaSlotsMax);
  if (probe_length == 0) probe_length = 4;  // main.cpp:81
#define NCK(call)                                                                                          \
  do {                                                                                                     \
    cudaError_t e_ = (call);                                                                               \
    if (e_ != cudaSuccess)                                                                                 \
      return na_fail(err, err_cap, HJ_ECUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
  } while (0)
  NCK(cudaSetDevice(device));
  if (!is_device_ptr(R) || (m && !is_device_ptr(S)) || !is_device_ptr(out))
    return na_fail(err, err_cap, HJ_EINVAL, "hj_join_noalloc: R, S and out must be device pointers");
  NaDevice nd;
  {
    std::lock_guard<std::mutex> lock(g_na_mutex);
    NaDevice& d = g_na_dev[device];
    if (!d.base) {  // set-up, once per device, with that device current
      void* p = nullptr;
      NCK(cudaGetSymbolAddress(&p, g_na_scratch));
      int sms = 0;
      NCK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
      d.sm_count = sms > 0 ? sms : 148;
      d.base = static_cast<u64*>(p);
    }
    nd = d;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  u64* const ctr = nd.base + kNaCtrOff;
  u64* const table = nd.base + kNaTableOff;
  u64* const ovf_table = nd.base + kNaOvfTableOff;
  u64* const ovf_list = nd.base + kNaOvfListOff;
  u32* const claim = reinterpret_cast<u32*>(nd.base + kNaClaimOff);
  u32* const ovfmap = reinterpret_cast<u32*>(nd.base + kNaOvfMapOff);
  const u64 slots = pow2_ceil_u64(n * 2 < 128 ? 128 : n * 2);  // prepare_table's geometry at scaleOutput 2
  const u64 mask = slots - 1, key_mask = ~0ull;                 // hj_config_default: all 64 bits are the key
  // what must start at zero: counters, table, claim bitmap
  NCK(cudaMemsetAsync(ctr, 0, 32 * sizeof(u64), st));
  NCK(cudaMemsetAsync(table, 0, slots * sizeof(u64), st));
  NCK(cudaMemsetAsync(claim, 0, slots / 8, st));
  {
    constexpr int IPT = 2;
    const u64 chunks = (n + kGtThreads * IPT - 1) / (kGtThreads * IPT);
    u64 grid = (u64)nd.sm_count * 4;
    if (grid > chunks) grid = chunks;
    gt_build_kernel<IPT, false><<<(int)grid, kGtThreads, 0, st>>>(reinterpret_cast<const u64*>(R), whole(n), table, claim, (u32)mask,
                                                                  (u32)(key_mask & mask), probe_length, ovf_list, ctr);
  }
  gt_ovf_clear_kernel<<<nd.sm_count * 2, 256, 0, st>>>(ovf_table, ovfmap, slots / 32, ctr, slots);
  gt_ovf_insert_kernel<<<nd.sm_count * 2, 256, 0, st>>>(ovf_list, ovf_table, ovfmap, ctr, mask, key_mask, slots);
  if (m) {
    {
      constexpr int HIPT = 8;
      const u64 chunks = (m + kGtThreads * HIPT - 1) / (kGtThreads * HIPT);
      u64 grid = (u64)nd.sm_count * 4;
      if (grid > chunks) grid = chunks;
      gt_probe_home_kernel<HIPT><<<(int)grid, kGtThreads, 0, st>>>(reinterpret_cast<const u64*>(S), m, table, mask, key_mask, ctr);
    }
    constexpr int IPT = 4;
    const u64 chunks = (m + kGtThreads * IPT - 1) / (kGtThreads * IPT);
    u64 grid = (u64)nd.sm_count * 3;
    if (grid > chunks) grid = chunks;
    gt_probe_kernel<IPT><<<(int)grid, kGtThreads, 0, st>>>(reinterpret_cast<const u64*>(S), m, table, mask, key_mask, probe_length,
                                                           ovf_table, slots, ovfmap, ctr);
  }
  NCK(cudaGetLastError());
  NCK(cudaMemcpyAsync(out, ctr + C_MATCHES, 2 * sizeof(u64), cudaMemcpyDeviceToDevice, st));
#undef NCK
  return HJ_OK;
}
