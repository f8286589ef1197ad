// tbb/tbb.h stand-in used ONLY to compile the unmodified reference headers into
// oracle/_ref (Intel TBB is not installed in this image and there is no
// network).  It provides the three TBB entry points the reference's build/probe
// path uses (AtomicHashBuild.hpp:37,71,90; DataGen.hpp:79) on top of
// std::thread.  TEST/BASELINE INFRASTRUCTURE -- never linked into the product.
//
// Scheduling model: the iteration space is cut into chunks of exactly
// `grainsize` elements (what TBB's simple_partitioner would produce for the
// power-of-two splits the reference uses) and the chunks are handed out to
// worker threads through one atomic counter.  HJ_REF_THREADS (env) sets the
// worker count (default: hardware_concurrency); with 1 worker the chunks run
// in index order, i.e. the serial schedule the oracle pins.
#pragma once
#include <algorithm>
#include <atomic>
#include <cstddef>
#include <cstdlib>
#include <thread>
#include <vector>

namespace tbb {

template <typename T>
class blocked_range {
 public:
  blocked_range(T b, T e, std::size_t grain = 1) : b_(b), e_(e), g_(grain ? grain : 1) {}
  T begin() const { return b_; }
  T end() const { return e_; }
  std::size_t grainsize() const { return g_; }
  std::size_t size() const { return std::size_t(e_ - b_); }
  bool empty() const { return !(b_ < e_); }

 private:
  T b_, e_;
  std::size_t g_;
};

namespace shim {
inline unsigned workers() {
  if (const char* e = std::getenv("HJ_REF_THREADS")) {
    int v = std::atoi(e);
    if (v > 0) return unsigned(v);
  }
  unsigned h = std::thread::hardware_concurrency();
  return h ? h : 1;
}

template <typename F>
void run_chunks(std::size_t nchunks, const F& f) {
  unsigned nw = unsigned(std::min<std::size_t>(workers(), nchunks));
  if (nw <= 1) {
    for (std::size_t c = 0; c < nchunks; c++) f(c);
    return;
  }
  std::atomic<std::size_t> next{0};
  std::vector<std::thread> pool;
  pool.reserve(nw);
  for (unsigned w = 0; w < nw; w++)
    pool.emplace_back([&] {
      for (;;) {
        std::size_t c = next.fetch_add(1, std::memory_order_relaxed);
        if (c >= nchunks) break;
        f(c);
      }
    });
  for (auto& t : pool) t.join();
}
}  // namespace shim

template <typename T, typename Body>
void parallel_for(const blocked_range<T>& r, const Body& body) {
  if (r.empty()) return;
  const std::size_t g = r.grainsize();
  const std::size_t n = (r.size() + g - 1) / g;
  shim::run_chunks(n, [&](std::size_t c) {
    T b = r.begin() + T(c * g);
    T e = std::min<T>(r.end(), b + T(g));
    body(blocked_range<T>(b, e, g));
  });
}

template <typename T, typename V, typename Body, typename Join>
V parallel_deterministic_reduce(const blocked_range<T>& r, const V& identity, const Body& body,
                                const Join& join) {
  if (r.empty()) return identity;
  const std::size_t g = r.grainsize();
  const std::size_t n = (r.size() + g - 1) / g;
  // fixed, thread-count independent decomposition: ~4096 super-chunks folded in order
  const std::size_t super = std::max<std::size_t>(1, (n + 4095) / 4096);
  const std::size_t ns = (n + super - 1) / super;
  std::vector<V> part(ns, identity);
  shim::run_chunks(ns, [&](std::size_t s) {
    V acc = identity;
    for (std::size_t c = s * super; c < std::min(n, (s + 1) * super); c++) {
      T b = r.begin() + T(c * g);
      T e = std::min<T>(r.end(), b + T(g));
      acc = body(blocked_range<T>(b, e, g), acc);
    }
    part[s] = acc;
  });
  V out = identity;
  for (std::size_t s = 0; s < ns; s++) out = join(out, part[s]);
  return out;
}

}  // namespace tbb
