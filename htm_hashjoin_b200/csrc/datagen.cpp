// datagen.cpp -- see datagen.hpp.  Written against the behaviour of
// include/DataGen.hpp:26-122 and mc/src/generator.c, not from their text.
#include "datagen.hpp"

#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

namespace {

// DataGen.hpp:46-53 / :63-70 / :107-115: one left-to-right sweep; a position that
// already received an element is skipped without drawing a random number; the
// partner lies rand() % min(range, remaining) ahead; the last element never draws.
void windowed_shuffle(uint64_t* a, uint64_t n, int range) {
  if (n < 2) return;
  std::vector<bool> taken(n, false);
  for (uint64_t i = 0; i + 1 < n; ++i) {
    if (taken[i]) continue;
    const uint64_t remaining = n - i;
    const int window = remaining < (uint64_t)range ? (int)remaining : range;
    const uint64_t j = i + (uint64_t)(rand() % window);
    std::swap(a[i], a[j]);
    taken[j] = true;
  }
}

void iota1(uint64_t* a, uint64_t n) {
  for (uint64_t i = 0; i < n; ++i) a[i] = i + 1;
}

// mc/src/generator.c:84-93 knuth_shuffle with RAND_RANGE (:20) on 1..n
void mc_permutation(uint64_t* a, uint64_t n) {
  iota1(a, n);
  for (uint64_t i = n; i-- > 1;) {
    const uint64_t j = (uint64_t)((double)rand() / ((double)RAND_MAX + 1.0) * (double)i);
    std::swap(a[i], a[j]);
  }
}

}  // namespace

extern "C" int hj_generate_data(const char* dist, uint64_t n, uint64_t distinct_keys, int range, uint64_t* out) {
  srand(0);                                              // DataGen.hpp:27
  const uint32_t mod_mask = (uint32_t)distinct_keys - 1; // :28
  if (!strcmp(dist, "uniform")) {                        // :30-54
    for (uint64_t i = 0; i < n; ++i) out[i] = (uint64_t)(((uint32_t)rand() & mod_mask) + 1);
    std::sort(out, out + n);
    windowed_shuffle(out, n, range);
  } else if (!strcmp(dist, "random")) {                  // :55-71
    for (uint64_t i = 0; i < n; ++i) {
      uint64_t v;
      do v = (uint64_t)rand(); while (v == 0);
      out[i] = v;
    }
    std::sort(out, out + n);
    windowed_shuffle(out, n, range);
  } else if (!strcmp(dist, "sorted")) {                  // :78-85
    iota1(out, n);
  } else if (!strcmp(dist, "shuffle")) {                 // :86-95, std::random_shuffle = rand() % (i+1)
    iota1(out, n);
    for (uint64_t i = 1; i < n; ++i) {
      const uint64_t j = (uint64_t)rand() % (i + 1);
      if (j != i) std::swap(out[i], out[j]);
    }
  } else if (!strcmp(dist, "local_shuffle")) {           // :96-115
    iota1(out, n);
    windowed_shuffle(out, n, range);
  } else {
    return -1;
  }
  return 0;
}

uint64_t* generate_data(std::string dist, uint64_t size, uint64_t distinct_keys, int local_shuffle_range) {
  uint64_t* a = new uint64_t[size ? size : 1];
  if (hj_generate_data(dist.c_str(), size, distinct_keys, local_shuffle_range, a) != 0) {
    delete[] a;
    return nullptr;
  }
  return a;
}

extern "C" void hj_mc_relation_pk(uint64_t* out, uint64_t n, unsigned seed) {
  srand(seed);
  mc_permutation(out, n);
}

extern "C" void hj_mc_relation_fk(uint64_t* out, uint64_t m, uint64_t maxid, unsigned seed) {
  srand(seed);
  const uint64_t full = m / maxid;
  for (uint64_t k = 0; k < full; ++k) mc_permutation(out + k * maxid, maxid);
  const uint64_t rest = m - full * maxid;
  if (rest) mc_permutation(out + full * maxid, rest);
}
