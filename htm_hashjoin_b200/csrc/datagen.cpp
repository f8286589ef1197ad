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

// ---- mc's remaining generators: --non-unique, --full-range, --skew, --local-shuffle-range ------
// Same libc calls in the same order as mc/src/generator.c / genzipf.c, so the arrays are the reference's arrays
// (checked against the reference's own code in tests/test_oracle_cpu.py via a binary built from the reference sources).

#include <math.h>

// generator.c:231-238 random_gen via create_relation_nonunique (:493): key = RAND_RANGE(maxid), i.e. keys in
// [0, maxid) -- 0 included (main.c:373-377: maxid = r_size for --non-unique, INT_MAX for --full-range).  The
// join paths treat key 0 as "no tuple" (hj_b200.h), so callers that want every tuple joined pass bias = 1.
extern "C" void hj_mc_relation_nonunique(uint64_t* out, uint64_t n, uint64_t maxid, unsigned seed, uint32_t bias) {
  srand(seed);
  for (uint64_t i = 0; i < n; ++i) {
    const int32_t key = (int32_t)((double)rand() / ((double)RAND_MAX + 1.0) * (double)maxid);
    out[i] = (uint64_t)(uint32_t)(key + (int32_t)bias);
  }
}

// generator.c:448-489 create_relation_fk_from_pk (--full-range S, main.c:395-397): copies of the PK relation back
// to back, then one Knuth shuffle of the whole FK relation under srand(seed)
extern "C" void hj_mc_relation_fk_from_pk(uint64_t* out, uint64_t m, const uint64_t* pk, uint64_t n, unsigned seed) {
  srand(seed);
  for (uint64_t i = 0; i < m; ++i) out[i] = pk[i % n];
  for (uint64_t i = m; i-- > 1;) {
    const uint64_t j = (uint64_t)((double)rand() / ((double)RAND_MAX + 1.0) * (double)i);
    std::swap(out[i], out[j]);
  }
}

// generator.c:95-110,263-282 create_relation_pk_lshuffle (--local-shuffle-range W): 1..n, every position swapped
// with one at most W-1 ahead; unlike DataGen's local_shuffle EVERY position draws a number
extern "C" void hj_mc_relation_pk_lshuffle(uint64_t* out, uint64_t n, int range, unsigned seed) {
  srand(seed);
  iota1(out, n);
  for (uint64_t i = 0; i < n; ++i) {
    const uint64_t runway = n - i;
    const int mod = runway > (uint64_t)range ? range : (int)runway;
    const uint64_t j = i + (uint64_t)(rand() % mod);
    std::swap(out[i], out[j]);
  }
}

// genzipf.c:27-153 gen_zipf via create_relation_zipf (generator.c:522-538; --skew z, main.c:404-407): a random
// permutation of 1..maxid as the alphabet (key >= 1, :39), the cumulated zipf density as a lookup table, one
// rand() and one binary search per tuple
extern "C" void hj_mc_relation_zipf(uint64_t* out, uint64_t m, uint64_t maxid, double z, unsigned seed) {
  srand(seed);
  const unsigned int size = (unsigned int)maxid;
  std::vector<uint32_t> alphabet(size);
  for (unsigned int i = 0; i < size; i++) alphabet[i] = i + 1;
  for (unsigned int i = size - 1; i > 0; i--) {
    const unsigned int k = (unsigned int)((unsigned long)i * (unsigned long)rand() / RAND_MAX);
    std::swap(alphabet[i], alphabet[k]);
  }
  std::vector<double> lut(size);
  double scaling = 0.0;
  for (unsigned int i = 1; i <= size; i++) scaling += 1.0 / pow((double)i, z);
  double sum = 0.0;
  for (unsigned int i = 1; i <= size; i++) {
    sum += 1.0 / pow((double)i, z);
    lut[i - 1] = sum / scaling;
  }
  for (uint64_t t = 0; t < m; t++) {
    const double r = ((double)rand()) / RAND_MAX;
    unsigned int left = 0, right = size - 1, pos;
    if (lut[0] >= r) pos = 0;
    else {
      while (right - left > 1) {
        const unsigned int mid = (left + right) / 2;
        if (lut[mid] < r) left = mid;
        else right = mid;
      }
      pos = right;
    }
    out[t] = alphabet[pos];
  }
}
