// datagen.hpp -- host-side input synthesis, bit-identical to the reference's.
//
// generate_data mirrors include/DataGen.hpp:26-122 (same name, same argument
// meaning, 64-bit sizes): it makes the same libc calls in the same order
// (srand(0), rand()), so the arrays are the reference's arrays.  The mc
// generators mirror mc/src/generator.c (create_relation_pk :241-260,
// create_relation_fk :409-445) and write tuples as key | payload<<32 with
// payload 0.  Arrays are allocated with hj_host_alloc (pinned) when pinned=true.
#pragma once
#include <stdint.h>

#include <string>

// Returns a new[]-allocated array (delete[] it), or nullptr for an unknown
// distribution (the reference prints "Unknown distribution" and exits,
// DataGen.hpp:117-120; the CLI does that on nullptr).
uint64_t* generate_data(std::string dist, uint64_t size, uint64_t distinct_keys, int local_shuffle_range = 16);

extern "C" {
// Same, into caller memory.  0 on success, -1 unknown distribution.
int hj_generate_data(const char* dist, uint64_t size, uint64_t distinct_keys, int local_shuffle_range, uint64_t* out);
// mc/src/generator.c: PK relation 1..n Knuth-shuffled with srand(seed) (main.c:362)
void hj_mc_relation_pk(uint64_t* out, uint64_t n, unsigned seed);
// mc/src/generator.c:409-445: FK relation of m tuples over keys 1..maxid
void hj_mc_relation_fk(uint64_t* out, uint64_t m, uint64_t maxid, unsigned seed);
// mc's other workloads (mc/src/main.c:498-513 --non-unique / --full-range / --skew / --local-shuffle-range):
// create_relation_nonunique (generator.c:493; keys in [0, maxid), + bias), create_relation_fk_from_pk (:448),
// create_relation_pk_lshuffle (:263), create_relation_zipf (:522 -> genzipf.c gen_zipf)
void hj_mc_relation_nonunique(uint64_t* out, uint64_t n, uint64_t maxid, unsigned seed, uint32_t bias);
void hj_mc_relation_fk_from_pk(uint64_t* out, uint64_t m, const uint64_t* pk, uint64_t n, unsigned seed);
void hj_mc_relation_pk_lshuffle(uint64_t* out, uint64_t n, int range, unsigned seed);
void hj_mc_relation_zipf(uint64_t* out, uint64_t m, uint64_t maxid, double z, unsigned seed);
}
