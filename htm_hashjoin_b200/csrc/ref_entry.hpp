// ref_entry.hpp -- the reference's own entry points, re-pointed at the GPU path.
//
// Same names and parameter lists as the functions main.cpp dispatches to
// (AtomicHashBuild.hpp:14-20, NoCCHashBuild.hpp:13-19, HTMHashBuild.hpp:54-60)
// and as mc's PRO (mc/src/parallel_radix_join.h:33-34).  The reference picks
// the build-only or build+probe form with the ENABLE_PROBE macro (config.h:4);
// here both forms exist as overloads.  Each call prints the reference's JSON
// line (same keys, same order) followed by extra keys, and returns nothing,
// like the reference.  On a library error the line carries "error" and the
// process exit code is left to the caller (hj_ref_last_status()).
#pragma once
#include <stdint.h>
// the standard headers the three operator headers pulled in for main.cpp (HTMHashBuild.hpp:20-31), so that
// swapping the includes at main.cpp:10-14 for this file leaves the rest of main.cpp compiling unchanged
#include <sys/time.h>

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <memory>

void AtomicHashBuild(uint64_t* relR, uint32_t rSize, uint32_t scaleOutput, uint32_t numPartitions,
                     uint32_t probeLength);
void AtomicHashBuild(uint64_t* relR, uint32_t rSize, uint64_t* relS, uint32_t sSize, uint32_t scaleOutput,
                     uint32_t numPartitions, uint32_t probeLength);
void NoCCHashBuild(uint64_t* relR, uint32_t rSize, uint32_t scaleOutput, uint32_t numPartitions,
                   uint32_t probeLength);
void NoCCHashBuild(uint64_t* relR, uint32_t rSize, uint64_t* relS, uint32_t sSize, uint32_t scaleOutput,
                   uint32_t numPartitions, uint32_t probeLength);
// HTM has no GPU meaning; its entry point runs the adaptive path: the sampling
// round of HTMHashBuild.hpp:100-156 decides between the global table and the
// radix join, transactionSize = inserts in flight per thread.
void HTMHashBuild(uint64_t* relR, uint32_t rSize, uint32_t transactionSize, uint32_t scaleOutput,
                  uint32_t numPartitions, uint32_t probeLength);
void HTMHashBuild(uint64_t* relR, uint32_t rSize, uint64_t* relS, uint32_t sSize, uint32_t transactionSize,
                  uint32_t scaleOutput, uint32_t numPartitions, uint32_t probeLength);
// SortMerge.hpp:5-6, the sort-merge comparator: both sides are radix-partitioned, each partition pair is sorted
// in shared memory and merged (rx_sortmerge.cuh).  Unlike SortMerge.cpp:23-33 the probe side need not arrive
// sorted.  numPartitions as for radix (64 = automatic).
void SortMerge(uint64_t* build, uint32_t buildSize, uint64_t* probe, uint32_t probeSize, uint32_t numPartitions);

// mc/src/types.h:31-46
struct tuple_t { int32_t key; int32_t payload; };
struct relation_t { tuple_t* tuples; uint32_t num_tuples; };
// mc/src/parallel_radix_join.h:33-34.  relS == nullptr or num_tuples <= 2 (the dummy S of
// experiments/motivation.sh:11, --s-size=2): partition + build only, returns the reference's
// "Results" checksum (parallel_radix_join.c:256); otherwise the join cardinality.  nthreads is
// accepted and ignored.  Fan-out / passes: hj_ref_set_radix().
int64_t PRO(relation_t* relR, relation_t* relS, int nthreads);

extern "C" {
// NUM_RADIX_BITS / NUM_PASSES (mc/src/prj_params.h:15-22) are macros in the
// reference; here they are process-wide settings.  0 = automatic.
void hj_ref_set_radix(uint32_t radix_bits, uint32_t passes);
void hj_ref_set_device(int device);
int hj_ref_last_status(void);
// The generic form behind all of the above: 64-bit sizes, any algo
// ("atomic","nocc","htm"/"adaptive","radix","sortmerge"), JSON line into json_out.
// Returns the hj_* status.
int64_t hj_ref_pro(const void* r_tuples, uint32_t nr, const void* s_tuples, uint32_t ns, int nthreads);
int hj_ref_run(const char* algo, const uint64_t* relR, uint64_t rSize, const uint64_t* relS, uint64_t sSize,
               uint32_t transactionSize, uint32_t scaleOutput, uint32_t numPartitions, uint32_t probeLength,
               uint32_t passes, char* json_out, uint32_t cap);
}
