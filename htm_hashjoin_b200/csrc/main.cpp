// main.cpp -- the reference's `./main` command line on the GPU path.
//
// Flags, defaults and behaviour follow main.cpp:21-128 of the reference:
//   --algo --rSize --transactionSize --probeLength --dataDistr --shuffleRange
//   --scaleOutput --numPartitions, one value each, any order; an unknown flag
// prints a message and exits 1 (:63-66); an unknown algo prints "Unknown Algo"
// and exits 0 (:107-108).  Differences, all additive: sizes accept K/M/G
// suffixes ("--rSize 16M" is atoi()==16 and a SIGFPE in the reference);
// --probeLength is actually stored (:53-54 stores it in dataDistr); the probe is
// a runtime flag --probe 0|1 (default 1) instead of the ENABLE_PROBE macro;
// --algo radix|adaptive|sortmerge, --sSize, --passes, --device, --sDistr fk.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>

#include "../../include/hj_b200.h"
#include "datagen.hpp"
#include "ref_entry.hpp"

struct param_t {
  std::string algo, dataDistr, sDistr;
  uint64_t rSize, sSize;
  uint32_t transactionSize, probeLength, scaleOutput, numPartitions, shuffleRange, passes;
  int probe, device;
};

static uint64_t parse_size(const char* s) {
  char* end = nullptr;
  double v = strtod(s, &end);
  if (end && *end) {
    switch (*end) {
      case 'k': case 'K': v *= 1024.0; break;
      case 'm': case 'M': v *= 1024.0 * 1024.0; break;
      case 'g': case 'G': v *= 1024.0 * 1024.0 * 1024.0; break;
      default: break;
    }
  }
  return (uint64_t)v;
}

static void parse_args(int argc, char** argv, param_t* p) {
  for (int i = 1; i < argc; i++) {
    const char* a = argv[i];
    const char* v = i + 1 < argc ? argv[i + 1] : nullptr;
    auto need = [&]() {
      if (!v) {
        printf("Missing value for %s\n", a);
        exit(1);
      }
      i++;
      return v;
    };
    if (!strcmp(a, "--algo")) p->algo = need();
    else if (!strcmp(a, "--rSize")) p->rSize = parse_size(need());
    else if (!strcmp(a, "--sSize")) p->sSize = parse_size(need());
    else if (!strcmp(a, "--transactionSize")) p->transactionSize = (uint32_t)parse_size(need());
    else if (!strcmp(a, "--probeLength")) p->probeLength = (uint32_t)parse_size(need());
    else if (!strcmp(a, "--dataDistr")) p->dataDistr = need();
    else if (!strcmp(a, "--sDistr")) p->sDistr = need();
    else if (!strcmp(a, "--shuffleRange")) p->shuffleRange = (uint32_t)parse_size(need());
    else if (!strcmp(a, "--scaleOutput")) p->scaleOutput = (uint32_t)parse_size(need());
    else if (!strcmp(a, "--numPartitions")) p->numPartitions = (uint32_t)parse_size(need());
    else if (!strcmp(a, "--passes")) p->passes = (uint32_t)parse_size(need());
    else if (!strcmp(a, "--probe")) p->probe = atoi(need());
    else if (!strcmp(a, "--device")) p->device = atoi(need());
    else {
      printf("Unknown option: %s\n", a);  // main.cpp:63-66
      exit(1);
    }
  }
}

int main(int argc, char* argv[]) {
  param_t p;
  p.algo = "htm";  // main.cpp:78-85
  p.rSize = 1ull << 28;
  p.sSize = 0;
  p.probeLength = 4;
  p.dataDistr = "shuffle";
  p.sDistr = "sorted";
  p.transactionSize = 16;
  p.scaleOutput = 2;
  p.numPartitions = 64;
  p.shuffleRange = 16;
  p.passes = 0;
  p.probe = 1;
  p.device = 0;
  parse_args(argc, argv, &p);

  if (p.algo != "atomic" && p.algo != "nocc" && p.algo != "htm" && p.algo != "adaptive" && p.algo != "radix" &&
      p.algo != "sortmerge") {
    printf("Unknown Algo: %s\n", p.algo.c_str());  // main.cpp:107-108 (exit code 0)
    return 0;
  }
  if (p.rSize == 0 || p.rSize >= (1ull << 32)) {
    printf("rSize out of range for the reference-shaped entry points (uint32_t): use the C ABI\n");
    return 1;
  }
  uint64_t* relR = generate_data(p.dataDistr, p.rSize, p.rSize, (int)p.shuffleRange);
  if (!relR) {
    printf("Unknown distribution\n");  // DataGen.hpp:117-120
    return 1;
  }
  uint64_t* relS = nullptr;
  uint64_t sSize = 0;
  if (p.probe) {
    sSize = p.sSize ? p.sSize : p.rSize;
    relS = new uint64_t[sSize];
    if (p.sDistr == "fk") {
      hj_mc_relation_fk(relS, sSize, p.rSize, 54321);  // mc/src/main.c:346 default s_seed
    } else if (p.dataDistr == "random") {               // main.cpp:94-97: S := copy of R
      for (uint64_t i = 0; i < sSize; i++) relS[i] = relR[i % p.rSize];
    } else {                                            // main.cpp:92-93: S = sorted 1..|R|
      for (uint64_t i = 0; i < sSize; i++) relS[i] = (i % p.rSize) + 1;
    }
  }
  hj_ref_set_device(p.device);
  hj_ref_set_radix(0, p.passes);
  const uint32_t n = (uint32_t)p.rSize, m = (uint32_t)sSize;
  if (p.algo == "atomic") {
    if (relS) AtomicHashBuild(relR, n, relS, m, p.scaleOutput, p.numPartitions, p.probeLength);
    else AtomicHashBuild(relR, n, p.scaleOutput, p.numPartitions, p.probeLength);
  } else if (p.algo == "nocc") {
    if (relS) NoCCHashBuild(relR, n, relS, m, p.scaleOutput, p.numPartitions, p.probeLength);
    else NoCCHashBuild(relR, n, p.scaleOutput, p.numPartitions, p.probeLength);
  } else if (p.algo == "htm" || p.algo == "adaptive") {
    if (relS) HTMHashBuild(relR, n, relS, m, p.transactionSize, p.scaleOutput, p.numPartitions, p.probeLength);
    else HTMHashBuild(relR, n, p.transactionSize, p.scaleOutput, p.numPartitions, p.probeLength);
  } else if (p.algo == "sortmerge") {  // the reference builds this comparator as its own binary (SortMerge.cpp)
    if (!relS) {
      printf("sortmerge needs a probe side (--probe 1)\n");
      delete[] relR;
      return 1;
    }
    SortMerge(relR, n, relS, m, p.numPartitions);
  } else {  // radix
    char buf[4096];
    int rc = hj_ref_run("radix", relR, p.rSize, relS, sSize, p.transactionSize, p.scaleOutput, p.numPartitions,
                        p.probeLength, p.passes, buf, sizeof(buf));
    puts(buf);
    if (rc) {
      delete[] relR;
      delete[] relS;
      return 2;
    }
  }
  delete[] relR;
  delete[] relS;
  return hj_ref_last_status() ? 2 : 0;
}
