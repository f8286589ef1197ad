/* ref_bench.c -- TEST / BASELINE INFRASTRUCTURE, not the product.
 *
 * A long-lived driver around the reference's own mc entry points, compiled UNMODIFIED from
 * /root/reference/mc/src (everything but main.c) by oracle/Makefile into oracle/_ref/mc_bench_b<B>_p<P>:
 * the relations are generated ONCE with the reference's generators and seeds
 * (mc/src/main.c:362-411: seed_generator(12345) + create_relation_pk, seed_generator(54321) +
 * create_relation_fk) and the join entry point is then called warmup + reps times, as
 * experiments/runner.sh:3 repeats its runs -- instead of paying ~40 s of serial rand() per repetition
 * through the mc CLI.  Nothing of the reference is copied here: this file only calls
 *   NPO  (mc/src/no_partitioning_join.h:32, no_partitioning_join.c:537)
 *   PRO  (mc/src/parallel_radix_join.h:33, parallel_radix_join.c:1305; partition + build of R only)
 * and prints, per call, what the callee printed itself plus one line
 *   RUN <i> ALGO <name> RESULT <n> WALL_US <gettimeofday around the call>
 *
 *   mc_bench_b14_p2 --algo=NPO --nthreads=16 --r-size=134217728 --s-size=1073741824 --reps=5 --warmup=1
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>

#include "types.h"
#include "generator.h"
#include "no_partitioning_join.h"
#include "parallel_radix_join.h"

static double now_us(void) {
  struct timeval tv;
  gettimeofday(&tv, NULL);
  return tv.tv_sec * 1e6 + tv.tv_usec;
}

int main(int argc, char** argv) {
  const char* algo = "NPO";
  int nthreads = 2, reps = 1, warmup = 0;
  long long r_size = 1 << 20, s_size = 1 << 22;
  unsigned r_seed = 12345, s_seed = 54321; /* mc/src/main.c:343-344 */
  for (int i = 1; i < argc; i++) {
    if (!strncmp(argv[i], "--algo=", 7)) algo = argv[i] + 7;
    else if (!strncmp(argv[i], "--nthreads=", 11)) nthreads = atoi(argv[i] + 11);
    else if (!strncmp(argv[i], "--r-size=", 9)) r_size = atoll(argv[i] + 9);
    else if (!strncmp(argv[i], "--s-size=", 9)) s_size = atoll(argv[i] + 9);
    else if (!strncmp(argv[i], "--reps=", 7)) reps = atoi(argv[i] + 7);
    else if (!strncmp(argv[i], "--warmup=", 9)) warmup = atoi(argv[i] + 9);
    else { fprintf(stderr, "unknown flag %s\n", argv[i]); return 2; }
  }
  if (r_size <= 0 || s_size <= 0 || r_size > 0x7fffffffLL || s_size > 0x7fffffffLL) {
    fprintf(stderr, "sizes must fit int32_t (mc/src/generator.h)\n");
    return 2;
  }
  relation_t relR, relS;
  double t0 = now_us();
  seed_generator(r_seed);
  create_relation_pk(&relR, (int32_t)r_size);
  seed_generator(s_seed);
  create_relation_fk(&relS, (int32_t)s_size, (int32_t)r_size);
  printf("GENERATED R %lld S %lld GEN_US %.0f\n", r_size, s_size, now_us() - t0);
  fflush(stdout);
  for (int i = 0; i < warmup + reps; i++) {
    double a = now_us();
    int64_t res;
    if (!strcmp(algo, "NPO")) res = NPO(&relR, &relS, nthreads);
    else if (!strcmp(algo, "PRO")) res = PRO(&relR, &relS, nthreads);
    else { fprintf(stderr, "unknown algo %s\n", algo); return 2; }
    double b = now_us();
    fflush(stdout);
    fflush(stderr);
    printf("\nRUN %d ALGO %s RESULT %lld WALL_US %.0f %s\n", i, algo, (long long)res, b - a, i < warmup ? "WARMUP" : "TIMED");
    fflush(stdout);
  }
  delete_relation(&relR);
  delete_relation(&relS);
  return 0;
}
