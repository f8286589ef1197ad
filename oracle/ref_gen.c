/* ref_gen.c -- TEST INFRASTRUCTURE, not the product.
 *
 * Dumps what the reference's own generators produce (mc/src/generator.c, genzipf.c, compiled UNMODIFIED by
 * oracle/Makefile into oracle/_ref/mc_gen) so that the product's bit-identical ports in csrc/datagen.cpp can be
 * pinned on them (tests/test_oracle_cpu.py).  Nothing of the reference is copied here; this file only calls
 *   seed_generator / create_relation_pk / _pk_lshuffle / _fk / _fk_from_pk / _nonunique / _zipf
 * with the argument patterns of mc/src/main.c:362-411 and writes the int32 keys to stdout (binary).
 *
 *   mc_gen pk N SEED | lshuffle N W SEED | fk M MAXID SEED | nonunique N MAXID SEED | zipf M MAXID Z SEED
 *   mc_gen fullrange_s M N RSEED SSEED      (R = nonunique(N, INT_MAX, RSEED); S = fk_from_pk(R, M) under SSEED)
 */
#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "types.h"
#include "generator.h"

static void dump(relation_t* r) {
  for (uint32_t i = 0; i < r->num_tuples; i++) fwrite(&r->tuples[i].key, sizeof(int32_t), 1, stdout);
}

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  relation_t rel, rel2;
  const char* w = argv[1];
  if (!strcmp(w, "pk")) {
    seed_generator(atoi(argv[3]));
    create_relation_pk(&rel, atoi(argv[2]));
  } else if (!strcmp(w, "lshuffle") && argc >= 5) {
    seed_generator(atoi(argv[4]));
    create_relation_pk_lshuffle(&rel, atoi(argv[2]), atoi(argv[3]));
  } else if (!strcmp(w, "fk") && argc >= 5) {
    seed_generator(atoi(argv[4]));
    create_relation_fk(&rel, atoi(argv[2]), atoi(argv[3]));
  } else if (!strcmp(w, "nonunique") && argc >= 5) {
    seed_generator(atoi(argv[4]));
    create_relation_nonunique(&rel, atoi(argv[2]), atoi(argv[3]));
  } else if (!strcmp(w, "zipf") && argc >= 6) {
    seed_generator(atoi(argv[5]));
    create_relation_zipf(&rel, atoi(argv[2]), atoi(argv[3]), atof(argv[4]));
  } else if (!strcmp(w, "fullrange_s") && argc >= 6) {
    seed_generator(atoi(argv[4]));
    create_relation_nonunique(&rel2, atoi(argv[3]), INT_MAX);
    seed_generator(atoi(argv[5]));
    create_relation_fk_from_pk(&rel, &rel2, atoi(argv[2]));
  } else {
    return 2;
  }
  dump(&rel);
  return 0;
}
