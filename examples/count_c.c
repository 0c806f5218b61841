/* Plain C host of the C ABI (no Python, no PyTorch): what the cgo shim of INTEGRATION.md does, in C.
 * Builds two synthetic fragments with the generator library, loads them, and runs
 * Count(Intersect(Row(f=1), Row(f=2))), Count(Union(...)), a materialised Row and a TopN.
 *
 *   gcc -O2 -Iinclude examples/count_c.c -Lpilosa_b200 -lpilosa_b200 -lpb200_synth -Wl,-rpath,$PWD/pilosa_b200 -o examples/count_c
 *
 * Exit code 0 and "C host ok" on success; any engine error is printed with pb200_last_error. */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "pilosa_b200.h"

/* generator library (pilosa_b200/csrc/synth.cpp) */
void* pbs_gen_bernoulli(uint64_t seed, uint64_t shard, uint64_t row_lo, uint64_t n_rows, double p);
const uint8_t* pbs_data(void* h);
size_t pbs_size(void* h);
void pbs_free(void* h);

#define CHECK(call)                                                                              \
  do {                                                                                           \
    int rc__ = (call);                                                                           \
    if (rc__ != PB200_OK) {                                                                      \
      fprintf(stderr, "%s -> %s: %s\n", #call, pb200_strerror(rc__), ctx ? pb200_last_error(ctx) : ""); \
      return 1;                                                                                  \
    }                                                                                            \
  } while (0)

static pb200_node row(pb200_family fam, uint64_t id) {
  pb200_node n;
  memset(&n, 0, sizeof n);
  n.op = PB200_OP_ROW; n.family = fam; n.row = id;
  return n;
}
static pb200_node nary(uint32_t op, uint32_t nargs) {
  pb200_node n;
  memset(&n, 0, sizeof n);
  n.op = op; n.nargs = nargs;
  return n;
}

int main(void) {
  pb200_ctx* ctx = NULL;
  CHECK(pb200_init(0, 0, &ctx));
  pb200_family f;
  CHECK(pb200_family_create(ctx, &f));
  const uint64_t shards[2] = {0, 1};
  for (int s = 0; s < 2; s++) {
    void* h = pbs_gen_bernoulli(20190416, shards[s], 0, 4, 0.10);
    CHECK(pb200_frag_load(ctx, f, shards[s], pbs_data(h), pbs_size(h)));
    pbs_free(h); /* the library copied: nothing of ours stays referenced */
  }
  pb200_node isect[3] = {row(f, 1), row(f, 2), nary(PB200_OP_INTERSECT, 2)};
  pb200_node uni[3] = {row(f, 1), row(f, 2), nary(PB200_OP_UNION, 2)};
  uint64_t ci = 0, cu = 0, c1 = 0, c2 = 0;
  pb200_node r1 = row(f, 1), r2 = row(f, 2);
  CHECK(pb200_count(ctx, isect, 3, shards, 2, &ci));
  CHECK(pb200_count(ctx, uni, 3, shards, 2, &cu));
  CHECK(pb200_count(ctx, &r1, 1, shards, 2, &c1));
  CHECK(pb200_count(ctx, &r2, 1, shards, 2, &c2));
  printf("|r1| = %llu  |r2| = %llu  |r1 & r2| = %llu  |r1 | r2| = %llu\n", (unsigned long long)c1, (unsigned long long)c2,
         (unsigned long long)ci, (unsigned long long)cu);
  if (c1 + c2 != ci + cu || ci == 0) { fprintf(stderr, "inclusion-exclusion violated\n"); return 1; }

  /* the intersection as per-shard Pilosa roaring streams (typed and serialised on the device) */
  size_t cap = 2 * (131072 + 536), len = 0;
  void* buf = malloc(cap);
  uint64_t off[3];
  CHECK(pb200_eval_shards(ctx, isect, 3, shards, 2, buf, cap, off, &len));
  printf("row segments: %zu bytes (%llu + %llu)\n", len, (unsigned long long)(off[1] - off[0]), (unsigned long long)(off[2] - off[1]));
  uint64_t n_cont = 0, n_bits = 0;
  char err[128];
  CHECK(pb200_host_check_roaring((const uint8_t*)buf + off[0], (size_t)(off[1] - off[0]), &n_cont, &n_bits, err, sizeof err));
  uint64_t per_shard[2];
  CHECK(pb200_count_shards(ctx, isect, 3, shards, 2, per_shard));
  if (n_bits != per_shard[0]) { fprintf(stderr, "segment of shard 0 holds %llu bits, count says %llu\n", (unsigned long long)n_bits, (unsigned long long)per_shard[0]); return 1; }
  free(buf);

  /* TopN(f, Row(f=1), n=2): row 1 itself must come first with its own cardinality */
  pb200_pair top[8];
  size_t n_top = 0;
  CHECK(pb200_topn(ctx, f, &r1, 1, 2, NULL, 0, 0, 0, 0, shards, 2, top, 8, &n_top));
  if (n_top != 2 || top[0].id != 1 || top[0].count != c1) { fprintf(stderr, "unexpected TopN result\n"); return 1; }
  pb200_timing t;
  CHECK(pb200_last_timing(ctx, &t));
  printf("TopN -> (%llu, %llu), (%llu, %llu); last call: %.3f ms on the device, %u launches\n", (unsigned long long)top[0].id,
         (unsigned long long)top[0].count, (unsigned long long)top[1].id, (unsigned long long)top[1].count, t.kernel_ms, t.launches);
  pb200_shutdown(ctx);
  printf("C host ok\n");
  return 0;
}
