/* Many host threads, one context: N threads each issue M single pb200_count calls (what the reference's worker
 * goroutines / concurrent requests would do through a cgo shim). Prints the aggregate rate for 1 thread and for N
 * threads and checks every answer against the single-threaded one.
 *
 *   gcc -O2 -pthread -Iinclude examples/concurrent_c.c -Lpilosa_b200 -lpilosa_b200 -lpb200_synth -Wl,-rpath,$PWD/pilosa_b200 -o examples/concurrent_c
 *   examples/concurrent_c [threads=16] [calls per thread=1000]
 */
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "pilosa_b200.h"

void* pbs_gen_bernoulli(uint64_t seed, uint64_t shard, uint64_t row_lo, uint64_t n_rows, double p);
const uint8_t* pbs_data(void* h);
size_t pbs_size(void* h);
void pbs_free(void* h);

enum { ROWS = 32, SHARDS = 16 };
static pb200_ctx* ctx;
static pb200_family fam;
static uint64_t shards[SHARDS];
static uint64_t expect[ROWS][ROWS];

typedef struct { int id, calls, bad; } job_t;

static void make_prog(pb200_node* p, uint64_t a, uint64_t b) {
  memset(p, 0, 3 * sizeof *p);
  p[0].op = PB200_OP_ROW; p[0].family = fam; p[0].row = a;
  p[1].op = PB200_OP_ROW; p[1].family = fam; p[1].row = b;
  p[2].op = PB200_OP_INTERSECT; p[2].nargs = 2;
}

static void* worker(void* arg) {
  job_t* j = (job_t*)arg;
  uint64_t x = 88172645463325252ull + (uint64_t)j->id * 7919;
  for (int k = 0; k < j->calls; k++) {
    x ^= x << 13; x ^= x >> 7; x ^= x << 17;
    const uint64_t a = x % ROWS, b = (x >> 20) % ROWS;
    pb200_node p[3];
    make_prog(p, a, b);
    uint64_t out = ~0ull;
    if (pb200_count(ctx, p, 3, shards, SHARDS, &out) != PB200_OK || out != expect[a][b]) j->bad++;
  }
  return NULL;
}

static double run(int threads, int calls) {
  pthread_t th[256];
  job_t jobs[256];
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (int i = 0; i < threads; i++) { jobs[i].id = i; jobs[i].calls = calls; jobs[i].bad = 0; pthread_create(&th[i], NULL, worker, &jobs[i]); }
  int bad = 0;
  for (int i = 0; i < threads; i++) { pthread_join(th[i], NULL); bad += jobs[i].bad; }
  clock_gettime(CLOCK_MONOTONIC, &t1);
  if (bad) { fprintf(stderr, "%d wrong answers with %d threads\n", bad, threads); exit(1); }
  const double s = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
  return (double)threads * calls / s;
}

int main(int argc, char** argv) {
  const int threads = argc > 1 ? atoi(argv[1]) : 16, calls = argc > 2 ? atoi(argv[2]) : 1000;
  if (threads < 1 || threads > 256) return 2;
  if (pb200_init(0, 0, &ctx) != PB200_OK) { fprintf(stderr, "pb200_init failed (a CUDA device is required)\n"); return 1; }
  pb200_family_create(ctx, &fam);
  for (int s = 0; s < SHARDS; s++) {
    shards[s] = (uint64_t)s;
    void* h = pbs_gen_bernoulli(20190416, (uint64_t)s, 0, ROWS, 0.10);
    if (pb200_frag_load(ctx, fam, (uint64_t)s, pbs_data(h), pbs_size(h)) != PB200_OK) { fprintf(stderr, "load: %s\n", pb200_last_error(ctx)); return 1; }
    pbs_free(h);
  }
  for (uint64_t a = 0; a < ROWS; a++)
    for (uint64_t b = 0; b < ROWS; b++) {
      pb200_node p[3];
      make_prog(p, a, b);
      if (pb200_count(ctx, p, 3, shards, SHARDS, &expect[a][b]) != PB200_OK) { fprintf(stderr, "count: %s\n", pb200_last_error(ctx)); return 1; }
    }
  run(1, 200); /* warm-up */
  const double r1 = run(1, calls);
  uint64_t l0 = 0, q0 = 0, l1 = 0, q1 = 0;
  pb200_combine_stats(ctx, &l0, &q0);
  const double rn = run(threads, calls);
  pb200_combine_stats(ctx, &l1, &q1);
  printf("{\"threads\": %d, \"calls_per_thread\": %d, \"queries_per_s_1_thread\": %.0f, \"queries_per_s_n_threads\": %.0f, \"speedup\": %.2f, "
         "\"call_us_1_thread\": %.2f, \"call_us_n_threads_per_query\": %.2f, \"combined_launches\": %llu, \"queries_per_combined_launch\": %.2f}\n",
         threads, calls, r1, rn, rn / r1, 1e6 / r1, 1e6 / rn, (unsigned long long)(l1 - l0), l1 > l0 ? (double)(q1 - q0) / (double)(l1 - l0) : 0.0);
  pb200_shutdown(ctx);
  return 0;
}
