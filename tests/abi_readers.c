/* abi_readers.c — the C ABI from plain C: N reader threads call ragb_search on one store (BASELINE configs[0]'s shape:
 * 10k x 384 fp32 rows, cosine + clamp, top-5, one query per call), as the reference's GUI does with one thread per
 * query (rag_query_interface.py:740-744).  Prints "threads N qps X" per run; tests/test_abi_c_gpu.py compiles this
 * with gcc against include/ragb200.h, links libragb200.so and checks that reader throughput scales (per-call stream +
 * workspace inside the library) and that every thread saw the single-thread answers. */
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "ragb200.h"

#define N 10000
#define D 384
#define K 5
#define NQ 64

static ragb_store* S;
static float Q[NQ][D];
static int64_t WANT[NQ][K];
static double SECONDS = 1.0;

static double now(void) {
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return t.tv_sec + 1e-9 * t.tv_nsec;
}
typedef struct { int t, nt; long count, bad; } Arg;

static void* reader(void* p) {
  Arg* a = (Arg*)p;
  const double stop = now() + SECONDS;
  float sc[K];
  int64_t ids[K];
  for (int i = a->t; now() < stop; i += a->nt) {
    const int j = i % NQ;
    if (ragb_search(S, Q[j], 1, K, RAGB_COSINE, -INFINITY, RAGB_CLAMP01, 0, -1, NULL, sc, ids) != RAGB_OK) { a->bad++; break; }
    if (memcmp(ids, WANT[j], sizeof ids) != 0) a->bad++;
    a->count++;
  }
  return NULL;
}

static double run(int nt) {
  pthread_t th[64];
  Arg arg[64];
  const double t0 = now();
  for (int t = 0; t < nt; ++t) { arg[t].t = t; arg[t].nt = nt; arg[t].count = arg[t].bad = 0; pthread_create(&th[t], NULL, reader, &arg[t]); }
  long total = 0, bad = 0;
  for (int t = 0; t < nt; ++t) { pthread_join(th[t], NULL); total += arg[t].count; bad += arg[t].bad; }
  const double qps = total / (now() - t0);
  printf("threads %d qps %.0f bad %ld\n", nt, qps, bad);
  return qps;
}

int main(int argc, char** argv) {
  if (argc > 1) SECONDS = atof(argv[1]);
  if (ragb_store_create(0, D, RAGB_F32, N, &S) != RAGB_OK) { fprintf(stderr, "%s\n", ragb_last_error()); return 2; }
  float* X = (float*)malloc(sizeof(float) * N * D);
  unsigned s = 12345u;
  for (long i = 0; i < (long)N * D; ++i) { s = s * 1664525u + 1013904223u; X[i] = (float)((s >> 8) & 0xffff) / 32768.0f - 1.0f; }
  if (ragb_store_add(S, X, N, NULL) != RAGB_OK) { fprintf(stderr, "%s\n", ragb_last_error()); return 2; }
  for (int j = 0; j < NQ; ++j) {
    memcpy(Q[j], X + (long)(j * 151 % N) * D, sizeof(float) * D);
    for (int i = 0; i < D; i += 7) Q[j][i] += 0.05f;
    float sc[K];
    if (ragb_search(S, Q[j], 1, K, RAGB_COSINE, -INFINITY, RAGB_CLAMP01, 0, -1, NULL, sc, WANT[j]) != RAGB_OK) return 2;
    if (WANT[j][0] != j * 151 % N) { fprintf(stderr, "query %d: nearest row %lld\n", j, (long long)WANT[j][0]); return 3; }
  }
  SECONDS = SECONDS / 4; run(2); SECONDS *= 4;   /* warm-up: contexts */
  run(1); run(2); run(4); run(8);
  ragb_store_destroy(S);
  free(X);
  return 0;
}
