/* Aggregate memcpy bandwidth of the host with 1..N threads (each copying its own 64 MB buffer pair, 8 rounds): the
 * ceiling of the pageable-plane staging in the *_host entry points.  gcc -O2 -pthread scripts/host_memcpy_bw.c */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define BYTES ((size_t)64 << 20)
#define ROUNDS 8
static pthread_barrier_t bar;
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }
static void* worker(void* arg) {
  char* a = malloc(BYTES); char* b = malloc(BYTES);
  memset(a, 1, BYTES); memset(b, 2, BYTES);
  pthread_barrier_wait(&bar);
  for (int r = 0; r < ROUNDS; r++) memcpy(r & 1 ? a : b, r & 1 ? b : a, BYTES);
  pthread_barrier_wait(&bar);
  free(a); free(b);
  return arg;
}
int main(void) {
  int counts[] = {1, 2, 4, 8, 12, 16};
  for (unsigned k = 0; k < sizeof counts / sizeof counts[0]; k++) {
    int n = counts[k];
    pthread_t th[64];
    pthread_barrier_init(&bar, NULL, n + 1);
    for (int i = 0; i < n; i++) pthread_create(&th[i], NULL, worker, NULL);
    pthread_barrier_wait(&bar);
    double t0 = now();
    pthread_barrier_wait(&bar);
    double dt = now() - t0;
    for (int i = 0; i < n; i++) pthread_join(th[i], NULL);
    pthread_barrier_destroy(&bar);
    printf("{\"threads\": %d, \"memcpy_gbs\": %.1f}\n", n, (double)n * ROUNDS * BYTES / dt / 1e9);
  }
  return 0;
}
