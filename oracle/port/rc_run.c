/* oracle/port/rc_run.c — TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into the product).
 * The pixel loop (Tracer::render, tests/regression/raycast/tracer.cpp:249-263) and per-ray hit dumps, single-threaded as
 * written or split over pthreads by rows / ray chunks (work distribution only; per-ray arithmetic is unchanged). */
#include <pthread.h>
#include <stdatomic.h>
#include <stdlib.h>
#include <unistd.h>
#include "rc_layout.h"

typedef void (*item_fn)(void *ctx, uint32_t item, rc_counters *cnt);
typedef struct {
  item_fn fn;
  void *ctx;
  uint32_t n;
  atomic_uint next;
  int want_counts;
  pthread_mutex_t mu;
  rc_counters total;
} pool_t;

static void *pool_worker(void *p) {
  pool_t *P = (pool_t *)p;
  rc_counters local = {0, 0, 0, 0, 0};
  for (;;) {
    uint32_t i = atomic_fetch_add(&P->next, 1u);
    if (i >= P->n) break;
    P->fn(P->ctx, i, P->want_counts ? &local : NULL);
  }
  pthread_mutex_lock(&P->mu);
  P->total.rays += local.rays; P->total.box_tests += local.box_tests; P->total.tri_tests += local.tri_tests;
  P->total.blas_entered += local.blas_entered; P->total.shaded_hits += local.shaded_hits;
  pthread_mutex_unlock(&P->mu);
  return NULL;
}

static void run_items(uint32_t n, int threads, item_fn fn, void *ctx, rc_counters *out) {
  pool_t P;
  P.fn = fn; P.ctx = ctx; P.n = n; P.want_counts = out != NULL;
  atomic_init(&P.next, 0u);
  pthread_mutex_init(&P.mu, NULL);
  memset(&P.total, 0, sizeof(P.total));
  if (threads <= 1) {
    pool_worker(&P);
  } else {
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
    for (int t = 0; t < threads; ++t) pthread_create(&th[t], NULL, pool_worker, &P);
    for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
    free(th);
  }
  pthread_mutex_destroy(&P.mu);
  if (out) *out = P.total;
}

/* ---- Tracer::render ---- */
typedef struct { const rc_kernel_arg *arg; uint32_t *out; uint32_t mode, y0; } render_ctx;

static void render_row(void *c, uint32_t r, rc_counters *cnt) {
  render_ctx *R = (render_ctx *)c;
  const rc_kernel_arg *arg = R->arg;
  uint32_t y = R->y0 + r;
  for (uint32_t x = 0; x < arg->dst_width; ++x) {
    uint32_t out_idx = y * arg->dst_width + x;
    v3 color = V3(0, 0, 0);
    for (uint32_t s = 0; s < arg->samples_per_pixel; ++s) {
      rc_ray ray = rc_generate_ray(x, y, arg);
      if (R->mode == 0) {
        color = vadd(color, rc_trace(&ray, arg, cnt));
      } else {
        uint32_t seed = rc_wang_hash(out_idx * arg->samples_per_pixel + s + 1);
        if (seed == 0) seed = 1;
        color = vadd(color, rc_trace_ext(&ray, arg, R->mode, &seed, cnt));
      }
    }
    R->out[out_idx] = rc_rgb32f_to_rgb8(color);
  }
}

/* rows [y0,y1).  Returns the number of rays traced when mode != 0 (the reference harness's convention), else 0. */
uint64_t port_render(const void *argp, uint32_t *out, uint32_t mode, uint32_t y0, uint32_t y1, int threads) {
  render_ctx R = {(const rc_kernel_arg *)argp, out, mode, y0};
  rc_counters c;
  run_items(y1 - y0, threads, render_row, &R, mode ? &c : NULL);
  return mode ? c.rays : 0;
}

/* same, with full work counters (SURVEY 8d algorithmic bytes/flops); counters: rays, box, tri, blas, shaded */
void port_render_counted(const void *argp, uint32_t *out, uint32_t mode, uint32_t y0, uint32_t y1, int threads, uint64_t *counters5) {
  render_ctx R = {(const rc_kernel_arg *)argp, out, mode, y0};
  rc_counters c;
  run_items(y1 - y0, threads, render_row, &R, &c);
  counters5[0] = c.rays; counters5[1] = c.box_tests; counters5[2] = c.tri_tests; counters5[3] = c.blas_entered; counters5[4] = c.shaded_hits;
}

/* ---- per-ray closest-hit records (what TLASIntersect leaves in ray_hit_t, render.h:226-227) ---- */
typedef struct { const rc_kernel_arg *arg; const float *rays; float *rays_out; uint64_t n; rc_hit *hits; } trace_ctx;
#define CHUNK 4096u

static void trace_chunk(void *c, uint32_t k, rc_counters *cnt) {
  trace_ctx *T = (trace_ctx *)c;
  uint64_t b = (uint64_t)k * CHUNK, e = b + CHUNK < T->n ? b + CHUNK : T->n;
  for (uint64_t i = b; i < e; ++i) {
    rc_ray r;
    if (T->rays) {
      const float *q = T->rays + 6 * i;
      r.orig = V3(q[0], q[1], q[2]);
      r.dir = V3(q[3], q[4], q[5]);
    } else {
      r = rc_generate_ray((uint32_t)(i % T->arg->dst_width), (uint32_t)(i / T->arg->dst_width), T->arg);
      if (T->rays_out) {
        float *q = T->rays_out + 6 * i;
        q[0] = r.orig.x; q[1] = r.orig.y; q[2] = r.orig.z; q[3] = r.dir.x; q[4] = r.dir.y; q[5] = r.dir.z;
      }
    }
    rc_hit hit = {RC_LARGE_FLOAT, {0, 0, 0}, 0, 0};
    rc_tlas_intersect(&r, T->arg, &hit, cnt);
    T->hits[i] = hit;
  }
}

void port_trace_rays(const void *argp, const float *rays, uint64_t n, void *hits, int threads) {
  trace_ctx T = {(const rc_kernel_arg *)argp, rays, NULL, n, (rc_hit *)hits};
  run_items((uint32_t)((n + CHUNK - 1) / CHUNK), threads, trace_chunk, &T, NULL);
}

void port_trace_primary(const void *argp, void *hits, float *rays_out, int threads) {
  const rc_kernel_arg *arg = (const rc_kernel_arg *)argp;
  uint64_t n = (uint64_t)arg->dst_width * arg->dst_height;
  trace_ctx T = {arg, NULL, rays_out, n, (rc_hit *)hits};
  run_items((uint32_t)((n + CHUNK - 1) / CHUNK), threads, trace_chunk, &T, NULL);
}

void port_trace_rays_counted(const void *argp, const float *rays, uint64_t n, void *hits, int threads, uint64_t *counters5) {
  trace_ctx T = {(const rc_kernel_arg *)argp, rays, NULL, n, (rc_hit *)hits};
  rc_counters c;
  run_items((uint32_t)((n + CHUNK - 1) / CHUNK), threads, trace_chunk, &T, &c);
  counters5[0] = c.rays; counters5[1] = c.box_tests; counters5[2] = c.tri_tests; counters5[3] = c.blas_entered; counters5[4] = c.shaded_hits;
}

int port_hw_threads(void) {
  long n = sysconf(_SC_NPROCESSORS_ONLN);
  return n > 0 ? (int)n : 1;
}
