/* oracle/port/rc_shade.c — TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into the product).
 *
 * Closest-hit shading and the bounce loop: Trace (render.h:210-276), texSample (:8-21), diffuseLighting (:58-70),
 * RGB packing (common.h:105-118), WangHash/xorshift (common.h:85-103); plus TraceExt, the shadow-ray / diffuse-bounce
 * extension of SURVEY.md 8(d) that has no reference counterpart (kept expression for expression in lock-step with
 * oracle/ref_harness/ref_raycast.cpp:TraceExt and the CUDA trace_ext).
 */
#include "rc_layout.h"

uint32_t rc_wang_hash(uint32_t s) {            /* common.h:85-91 */
  s = (s ^ 61) ^ (s >> 16);
  s *= 9, s = s ^ (s >> 4);
  s *= 0x27d4eb2d;
  s = s ^ (s >> 15);
  return s;
}
static inline uint32_t random_int(uint32_t *s) {  /* common.h:93-99 */
  *s ^= *s << 13;
  *s ^= *s >> 17;
  *s ^= *s << 5;
  return *s;
}
static inline float random_float(uint32_t *s) { return random_int(s) * 2.3283064365387e-10f; }   /* common.h:101-103 */

uint32_t rc_rgb32f_to_rgb8(v3 c) {             /* common.h:105-110 */
  int r = (int)(std_minf(c.x, 1.f) * 255);
  int g = (int)(std_minf(c.y, 1.f) * 255);
  int b = (int)(std_minf(c.z, 1.f) * 255);
  return (uint32_t)((r << 16) + (g << 8) + b);
}
static inline v3 rgb8_to_rgb32f(uint32_t c) {  /* common.h:112-118 — divides by 256 */
  float s = 1 / 256.0f;
  int r = (c >> 16) & 255, g = (c >> 8) & 255, b = c & 255;
  return V3(r * s, g * s, b * s);
}
static inline v3 tex_sample(v2 uv, const uint32_t *pixels, uint32_t width, uint32_t height) {   /* render.h:8-21 */
  uint32_t iu = (uint32_t)(uv.x * width);
  uint32_t iv = (uint32_t)(uv.y * height);
  iu %= width;
  iv %= height;
  return rgb8_to_rgb32f(pixels[iu + iv * width]);
}

typedef struct { v3 I, N, texColor; float reflectivity; } surface_t;

static surface_t surface_at(const rc_ray *ray, const rc_hit *hit, const rc_kernel_arg *arg) {   /* render.h:233-250 */
  const rc_blas_node *blas = (const rc_blas_node *)(uintptr_t)arg->blas_addr + hit->blasIdx;
  const rc_tri_ex *ex = (const rc_tri_ex *)(uintptr_t)arg->triEx_addr + hit->triIdx;
  const uint8_t *tex = (const uint8_t *)(uintptr_t)arg->tex_addr;
  surface_t s;
  s.I = vadd(ray->orig, vscale(ray->dir, hit->dist));
  v3 N = vadd(vadd(vscale(ex->N1, hit->bcoords.x), vscale(ex->N2, hit->bcoords.y)), vscale(ex->N0, hit->bcoords.z));
  m4 invT = m4_transposed3(&blas->invTransform);
  s.N = vnormalize(xform_vec(N, &invT));
  v2 uv;
  uv.x = ex->uv1.x * hit->bcoords.x + ex->uv2.x * hit->bcoords.y + ex->uv0.x * hit->bcoords.z;
  uv.y = ex->uv1.y * hit->bcoords.x + ex->uv2.y * hit->bcoords.y + ex->uv0.y * hit->bcoords.z;
  s.texColor = tex_sample(uv, (const uint32_t *)(tex + blas->tex_offset), blas->tex_width, blas->tex_height);
  s.reflectivity = blas->reflectivity;
  return s;
}

/* Trace — render.h:210-276 */
v3 rc_trace(const rc_ray *ray, const rc_kernel_arg *arg, rc_counters *cnt) {
  rc_ray cur = *ray;
  v3 radiance = V3(0, 0, 0);
  float throughput = 1.0f;
  for (uint32_t bounce = 0; bounce < arg->max_depth; ++bounce) {
    rc_hit hit = {RC_LARGE_FLOAT, {0, 0, 0}, 0, 0};
    rc_tlas_intersect(&cur, arg, &hit, cnt);
    if (hit.dist == RC_LARGE_FLOAT) {
      radiance = vadd(radiance, vscale(arg->background_color, throughput));
      break;
    }
    if (cnt) cnt->shaded_hits++;
    surface_t s = surface_at(&cur, &hit, arg);
    /* diffuseLighting (render.h:58-70) */
    v3 L = vsub(arg->light_pos, s.I);
    float dist = vlength(L);
    L = vscale(L, 1.0f / dist);
    float att = 1.0f / (1.0f + dist * 0.1f);
    float NdotL = std_maxf(0.0f, vdot(s.N, L));
    v3 diffuse = vmul(s.texColor, vadd(arg->ambient_color, vscale(sscale(att, arg->light_color), NdotL)));
    radiance = vadd(radiance, vscale(sscale(throughput, diffuse), 1 - s.reflectivity));
    throughput *= s.reflectivity;
    if (s.reflectivity > 0.0f && bounce + 1 < arg->max_depth) {
      v3 R = vnormalize(vsub(cur.dir, vscale(sscale(2.0f, s.N), vdot(s.N, cur.dir))));
      cur.orig = vadd(s.I, vscale(R, 0.001f));
      cur.dir = R;
      continue;
    }
    radiance = vadd(radiance, sscale(throughput, arg->background_color));
    break;
  }
  return radiance;
}

static int occluded(v3 O, v3 D, float maxd, const rc_kernel_arg *arg, rc_counters *cnt) {
  rc_hit sh = {maxd, {0, 0, 0}, 0, 0};
  rc_ray r = {O, D};
  rc_tlas_intersect(&r, arg, &sh, cnt);
  return sh.dist < maxd;
}

/* TraceExt — extension: mode bit0 = shadow rays, bit1 = diffuse bounces */
v3 rc_trace_ext(const rc_ray *ray, const rc_kernel_arg *arg, uint32_t mode, uint32_t *seed, rc_counters *cnt) {
  rc_ray cur = *ray;
  v3 radiance = V3(0, 0, 0);
  v3 thr = V3(1, 1, 1);
  for (uint32_t bounce = 0; bounce < arg->max_depth; ++bounce) {
    rc_hit hit = {RC_LARGE_FLOAT, {0, 0, 0}, 0, 0};
    rc_tlas_intersect(&cur, arg, &hit, cnt);
    if (hit.dist == RC_LARGE_FLOAT) {
      radiance = vadd(radiance, vmul(arg->background_color, thr));
      break;
    }
    if (cnt) cnt->shaded_hits++;
    surface_t s = surface_at(&cur, &hit, arg);
    v3 Ns = (vdot(s.N, cur.dir) > 0.0f) ? vneg(s.N) : s.N;
    v3 O = vadd(s.I, vscale(Ns, 0.001f));
    v3 L = vsub(arg->light_pos, s.I);
    float dist = vlength(L);
    L = vscale(L, 1.0f / dist);
    float att = 1.0f / (1.0f + dist * 0.1f);
    float NdotL = std_maxf(0.0f, vdot(s.N, L));
    if ((mode & 1u) && NdotL > 0.0f) {
      v3 Lv = vsub(arg->light_pos, O);
      float d2 = vlength(Lv);
      v3 D = vscale(Lv, 1.0f / d2);
      if (occluded(O, D, d2, arg, cnt)) NdotL = 0.0f;
    }
    v3 diffuse = vmul(s.texColor, vadd(arg->ambient_color, vscale(sscale(att, arg->light_color), NdotL)));
    if (mode & 2u) {
      radiance = vadd(radiance, vmul(thr, diffuse));
      thr = vmul(thr, s.texColor);
      if (bounce + 1 < arg->max_depth) {
        float px, py, pz, l2;
        do {
          px = random_float(seed) * 2.0f - 1.0f;
          py = random_float(seed) * 2.0f - 1.0f;
          pz = random_float(seed) * 2.0f - 1.0f;
          l2 = px * px + py * py + pz * pz;
        } while (l2 > 1.0f || l2 < 1e-6f);
        float il = 1.0f / sqrtf(l2);
        v3 D = vadd(Ns, V3(px * il, py * il, pz * il));
        float dl2 = vdot(D, D);
        if (dl2 < 1e-12f) D = Ns;
        else D = vscale(D, 1.0f / sqrtf(dl2));
        cur.orig = O;
        cur.dir = D;
        continue;
      }
      break;
    }
    radiance = vadd(radiance, vscale(vmul(thr, diffuse), 1 - s.reflectivity));
    thr = vscale(thr, s.reflectivity);
    if (s.reflectivity > 0.0f && bounce + 1 < arg->max_depth) {
      v3 R = vnormalize(vsub(cur.dir, vscale(sscale(2.0f, s.N), vdot(s.N, cur.dir))));
      cur.orig = vadd(s.I, vscale(R, 0.001f));
      cur.dir = R;
      continue;
    }
    radiance = vadd(radiance, vmul(thr, arg->background_color));
    break;
  }
  return radiance;
}
