/* oracle/port/rc_math.h — TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into the product).
 *
 * Plain-C restatement of the FP32 arithmetic of the reference's raycast path.  Every function cites the
 * reference line it follows (paths relative to /root/reference/tests/regression/raycast/).  Compiled with
 * -O2 -ffp-contract=off on x86-64 (SSE2 scalar float, no FMA), i.e. the same rounding as the reference's
 * own host build (g++ -O2, no -march).  Parity pin: oracle/_ref (the reference sources compiled in place)
 * is compared with this port bit for bit by tests/test_oracle.py; golden fixtures in tests/golden/.
 */
#ifndef RC_MATH_H
#define RC_MATH_H
#include <math.h>
#include <stdint.h>
#include <string.h>

#define RC_LARGE_FLOAT 1e30f              /* geometry.h:15 */
#define RC_EPSILON 1e-6f                  /* geometry.h:17 */
#define RC_PI 3.14159265358979323846264f  /* geometry.h:10 */
#define RC_DEG2RAD 0.017453292519943295769236907684886f /* geometry.h:16 */
#define RC_STACK 64                       /* render.h:5 */

typedef struct { float x, y, z; } v3;
typedef struct { float x, y; } v2;
typedef struct { float c[16]; } m4;       /* row-major, translation in cells 3,7,11 (geometry.h:1123-1129) */

static inline v3 V3(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 v3s(float s) { return V3(s, s, s); }
static inline float v3get(v3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
static inline v3 vadd(v3 a, v3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }      /* geometry.h:346 */
static inline v3 vsub(v3 a, v3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }      /* geometry.h:542 */
static inline v3 vmul(v3 a, v3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }      /* geometry.h:715 */
static inline v3 vscale(v3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }       /* geometry.h:721 */
static inline v3 sscale(float s, v3 a) { return V3(s * a.x, s * a.y, s * a.z); }       /* geometry.h:722 */
static inline v3 vdivs(v3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }        /* geometry.h:817 */
static inline v3 vneg(v3 a) { return V3(-a.x, -a.y, -a.z); }                           /* geometry.h:280 */
static inline float vdot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }     /* geometry.h:888 */
static inline v3 vcross(v3 a, v3 b) {                                                  /* geometry.h:952 */
  return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline float vlength(v3 v) { return sqrtf(vdot(v, v)); }                        /* geometry.h:902 */
static inline v3 vnormalize(v3 v) {                                                    /* geometry.h:913-916, rsqrtf :180 */
  float invLen = 1.0f / sqrtf(vdot(v, v));
  return vscale(v, invLen);
}
/* the reference's own scalar fminf/fmaxf (geometry.h:178-179), NOT libm's */
static inline float rc_fminf(float a, float b) { return a < b ? a : b; }
static inline float rc_fmaxf(float a, float b) { return a > b ? a : b; }
static inline v3 vmin(v3 a, v3 b) { return V3(rc_fminf(a.x, b.x), rc_fminf(a.y, b.y), rc_fminf(a.z, b.z)); }  /* :841 */
static inline v3 vmax(v3 a, v3 b) { return V3(rc_fmaxf(a.x, b.x), rc_fmaxf(a.y, b.y), rc_fmaxf(a.z, b.z)); }  /* :851 */
/* std::min / std::max: return the FIRST argument unless the comparison holds (NaN behaviour) */
static inline float std_minf(float a, float b) { return (b < a) ? b : a; }
static inline float std_maxf(float a, float b) { return (a < b) ? b : a; }
static inline int std_mini(int a, int b) { return (b < a) ? b : a; }
static inline int std_maxi(int a, int b) { return (a < b) ? b : a; }

static inline float surface_area(v3 mn, v3 mx) {                                       /* geometry.h:971-974 */
  v3 e = vsub(mx, mn);
  return 2 * (e.x * e.y + e.y * e.z + e.z * e.x);
}

/* ---- mat4_t ---- */
static inline m4 m4_identity(void) {
  m4 r;
  memset(&r, 0, sizeof(r));
  r.c[0] = r.c[5] = r.c[10] = r.c[15] = 1.0f;
  return r;
}
static inline m4 m4_rotate_x(float a) {                                                /* geometry.h:1023-1030 */
  m4 r = m4_identity();
  r.c[5] = cosf(a); r.c[6] = -sinf(a); r.c[9] = sinf(a); r.c[10] = cosf(a);
  return r;
}
static inline m4 m4_rotate_y(float a) {                                                /* geometry.h:1032-1039 */
  m4 r = m4_identity();
  r.c[0] = cosf(a); r.c[2] = sinf(a); r.c[8] = -sinf(a); r.c[10] = cosf(a);
  return r;
}
static inline m4 m4_translate(v3 p) {                                                  /* geometry.h:1131-1137 */
  m4 r = m4_identity();
  r.c[3] = p.x; r.c[7] = p.y; r.c[11] = p.z;
  return r;
}
static inline m4 m4_mul(const m4 *a, const m4 *b) {                                    /* geometry.h:1220-1231 */
  m4 r;
  for (int i = 0; i < 16; i += 4)
    for (int j = 0; j < 4; ++j)
      r.c[i + j] = (a->c[i + 0] * b->c[j + 0]) + (a->c[i + 1] * b->c[j + 4]) + (a->c[i + 2] * b->c[j + 8]) +
                   (a->c[i + 3] * b->c[j + 12]);
  return r;
}
static inline m4 m4_transposed3(const m4 *m) {                                         /* geometry.h:1141-1147: 3x3 only */
  m4 M = m4_identity();
  M.c[0] = m->c[0]; M.c[1] = m->c[4]; M.c[2] = m->c[8];
  M.c[4] = m->c[1]; M.c[5] = m->c[5]; M.c[6] = m->c[9];
  M.c[8] = m->c[2]; M.c[9] = m->c[6]; M.c[10] = m->c[10];
  return M;
}
m4 m4_inverted(const m4 *m);                                                           /* geometry.h:1149-1192 (rc_scene.c) */

/* float4(a,w) * M (geometry.h:1280-1293): all four products per row are kept, "+ m3*w" included */
static inline v3 xform_vec(v3 a, const m4 *M) {                                        /* TransformVector, w = 0 */
  const float *c = M->c;
  return V3(c[0] * a.x + c[1] * a.y + c[2] * a.z + c[3] * 0.0f, c[4] * a.x + c[5] * a.y + c[6] * a.z + c[7] * 0.0f,
            c[8] * a.x + c[9] * a.y + c[10] * a.z + c[11] * 0.0f);
}
static inline v3 xform_pos(v3 a, const m4 *M) {                                        /* TransformPosition, w = 1 */
  const float *c = M->c;
  return V3(c[0] * a.x + c[1] * a.y + c[2] * a.z + c[3] * 1.0f, c[4] * a.x + c[5] * a.y + c[6] * a.z + c[7] * 1.0f,
            c[8] * a.x + c[9] * a.y + c[10] * a.z + c[11] * 1.0f);
}
#endif
