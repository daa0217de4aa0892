/* oracle/icon_oracle.c -- TEST INFRASTRUCTURE, not product code.  See icon_oracle.h.
 *
 * Plain-C restatement of the reference's icon_rt hot path.  Every arithmetic
 * expression keeps the reference's operand order and is compiled with
 * -O2 -ffp-contract=off (oracle/Makefile), so results are bit-identical to the
 * reference's own sources compiled with the same flags (checked by
 * tests/test_oracle_vs_ref.py against oracle/_ref/libicon_ref.so).
 * File:line citations are relative to /root/reference.
 */
#include "icon_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { float x, y, z; } v3;
typedef struct { float x, y, z, w; } v4;

/* common/vecmath.h:437-555 -- component-wise ops, dot as (x+y)+z, cross, length */
static inline v3 v3mk(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 v3add(v3 a, v3 b) { return v3mk(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 v3sub(v3 a, v3 b) { return v3mk(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 v3mul(v3 a, v3 b) { return v3mk(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 v3div(v3 a, v3 b) { return v3mk(a.x / b.x, a.y / b.y, a.z / b.z); }
static inline v3 v3scale(v3 a, float s) { return v3mk(a.x * s, a.y * s, a.z * s); }
static inline v3 v3divs(v3 a, float s) { return v3mk(a.x / s, a.y / s, a.z / s); }
static inline v3 v3neg(v3 a) { return v3mk(-a.x, -a.y, -a.z); }
static inline float v3dot(v3 u, v3 v) { return u.x * v.x + u.y * v.y + u.z * v.z; }
static inline v3 v3cross(v3 u, v3 v)
{
  return v3mk(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x);
}
static inline float v3length(v3 u) { return sqrtf(v3dot(u, u)); }
static inline v3 v3normalize(v3 u) { return v3divs(u, sqrtf(v3dot(u, u))); }
static inline v3 v3min(v3 a, v3 b) { return v3mk(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
static inline v3 v3max(v3 a, v3 b) { return v3mk(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }

int orc_sizeof_cell(void) { return (int)sizeof(orc_cell); }

int orc_num_threads(void)
{
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* icon_rt/ICONGrid.h:44-54 toCartesian */
static inline v3 to_cartesian(float r, float lat, float lon)
{
  float x = r * cosf(lat) * cosf(lon);
  float y = r * cosf(lat) * sinf(lon);
  float z = r * sinf(lat);
  return v3mk(x, y, z);
}

void orc_to_cartesian(const float *sph, int n, float *out)
{
  for (int i = 0; i < n; ++i) {
    v3 c = to_cartesian(sph[3 * i], sph[3 * i + 1], sph[3 * i + 2]);
    out[3 * i] = c.x; out[3 * i + 1] = c.y; out[3 * i + 2] = c.z;
  }
}

/* icon_rt/ICONGrid.h:168-174 makePlane: N = cross(b-a,c-a), w = dot(a,N) */
static inline v4 make_plane(v3 a, v3 b, v3 c)
{
  v3 N = v3cross(v3sub(b, a), v3sub(c, a));
  v4 p = {N.x, N.y, N.z, v3dot(a, N)};
  return p;
}

/* icon_rt/ICONGrid.h:176-179 evalPlane: dot(pos,N) - w */
static inline float eval_plane(v4 p, v3 pos)
{
  return v3dot(pos, v3mk(p.x, p.y, p.z)) - p.w;
}

/* icon_rt/ICONGrid.h:187-203 -- the three side planes of a record */
static inline void cell_planes(const orc_cell *cell, v4 p[3])
{
  const float h0 = cell->height[0], h1 = cell->height[cell->numLayers];
  v3 bv1 = to_cartesian(h0, cell->lat[0], cell->lon[0]);
  v3 bv2 = to_cartesian(h0, cell->lat[1], cell->lon[1]);
  v3 bv3 = to_cartesian(h0, cell->lat[2], cell->lon[2]);
  v3 tv1 = to_cartesian(h1, cell->lat[0], cell->lon[0]);
  v3 tv2 = to_cartesian(h1, cell->lat[1], cell->lon[1]);
  v3 tv3 = to_cartesian(h1, cell->lat[2], cell->lon[2]);
  p[0] = make_plane(bv1, bv2, tv2);
  p[1] = make_plane(bv2, bv3, tv3);
  p[2] = make_plane(bv3, bv1, tv1);
}

void orc_cell_planes(const orc_cell *cells, int n, float *planes)
{
  for (int i = 0; i < n; ++i) {
    v4 p[3];
    cell_planes(&cells[i], p);
    memcpy(planes + 12 * i, p, sizeof p);
  }
}

/* icon_rt/ICONGrid.h:117-144 findHeight (the live #if 1 branch): lower-bound search,
 * first i with hpos <= height[i+1], probing exactly like the reference. */
int orc_find_height(const orc_cell *cell, float hpos)
{
  const float *it;
  const float *first = cell->height;
  const float *last = cell->height + cell->numLayers;
  int count, step;
  count = (int)(last - first);
  while (count > 0) {
    it = first;
    step = count / 2;
    it = it + step;
    if (!(hpos <= *(it + 1))) {
      first = ++it;
      count -= step + 1;
    } else
      count = step;
  }
  return (int)(first - cell->height);
}

void orc_find_height_n(const orc_cell *cell, const float *r, int n, int32_t *layer)
{
  for (int i = 0; i < n; ++i) layer[i] = orc_find_height(cell, r[i]);
}

/* icon_rt/ICONGrid.h:181-208 sample(): radius test (closed), three side planes
 * (reject on > 0), getValue -> value[findHeight(r)] (ICONGrid.h:146-153).
 * toSpherical's asinf/atan2f results (ICONGrid.h:36-42) are unused by sample(), only r. */
int orc_sample(const orc_cell *cell, const float *pos, float *value)
{
  v3 P = v3mk(pos[0], pos[1], pos[2]);
  float r = v3length(P);
  if (r < cell->height[0] || r > cell->height[cell->numLayers]) return 0;
  v4 p[3];
  cell_planes(cell, p);
  if (eval_plane(p[0], P) > 0.f) return 0;
  if (eval_plane(p[1], P) > 0.f) return 0;
  if (eval_plane(p[2], P) > 0.f) return 0;
  *value = cell->value[orc_find_height(cell, r)];
  return 1;
}

/* icon_rt/deviceCode.cu:116-123 -- CPU sampleVolume: first record index that accepts */
static inline int locate1(const orc_cell *cells, int numCells, const float *pos, float *value)
{
  for (int i = 0; i < numCells; ++i)
    if (orc_sample(&cells[i], pos, value)) return i;
  return -1;
}

void orc_locate(const orc_cell *cells, int numCells, const float *pos, int npos,
                int32_t *ids, float *values)
{
#pragma omp parallel for schedule(dynamic, 16)
  for (int i = 0; i < npos; ++i) {
    float v = 0.f;
    int id = locate1(cells, numCells, pos + 3 * i, &v);
    ids[i] = id;
    values[i] = id >= 0 ? v : 0.f;
  }
}

/* icon_rt/deviceCode.cu:127-135 postClassify (swapped lerp weights; opacityScale on
 * the v2 term only; idx = int(v*size), clamped to [0,size-1]) */
static inline int clampi(int x, int a, int b) { int m = x < b ? x : b; return a < m ? m : a; } /* vecmath.h:1338-1341 */
static inline v4 post_classify(const float *lut, int size, float lo, float hi, float opacityScale, float v)
{
  v = (v - lo) / (hi - lo);
  int idx = (int)(v * (float)size);
  float frac = (v * (float)size) - (float)idx;
  const float *a = lut + 4 * clampi(idx, 0, size - 1);
  const float *b = lut + 4 * clampi(idx + 1, 0, size - 1);
  float om = 1.f - frac;
  v4 r;
  r.x = a[0] * frac + b[0] * om * 1.f;
  r.y = a[1] * frac + b[1] * om * 1.f;
  r.z = a[2] * frac + b[2] * om * 1.f;
  r.w = a[3] * frac + b[3] * om * opacityScale;
  return r;
}

void orc_post_classify(const float *lut, int size, float lo, float hi, float opacityScale,
                       const float *v, int n, float *rgba)
{
  for (int i = 0; i < n; ++i) {
    v4 s = post_classify(lut, size, lo, hi, opacityScale, v[i]);
    rgba[4 * i] = s.x; rgba[4 * i + 1] = s.y; rgba[4 * i + 2] = s.z; rgba[4 * i + 3] = s.w;
  }
}

/* icon_rt/ShellAccel.h:34-53 intersectSphere (origin-centred, stable quadratic) */
static inline int intersect_sphere(v3 org, v3 dir, float radius, float *tnear, float *tfar)
{
  float A = v3dot(dir, dir);
  float B = v3dot(dir, org) * 2.f;
  float C = v3dot(org, org) - radius * radius;
  float d = B * B - 4.f * A * C;
  if (d < 0.f) return 0;
  d = sqrtf(d);
  float q = B < 0.f ? -0.5f * (B - d) : -0.5f * (B + d);
  float t1 = q / A;
  float t2 = C / q;
  *tnear = fminf(t1, t2);
  *tfar = fmaxf(t1, t2);
  return 1;
}

void orc_intersect_sphere(const float *org, const float *dir, float radius, int n,
                          int32_t *hit, float *tnear, float *tfar)
{
  for (int i = 0; i < n; ++i) {
    float a = 0.f, b = 0.f;
    hit[i] = intersect_sphere(v3mk(org[3 * i], org[3 * i + 1], org[3 * i + 2]),
                              v3mk(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]), radius, &a, &b);
    tnear[i] = a; tfar[i] = b;
  }
}

/* common/dvr_course-common-both.h:42-88 LCG<4>: 4-round TEA seed, then LCG */
typedef struct { uint32_t state; } lcg;
static inline void lcg_init(lcg *g, uint32_t val0, uint32_t val1)
{
  uint32_t v0 = val0, v1 = val1, s0 = 0;
  for (unsigned n = 0; n < 4; n++) {
    s0 += 0x9e3779b9u;
    v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
    v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
  }
  g->state = v0;
}
static inline float lcg_next(lcg *g)
{
  g->state = 1664525u * g->state + 1013904223u;
  return (float)(g->state & 0x00FFFFFFu) / (float)0x01000000;
}

void orc_lcg(unsigned v0, unsigned v1, int n, float *out)
{
  lcg g;
  lcg_init(&g, v0, v1);
  for (int i = 0; i < n; ++i) out[i] = lcg_next(&g);
}

/* icon_rt/deviceCode.cu:36-49 generateRay; jitter==0 replaces both rnd() draws by 0
 * (SURVEY.md A.3). */
static inline void generate_ray(const orc_scene *sc, float su, float sv, lcg *rnd, v3 *org, v3 *dir)
{
  v3 d00 = v3mk(sc->dir00[0], sc->dir00[1], sc->dir00[2]);
  v3 du = v3mk(sc->dirDu[0], sc->dirDu[1], sc->dirDu[2]);
  v3 dv = v3mk(sc->dirDv[0], sc->dirDv[1], sc->dirDv[2]);
  /* The two rnd() calls of deviceCode.cu:42-43 are unsequenced in C++; the reference as built
   * by g++ (oracle/_ref) draws the dir_dv term's sample first, then the dir_du term's. */
  float jv = sc->jitter ? lcg_next(rnd) : 0.f;
  float b = sv + jv;
  float ju = sc->jitter ? lcg_next(rnd) : 0.f;
  float a = su + ju;
  /* float*vec3f goes through vec3f(float) (vecmath.h:414): component-wise a*du */
  v3 d = v3add(v3add(d00, v3mul(v3mk(a, a, a), du)), v3mul(v3mk(b, b, b), dv));
  d = v3normalize(d);
  if (fabsf(d.x) < 1e-5f) d.x = 1e-5f;
  if (fabsf(d.y) < 1e-5f) d.y = 1e-5f;
  if (fabsf(d.z) < 1e-5f) d.z = 1e-5f;
  *org = v3mk(sc->org[0], sc->org[1], sc->org[2]);
  *dir = d;
}

/* seeds as in the ray-gen programs, deviceCode.cu:246-249 / 288-291 */
static inline void pixel_ray(const orc_scene *sc, int x, int y, int W, int H, lcg *rnd, v3 *org, v3 *dir)
{
  lcg_init(rnd, (uint32_t)(sc->accumID * W * H) + (uint32_t)x, (uint32_t)y);
  generate_ray(sc, (float)x + .5f, (float)y + .5f, rnd, org, dir);
}

void orc_generate_rays(const orc_scene *sc, const int32_t *xy, int n, int W, int H,
                       float *org, float *dir)
{
  for (int i = 0; i < n; ++i) {
    lcg rnd;
    v3 o, d;
    pixel_ray(sc, xy[2 * i], xy[2 * i + 1], W, H, &rnd, &o, &d);
    org[3 * i] = o.x; org[3 * i + 1] = o.y; org[3 * i + 2] = o.z;
    dir[3 * i] = d.x; dir[3 * i + 1] = d.y; dir[3 * i + 2] = d.z;
  }
}

/* Segment selection of sdda, icon_rt/ShellAccel.h:85-111 (+ :114 empty test
 * `upper <= lower`, vecmath.h:981-983), then the A.3 shrink by sceneEPS and clamp to
 * [max(tmin,0), tmax] (the clamp is the documented divergence for cameras inside the shell). */
static inline int shell_segments(v3 org, v3 dir, float tmin, float tmax, float rInner, float rOuter,
                                 float seg[2][2])
{
  const float sceneEPS = rInner * 1e-6f;
  float t1 = 0.f, t2 = 0.f, t3 = 0.f, t4 = 0.f;
  int s1 = intersect_sphere(org, dir, rOuter, &t1, &t4);
  int s2 = intersect_sphere(org, dir, rInner, &t2, &t3);
  if (!s1) return 0;
  if (t4 < tmin) return 0;
  float r[2][2] = {{INFINITY, -INFINITY}, {INFINITY, -INFINITY}};
  if (s1 && !s2) { r[0][0] = t1; r[0][1] = t4; }
  else if (tmin < t2) { r[0][0] = t1; r[0][1] = t2; r[1][0] = t3; r[1][1] = t4; }
  else { r[0][0] = t3; r[0][1] = t4; }
  int n = 0;
  const float tlo = fmaxf(tmin, 0.f);
  for (int i = 0; i < 2; ++i) {
    if (r[i][1] <= r[i][0]) break;
    seg[n][0] = fmaxf(r[i][0] + sceneEPS, tlo);
    seg[n][1] = fminf(r[i][1] - sceneEPS, tmax);
    ++n;
  }
  return n;
}

int orc_shell_segments(const float *org, const float *dir, float tmin, float tmax,
                       float rInner, float rOuter, float *seg4)
{
  float seg[2][2];
  int n = shell_segments(v3mk(org[0], org[1], org[2]), v3mk(dir[0], dir[1], dir[2]), tmin, tmax,
                         rInner, rOuter, seg);
  for (int i = 0; i < n; ++i) { seg4[2 * i] = seg[i][0]; seg4[2 * i + 1] = seg[i][1]; }
  return n;
}

/* common/dvr_course-common-both.h:30-35 linear_to_srgb, :89-110 make_8bit / make_rgba(vec4f) */
static inline float linear_to_srgb(float x)
{
  if (x <= 0.0031308f) return 12.92f * x;
  return 1.055f * powf(x, 1.f / 2.4f) - 0.055f;
}
static inline uint32_t make_8bit(float f) { return (uint32_t)fminf(255.f, fmaxf(0.f, (float)(int)(f * 256.f))); }
static inline uint32_t srgb_pack(v4 c)
{
  c.x = linear_to_srgb(c.x); c.y = linear_to_srgb(c.y); c.z = linear_to_srgb(c.z);
  return (make_8bit(c.x) << 0) + (make_8bit(c.y) << 8) + (make_8bit(c.z) << 16) + (make_8bit(c.w) << 24);
}

void orc_srgb_pack(const float *rgba, int n, uint32_t *out)
{
  for (int i = 0; i < n; ++i) {
    v4 c = {rgba[4 * i], rgba[4 * i + 1], rgba[4 * i + 2], rgba[4 * i + 3]};
    out[i] = srgb_pack(c);
  }
}

/* icon_rt/ICONGrid.h:78-115 ICONCell::getBounds and the host loop icon_rt/hostCode.cu:792-808 */
void orc_grid_bounds(const orc_cell *cells, int n, float *volLo, float *volHi,
                     float *sphLo, float *sphHi, float *dataRange)
{
  v3 vlo = v3mk(INFINITY, INFINITY, INFINITY), vhi = v3mk(-INFINITY, -INFINITY, -INFINITY);
  v3 slo = vlo, shi = vhi;
  float dlo = INFINITY, dhi = -INFINITY;
  for (int i = 0; i < n; ++i) {
    const orc_cell *c = &cells[i];
    float minLat = fminf(c->lat[0], fminf(c->lat[1], c->lat[2]));
    float maxLat = fmaxf(c->lat[0], fmaxf(c->lat[1], c->lat[2]));
    float minLon = fminf(c->lon[0], fminf(c->lon[1], c->lon[2]));
    float maxLon = fmaxf(c->lon[0], fmaxf(c->lon[1], c->lon[2]));
    slo.x = fminf(slo.x, c->height[0]);
    shi.x = fmaxf(shi.x, c->height[c->numLayers]);
    slo.y = fminf(slo.y, minLat); shi.y = fmaxf(shi.y, maxLat);
    slo.z = fminf(slo.z, minLon); shi.z = fmaxf(shi.z, maxLon);
    /* getBounds */
    v3 blo = v3mk(INFINITY, INFINITY, INFINITY), bhi = v3mk(-INFINITY, -INFINITY, -INFINITY);
    const float h0 = c->height[0], R = c->height[c->numLayers];
    for (int k = 0; k < 3; ++k) {
      v3 bv = to_cartesian(h0, c->lat[k], c->lon[k]);
      blo = v3min(blo, bv); bhi = v3max(bhi, bv);
    }
    v3 tv1 = to_cartesian(R, c->lat[0], c->lon[0]);
    v3 tv2 = to_cartesian(R, c->lat[1], c->lon[1]);
    v3 tv3 = to_cartesian(R, c->lat[2], c->lon[2]);
    v3 bary = v3divs(v3add(v3add(tv1, tv2), tv3), 3.f);
    float D = R - v3length(bary);
    float off = D / R;
    tv1 = v3add(tv1, v3scale(tv1, off));
    tv2 = v3add(tv2, v3scale(tv2, off));
    tv3 = v3add(tv3, v3scale(tv3, off));
    blo = v3min(blo, tv1); bhi = v3max(bhi, tv1);
    blo = v3min(blo, tv2); bhi = v3max(bhi, tv2);
    blo = v3min(blo, tv3); bhi = v3max(bhi, tv3);
    vlo = v3min(vlo, blo); vhi = v3max(vhi, bhi);
    for (int j = 0; j < c->numLayers; ++j) {   /* box1f::extend */
      dlo = fminf(dlo, c->value[j]); dhi = fmaxf(dhi, c->value[j]);
    }
  }
  volLo[0] = vlo.x; volLo[1] = vlo.y; volLo[2] = vlo.z;
  volHi[0] = vhi.x; volHi[1] = vhi.y; volHi[2] = vhi.z;
  sphLo[0] = slo.x; sphLo[1] = slo.y; sphLo[2] = slo.z;
  sphHi[0] = shi.x; sphHi[1] = shi.y; sphHi[2] = shi.z;
  dataRange[0] = dlo; dataRange[1] = dhi;
}

/* common/camera.h:34-60 setOrientation + forceUpFrame, :86-96 getScreen, and the
 * launch-param camera of icon_rt/hostCode.cu:936-945 (du = horizontal/W, dv = vertical/H) */
static void camera_screen(v3 origin, v3 poi, v3 up, float fovy, float aspect, int W, int H,
                          float *org, float *d00, float *du, float *dv)
{
  v3 vx, vy, vz;
  if (poi.x == origin.x && poi.y == origin.y && poi.z == origin.z) vz = v3mk(0, 0, 1);
  else vz = v3neg(v3normalize(v3sub(poi, origin)));
  vx = v3cross(up, vz);
  if (v3dot(vx, vx) < 1e-8f) vx = v3mk(0, 1, 0);
  else vx = v3normalize(vx);
  vy = v3normalize(v3cross(vz, vx));
  if (!(fabsf(v3dot(vz, up)) < 1e-6f)) {          /* forceUpFrame */
    vx = v3normalize(v3cross(up, vz));
    vy = v3normalize(v3cross(vz, vx));
  }
  float screen_height = 2.f * tanf(0.5f * fovy);
  v3 vertical = v3mul(v3mk(screen_height, screen_height, screen_height), vy);
  float sa = screen_height * aspect;
  v3 horizontal = v3mul(v3mk(sa, sa, sa), vx);
  v3 ll = v3sub(v3sub(v3neg(vz), v3mul(v3mk(.5f, .5f, .5f), vertical)),
                v3mul(v3mk(.5f, .5f, .5f), horizontal));
  v3 u = v3divs(horizontal, (float)W), v = v3divs(vertical, (float)H);
  org[0] = origin.x; org[1] = origin.y; org[2] = origin.z;
  d00[0] = ll.x; d00[1] = ll.y; d00[2] = ll.z;
  du[0] = u.x; du[1] = u.y; du[2] = u.z;
  dv[0] = v.x; dv[1] = v.y; dv[2] = v.z;
}

void orc_camera_orient(const float *eye, const float *poi, const float *up, float fovy,
                       float aspect, int W, int H, float *org, float *d00, float *du, float *dv)
{
  camera_screen(v3mk(eye[0], eye[1], eye[2]), v3mk(poi[0], poi[1], poi[2]), v3mk(up[0], up[1], up[2]),
                fovy, aspect, W, H, org, d00, du, dv);
}

/* common/camera.h:98-104 viewAll: eye = centre + (0,0, r + r/atan(fovy)), up = +y */
void orc_camera_view_all(const float *volLo, const float *volHi, float fovy, float aspect,
                         int W, int H, float *org, float *d00, float *du, float *dv)
{
  v3 lo = v3mk(volLo[0], volLo[1], volLo[2]), hi = v3mk(volHi[0], volHi[1], volHi[2]);
  float diagonal = v3length(v3sub(hi, lo));
  float r = diagonal * 0.5f;
  v3 center = v3divs(v3add(lo, hi), 2.f);
  v3 eye = v3add(center, v3mk(0.f, 0.f, r + r / atanf(fovy)));
  camera_screen(eye, center, v3mk(0, 1, 0), fovy, aspect, W, H, org, d00, du, dv);
}

/* common/dvr_course-common.h:44-70 resampleLUT (lerp(a,b,x) = x*a + (1-x)*b, vecmath.h:56-59) */
void orc_resample_lut(const float *src, int nsrc, float *dst, int ndst)
{
  for (int i = 0; i < ndst; ++i) {
    float indexf = (float)i / (float)ndst * (float)(nsrc - 1);
    int indexa = (int)indexf;
    int indexb = indexa + 1 < nsrc - 1 ? indexa + 1 : nsrc - 1;
    float frac = indexf - (float)indexa;
    float x = 1.f - frac;
    for (int k = 0; k < 4; ++k)
      dst[4 * i + k] = x * src[4 * indexa + k] + (1.f - x) * src[4 * indexb + k];
  }
}

/* ---- the new integrator, SURVEY.md Appendix A.3 --------------------------------- */

typedef struct {
  v4       C;
  uint64_t hash;
  int      nSteps, nHits;
} pixel_out;

static inline uint64_t fnv1a_step(uint64_t h, int32_t id)
{
  uint32_t u = (uint32_t)id;
  for (int b = 0; b < 4; ++b) {
    h ^= (u >> (8 * b)) & 0xffu;
    h *= 1099511628211ull;
  }
  return h;
}

/* Chunked compositing (sc->chunk > 0; what the CUDA marcher computes, DESIGN.md "split rays").  A ray's samples are
 * grouped into chunks of `chunk` consecutive sample indices per shell segment -- the cut positions are a function of the
 * ray alone.  Each chunk is composited front to back from zero into a LOCAL accumulator (the A.3 recurrence), and the
 * chunks are merged in order with the `over` operator:  C += (1-A)*c_loc,  A = A + (1-A)*a_loc.  Early termination is
 * tested on the merged opacity after every contributing sample, so the ray stops at the same sample as the plain
 * statement would (up to rounding at the threshold).  Mathematically the plain recurrence (over is associative); in fp32
 * the two differ by a few ulps.  Because the arithmetic does not depend on WHO marches a chunk, the device may hand the
 * chunks of a long limb ray to different lanes and still produce these bits. */
static pixel_out march_pixel(const orc_scene *sc, int x, int y, int W, int H, int32_t *trace, int maxTrace)
{
  lcg rnd;
  v3 org, dir;
  pixel_ray(sc, x, y, W, H, &rnd, &org, &dir);
  pixel_out out;
  out.C.x = out.C.y = out.C.z = out.C.w = 0.f;
  out.hash = 1469598103934665603ull;
  out.nSteps = out.nHits = 0;
  float seg[2][2];
  int nseg = shell_segments(org, dir, 0.f, 1e10f, sc->rInner, sc->rOuter, seg);   /* Ray(org,dir,0,1e10) */
  const v3 amb = v3mk(sc->ambientColor[0], sc->ambientColor[1], sc->ambientColor[2]);
  const int chunk = sc->chunk;
  v4 loc = {0.f, 0.f, 0.f, 0.f};      /* the current chunk's own accumulator (chunk > 0) */
  int done = 0;
  for (int s = 0; s < nseg && !done; ++s) {
    for (int i = 0;; ++i) {
      float t = seg[s][0] + ((float)i + 0.5f) * sc->dt;
      if (t >= seg[s][1]) break;
      if (chunk > 0 && i % chunk == 0 && i > 0) {       /* chunk boundary: merge, start the next chunk from zero */
        const float T = 1.f - out.C.w;
        out.C.x += T * loc.x; out.C.y += T * loc.y; out.C.z += T * loc.z; out.C.w += T * loc.w;
        loc.x = loc.y = loc.z = loc.w = 0.f;
      }
      v3 P = v3add(org, v3scale(dir, t));            /* Ray::eval, vecmath.h:1917-1918 */
      float pos[3] = {P.x, P.y, P.z};
      float value = 0.f;
      int id = locate1(sc->cells, sc->numCells, pos, &value);
      if (trace && out.nSteps < maxTrace) trace[out.nSteps] = id;
      out.hash = fnv1a_step(out.hash, id);
      out.nSteps++;
      if (id < 0) continue;
      out.nHits++;
      if (sc->prismVtx) {     /* prism sampler: a value that is not a number adds nothing */
        const int layer = orc_find_height(&sc->cells[id], v3length(P));
        if (!orc_prism_value(&sc->cells[id], sc->prismVtx + (size_t)id * 96, sc->prismGeom + (size_t)id * 6, layer, pos, &value, NULL))
          continue;
      }
      v4 sm = post_classify(sc->lut, sc->lutSize, sc->tfLo, sc->tfHi, sc->opacityScale, value);
      float a = 1.f - expf(-sm.w * sc->dt / sc->unitDistance);
      v3 col = v3scale(v3mul(v3mk(sm.x, sm.y, sm.z), amb), sc->ambientRadiance);
      if (chunk > 0) {
        float w = (1.f - loc.w) * a;
        loc.x += w * col.x;
        loc.y += w * col.y;
        loc.z += w * col.z;
        loc.w += w;
        if (out.C.w + (1.f - out.C.w) * loc.w >= sc->termAlpha) { done = 1; break; }
      } else {
        float w = (1.f - out.C.w) * a;
        out.C.x += w * col.x;
        out.C.y += w * col.y;
        out.C.z += w * col.z;
        out.C.w += w;
        if (out.C.w >= sc->termAlpha) { done = 1; break; }
      }
    }
    if (chunk > 0) {                                   /* end of the segment (or of the ray): merge the open chunk */
      const float T = 1.f - out.C.w;
      out.C.x += T * loc.x; out.C.y += T * loc.y; out.C.z += T * loc.z; out.C.w += T * loc.w;
      loc.x = loc.y = loc.z = loc.w = 0.f;
    }
  }
  return out;
}

/* icon_rt/deviceCode.cu:333-340: running mean into accumBuffer, sRGB + RGBA8 into fbPointer */
static inline void write_pixel(const orc_scene *sc, int p, v4 C, uint32_t *fbPointer, float *accumBuffer)
{
  float accum = 1.f / (float)(sc->accumID + 1);
  float om = 1.f - accum;
  v4 o = {accumBuffer[4 * p], accumBuffer[4 * p + 1], accumBuffer[4 * p + 2], accumBuffer[4 * p + 3]};
  v4 n = {accum * C.x + om * o.x, accum * C.y + om * o.y, accum * C.z + om * o.z, accum * C.w + om * o.w};
  accumBuffer[4 * p] = n.x; accumBuffer[4 * p + 1] = n.y; accumBuffer[4 * p + 2] = n.z; accumBuffer[4 * p + 3] = n.w;
  fbPointer[p] = srgb_pack(n);
}

static double now_s(void)
{
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

long orc_march_frame(const orc_scene *sc, int W, int H, int x0, int y0, int x1, int y1,
                     uint32_t *fbPointer, float *accumBuffer, int32_t *perPixel,
                     double deadline_s, double *totals)
{
  double t0 = now_s();
  long steps = 0, hits = 0, pixels = 0;
  if (sc->accumID == 0) /* common/pipeline.cu:1058-1059 clearFramebuffer */
    for (int y = y0; y < y1; ++y)
      for (int x = x0; x < x1; ++x) {
        int p = x + W * y;
        fbPointer[p] = 0;
        accumBuffer[4 * p] = accumBuffer[4 * p + 1] = accumBuffer[4 * p + 2] = accumBuffer[4 * p + 3] = 0.f;
      }
  const int nx = x1 - x0, ny = y1 - y0;
#pragma omp parallel for schedule(dynamic, 8) reduction(+ : steps, hits, pixels)
  for (int k = 0; k < nx * ny; ++k) {
    if (deadline_s > 0 && now_s() - t0 > deadline_s) continue;
    int x = x0 + k % nx, y = y0 + k / nx;
    pixel_out o = march_pixel(sc, x, y, W, H, NULL, 0);
    int p = x + W * y;
    write_pixel(sc, p, o.C, fbPointer, accumBuffer);
    if (perPixel) {
      perPixel[4 * p + 0] = o.nSteps;
      perPixel[4 * p + 1] = o.nHits;
      perPixel[4 * p + 2] = (int32_t)(uint32_t)(o.hash & 0xffffffffu);
      perPixel[4 * p + 3] = (int32_t)(uint32_t)(o.hash >> 32);
    }
    steps += o.nSteps; hits += o.nHits; pixels += 1;
  }
  if (totals) { totals[0] = (double)steps; totals[1] = (double)hits; totals[2] = now_s() - t0; }
  return pixels;
}

int orc_march_trace(const orc_scene *sc, int W, int H, int x, int y,
                    int32_t *ids, int maxIds, float *rgba)
{
  pixel_out o = march_pixel(sc, x, y, W, H, ids, maxIds);
  rgba[0] = o.C.x; rgba[1] = o.C.y; rgba[2] = o.C.z; rgba[3] = o.C.w;
  return o.nSteps;
}

/* ---- prism sampler ---------------------------------------------
 *
 * The reference's wedge sampler (CUBQL mode, deviceCode.cu:90-115) evaluates, for a prism with six
 * vertices V0..V5 (bottom corners 1,2,3 then top corners 1,2,3: hostCode.cu:562-597), the shape functions of
 * UElems.h:176-187 at the parametric coordinates (r,s,t) that a Newton iteration finds for P
 * (UElems.h:214-311), and returns sum_i w_i * V_i.w.  Its six vertices are r_bottom*U_k and r_top*U_k for the
 * three corner directions U_k, so the wedge map is X(r,s,t) = rho(t) * ((1-r-s) U_0 + r U_1 + s U_2) with
 * rho(t) = h_i + t (h_{i+1} - h_i), and the inverse has a closed form: central projection of P onto the chord
 * triangle gives (r,s), the scale factor gives t.  This restatement computes that closed form; the tests
 * pin it against the reference's own intersectWedgeEXT (oracle/_ref) wherever the Newton iteration accepts.
 *
 * Location stays the predicate of ICONGrid.h:181-208 (record id and layer are those of the constant sampler);
 * t is clamped to [0,1] because the predicate's layers are spherical shells while the wedge's are flat
 * (the sag between chord and sphere).
 *
 * Vertex scalars: per column (records with bit-identical lat/lon triples, stacked by height[0]) the
 * interface values I[0] = v[0], I[j] = (v[j-1] + v[j]) * 0.5f, I[L] = v[L-1] (mid-level values in the sense of
 * hostCode.cu:574-575; the reference's getValue(height[h]) resolves to the layer BELOW the interface,
 * ICONGrid.h:129, which shifts its own average by one layer -- not reproduced); a vertex's scalar at
 * interface j is the mean of I_c[j] over the (column, corner) pairs welded to that vertex, summed in
 * ascending (column, corner) order in fp32 and divided by the count.
 */
static inline uint32_t fbits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

/* canonical (lat,lon) of a corner: poles lose their longitude, -pi folds onto +pi, -0 onto +0 */
static uint64_t vertex_key(float lat, float lon)
{
  const float halfPi = 1.57079637f, pi = 3.14159274f;
  if (lat == 0.f) lat = 0.f;
  if (lon == 0.f) lon = 0.f;
  if (fabsf(lat) >= halfPi) lon = 0.f;
  if (lon <= -pi) lon = pi;
  return ((uint64_t)fbits(lat) << 32) | fbits(lon);
}

typedef struct { uint64_t key; uint32_t idx; } keyed;
static int cmp_keyed(const void *a, const void *b)
{
  const keyed *x = (const keyed *)a, *y = (const keyed *)b;
  if (x->key != y->key) return x->key < y->key ? -1 : 1;
  return x->idx < y->idx ? -1 : (x->idx > y->idx ? 1 : 0);
}

static const orc_cell *g_cmpCells;
static int cmp_latlon(const void *a, const void *b)
{
  const int32_t x = *(const int32_t *)a, y = *(const int32_t *)b;
  const int c = memcmp(g_cmpCells[x].lat, g_cmpCells[y].lat, 24);
  if (c) return c;
  return x < y ? -1 : (x > y ? 1 : 0);
}
typedef struct { int32_t col; float h0; int32_t idx; } colrec;
static int cmp_colrec(const void *a, const void *b)
{
  const colrec *x = (const colrec *)a, *y = (const colrec *)b;
  if (x->col != y->col) return x->col < y->col ? -1 : 1;
  if (x->h0 < y->h0) return -1;
  if (y->h0 < x->h0) return 1;
  return x->idx < y->idx ? -1 : (x->idx > y->idx ? 1 : 0);
}
static int cmp_i32(const void *a, const void *b)
{
  const int32_t x = *(const int32_t *)a, y = *(const int32_t *)b;
  return x < y ? -1 : (x > y ? 1 : 0);
}

static inline void unit_from_latlon(float lat, float lon, double v[3])
{
  const double cl = cos((double)lat);
  v[0] = cl * cos((double)lon);
  v[1] = cl * sin((double)lon);
  v[2] = sin((double)lat);
}

/* geom[0..2] = kappa_k = 1 / (U_opp(k) . N_k): evalPlane(p_k, P) * kappa_k is the (unnormalised) barycentric
 * weight of the corner opposite side plane k; geom[3..5] = M = Nt / (U_0 . Nt), Nt the chord triangle's normal:
 * dot(P, M) is the wedge's scale factor rho. */
static void prism_geom(const orc_cell *cell, float *geom)
{
  v4 p[3];
  cell_planes(cell, p);
  double U[3][3];
  for (int k = 0; k < 3; ++k) unit_from_latlon(cell->lat[k], cell->lon[k], U[k]);
  for (int k = 0; k < 3; ++k) {
    const int opp = (k + 2) % 3;       /* p0 spans corners (0,1), p1 (1,2), p2 (2,0) */
    const double d = U[opp][0] * (double)p[k].x + U[opp][1] * (double)p[k].y + U[opp][2] * (double)p[k].z;
    const float kap = (float)(1.0 / d);
    geom[k] = isfinite(kap) ? kap : 0.f;
  }
  const double ax = U[1][0] - U[0][0], ay = U[1][1] - U[0][1], az = U[1][2] - U[0][2];
  const double bx = U[2][0] - U[0][0], by = U[2][1] - U[0][1], bz = U[2][2] - U[0][2];
  const double nx = ay * bz - az * by, ny = az * bx - ax * bz, nz = ax * by - ay * bx;
  const double w = U[0][0] * nx + U[0][1] * ny + U[0][2] * nz;
  const float m[3] = {(float)(nx / w), (float)(ny / w), (float)(nz / w)};
  for (int k = 0; k < 3; ++k) geom[3 + k] = isfinite(m[k]) ? m[k] : 0.f;
}

void orc_prism_build(const orc_cell *cells, int n, float *vtx, float *geom)
{
  memset(vtx, 0, (size_t)n * 96 * sizeof(float));
  for (int i = 0; i < n; ++i) prism_geom(&cells[i], geom + 6 * (size_t)i);
  if (n <= 0) return;
  /* columns: records with bit-identical lat/lon triples, numbered by ascending lowest record index */
  int32_t *idx = (int32_t *)malloc(sizeof(int32_t) * n), *rep = (int32_t *)malloc(sizeof(int32_t) * n);
  int32_t *colOfRep = (int32_t *)malloc(sizeof(int32_t) * n), *reps = (int32_t *)malloc(sizeof(int32_t) * n);
  for (int i = 0; i < n; ++i) idx[i] = i;
  g_cmpCells = cells;
  qsort(idx, n, sizeof(int32_t), cmp_latlon);
  int nCol = 0;
  for (int i = 0; i < n;) {
    int j = i + 1;
    while (j < n && memcmp(cells[idx[j]].lat, cells[idx[i]].lat, 24) == 0) ++j;
    for (int q = i; q < j; ++q) rep[idx[q]] = idx[i];
    reps[nCol++] = idx[i];
    i = j;
  }
  qsort(reps, nCol, sizeof(int32_t), cmp_i32);
  for (int c = 0; c < nCol; ++c) colOfRep[reps[c]] = c;
  /* records of each column stacked by (height[0], index) */
  colrec *cr = (colrec *)malloc(sizeof(colrec) * n);
  for (int i = 0; i < n; ++i) { cr[i].col = colOfRep[rep[i]]; cr[i].h0 = cells[i].height[0]; cr[i].idx = i; }
  qsort(cr, n, sizeof(colrec), cmp_colrec);
  int32_t *colBegin = (int32_t *)calloc((size_t)nCol + 1, sizeof(int32_t));
  int64_t *valOff = (int64_t *)calloc((size_t)nCol + 1, sizeof(int64_t));
  int32_t *base = (int32_t *)malloc(sizeof(int32_t) * n);       /* layers below the record in its column */
  for (int i = 0; i < n; ++i) colBegin[cr[i].col + 1]++;
  for (int c = 0; c < nCol; ++c) colBegin[c + 1] += colBegin[c];
  int64_t total = 0;
  for (int c = 0; c < nCol; ++c) {
    valOff[c] = total;
    int b = 0;
    for (int q = colBegin[c]; q < colBegin[c + 1]; ++q) { base[cr[q].idx] = b; b += cells[cr[q].idx].numLayers; }
    total += b;
  }
  valOff[nCol] = total;
  float *vals = (float *)malloc(sizeof(float) * (size_t)(total > 0 ? total : 1));
  for (int c = 0; c < nCol; ++c) {
    int64_t o = valOff[c];
    for (int q = colBegin[c]; q < colBegin[c + 1]; ++q) {
      const orc_cell *cl = &cells[cr[q].idx];
      for (int l = 0; l < cl->numLayers; ++l) vals[o++] = cl->value[l];
    }
  }
  /* vertex welding */
  keyed *ks = (keyed *)malloc(sizeof(keyed) * (size_t)nCol * 3);
  for (int c = 0; c < nCol; ++c)
    for (int k = 0; k < 3; ++k) {
      ks[c * 3 + k].key = vertex_key(cells[reps[c]].lat[k], cells[reps[c]].lon[k]);
      ks[c * 3 + k].idx = (uint32_t)(c * 3 + k);
    }
  qsort(ks, (size_t)nCol * 3, sizeof(keyed), cmp_keyed);
  int32_t *runOf = (int32_t *)malloc(sizeof(int32_t) * (size_t)nCol * 3);     /* (col,corner) -> start of its vertex run */
  int32_t *runEnd = (int32_t *)malloc(sizeof(int32_t) * (size_t)nCol * 3);
  for (int i = 0; i < nCol * 3;) {
    int j = i + 1;
    while (j < nCol * 3 && ks[j].key == ks[i].key) ++j;
    for (int q = i; q < j; ++q) { runOf[ks[q].idx] = i; runEnd[ks[q].idx] = j; }
    i = j;
  }
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < n; ++i) {
    const int c = colOfRep[rep[i]];
    for (int il = 0; il <= cells[i].numLayers && il < 32; ++il) {
      const int j = base[i] + il;
      for (int k = 0; k < 3; ++k) {
        float acc = 0.f;
        int cnt = 0;
        for (int q = runOf[c * 3 + k]; q < runEnd[c * 3 + k]; ++q) {
          const int oc = (int)(ks[q].idx / 3);
          const int64_t L = valOff[oc + 1] - valOff[oc];
          if (L < 1 || j > L) continue;
          const float *v = vals + valOff[oc];
          const float I = j == 0 ? v[0] : (j == L ? v[L - 1] : (v[j - 1] + v[j]) * 0.5f);
          acc += I;
          ++cnt;
        }
        vtx[(size_t)i * 96 + il * 3 + k] = cnt > 0 ? acc / (float)cnt : 0.f;
      }
    }
  }
  free(idx); free(rep); free(colOfRep); free(reps); free(cr); free(colBegin); free(valOff); free(base); free(vals);
  free(ks); free(runOf); free(runEnd);
}

/* closed-form wedge coordinates + UElems.h:176-187 shape functions + UElems.h:299-303 evaluation.
 * Returns 0 (value untouched) when the value is not a number. */
int orc_prism_value(const orc_cell *cell, const float *vtxRec, const float *geomRec, int layer,
                    const float *pos, float *value, float *pcoords)
{
  v3 P = v3mk(pos[0], pos[1], pos[2]);
  v4 p[3];
  cell_planes(cell, p);
  const float e0 = eval_plane(p[0], P), e1 = eval_plane(p[1], P), e2 = eval_plane(p[2], P);
  const float a0 = e1 * geomRec[1], a1 = e2 * geomRec[2], a2 = e0 * geomRec[0];
  const float sum = (a0 + a1) + a2;
  const float r = a1 / sum, s = a2 / sum;
  const float rho = v3dot(P, v3mk(geomRec[3], geomRec[4], geomRec[5]));
  const int li = layer < 0 ? 0 : (layer > 30 ? 30 : layer);
  const float h0 = cell->height[li], h1 = cell->height[li + 1];
  const float d = h1 - h0;
  float t = d > 0.f ? (rho - h0) / d : 0.f;
  t = fminf(fmaxf(t, 0.f), 1.f);
  const float *B = vtxRec + 3 * li, *T = vtxRec + 3 * (li + 1);
  const float w0 = (1.f - r) - s, om = 1.f - t;
  float val = 0.f;
  val += (w0 * om) * B[0];
  val += (r * om) * B[1];
  val += (s * om) * B[2];
  val += (w0 * t) * T[0];
  val += (r * t) * T[1];
  val += (s * t) * T[2];
  if (pcoords) { pcoords[0] = r; pcoords[1] = s; pcoords[2] = t; }
  if (!(val == val)) return 0;
  *value = val;
  return 1;
}

/* the six wedge vertices (xyz + scalar) of layer `layer`, as buildCuBQLAccel lays them out (hostCode.cu:562-597) */
void orc_prism_wedge(const orc_cell *cell, const float *vtxRec, int layer, float *V24)
{
  for (int half = 0; half < 2; ++half)
    for (int k = 0; k < 3; ++k) {
      v3 c = to_cartesian(cell->height[layer + half], cell->lat[k], cell->lon[k]);
      float *o = V24 + 4 * (3 * half + k);
      o[0] = c.x; o[1] = c.y; o[2] = c.z; o[3] = vtxRec[3 * (layer + half) + k];
    }
}

void orc_locate_prism(const orc_cell *cells, int numCells, const float *vtx, const float *geom,
                      const float *pos, int npos, int32_t *ids, float *values, int32_t *layers, float *pcoords)
{
#pragma omp parallel for schedule(dynamic, 16)
  for (int i = 0; i < npos; ++i) {
    float v = 0.f;
    const int id = locate1(cells, numCells, pos + 3 * i, &v);
    ids[i] = id;
    values[i] = 0.f;
    if (layers) layers[i] = -1;
    if (pcoords) pcoords[3 * i] = pcoords[3 * i + 1] = pcoords[3 * i + 2] = 0.f;
    if (id < 0) continue;
    const float *P = pos + 3 * i;
    const float r = v3length(v3mk(P[0], P[1], P[2]));
    const int layer = orc_find_height(&cells[id], r);
    if (layers) layers[i] = layer;
    float pv = NAN;
    orc_prism_value(&cells[id], vtx + (size_t)id * 96, geom + (size_t)id * 6, layer, P, &pv, pcoords ? pcoords + 3 * i : NULL);
    values[i] = pv;
  }
}

/* ---- the reference's own integrator ------------------------------------------------ */

/* common/vecmath.h:1927-1938 boxTest */
static inline int box_test(v3 org, v3 dir, float tmin, float tmax, v3 lo, v3 hi, float *t0, float *t1)
{
  v3 tlo = v3div(v3sub(lo, org), dir);
  v3 thi = v3div(v3sub(hi, org), dir);
  v3 tnr = v3min(tlo, thi), tfr = v3max(tlo, thi);
  *t0 = fmaxf(tmin, fmaxf(fmaxf(tnr.x, tnr.y), tnr.z));
  *t1 = fminf(tmax, fminf(fminf(tfr.x, tfr.y), tfr.z));
  return *t0 < *t1;
}

/* icon_rt/deviceCode.cu:239-275 woodcockTrackingAE with :149-186 woodcockTracking (majorant 1) */
void orc_woodcock_frame(const orc_scene *sc, int W, int H, const float *volLo, const float *volHi,
                        uint32_t *fbPointer, float *accumBuffer)
{
  const v3 lo = v3mk(volLo[0], volLo[1], volLo[2]), hi = v3mk(volHi[0], volHi[1], volHi[2]);
  const v3 amb = v3mk(sc->ambientColor[0], sc->ambientColor[1], sc->ambientColor[2]);
  if (sc->accumID == 0) {
    memset(fbPointer, 0, sizeof(uint32_t) * (size_t)W * H);
    memset(accumBuffer, 0, sizeof(float) * 4 * (size_t)W * H);
  }
#pragma omp parallel for schedule(dynamic, 8)
  for (int p = 0; p < W * H; ++p) {
    int x = p % W, y = p / W;
    lcg rnd;
    v3 org, dir;
    orc_scene s2 = *sc;
    s2.jitter = 1;
    pixel_ray(&s2, x, y, W, H, &rnd, &org, &dir);
    float t0, t1;
    if (!box_test(org, dir, 0.f, 1e10f, lo, hi, &t0, &t1)) continue;
    const float majorant = 1.f;
    v3 albedo = v3mk(0, 0, 0);
    float extinction = 0.f;
    float t = t0;
    while (1) {
      if (majorant <= 0.f) break;
      t -= (logf(1.f - lcg_next(&rnd)) / (majorant / sc->unitDistance));
      if (t > t1) break;
      v3 P = v3add(org, v3scale(dir, t));
      float pos[3] = {P.x, P.y, P.z};
      float value = 0.f;
      if (locate1(sc->cells, sc->numCells, pos, &value) < 0) continue;
      v4 sm = post_classify(sc->lut, sc->lutSize, sc->tfLo, sc->tfHi, sc->opacityScale, value);
      float u = lcg_next(&rnd);
      if (sm.w >= u * majorant) {
        albedo = v3mk(sm.x, sm.y, sm.z);
        extinction = sm.w;
        break;
      }
    }
    v3 color = v3scale(v3mul(albedo, amb), sc->ambientRadiance);
    v4 C = {color.x, color.y, color.z, extinction > 0.f ? 1.f : 0.f};
    write_pixel(sc, p, C, fbPointer, accumBuffer);
  }
}

/* ---- macrocell majorants + the reference's default ray-gen ---------------------------------
 *
 * icon_rt/hostCode.cu:205-255 buildGrid_ICON (+ :36-56 float atomicMin/Max, DDA.h:24-32 projectOnGrid,
 * :15-22 linearIndex), hostCode.cu:400-434 computeMaxOpacities(Grid) -- the builders are CUDA kernels inside
 * hostCode.cu, which cannot be compiled here (nvcc + OWL), so they are RESTATED (parity of the builders:
 * unpinned; the ray-gen that consumes their output, woodcockTrackingWithAccel, is the reference's own code
 * in oracle/_ref and pins orc_woodcock_grid_frame bit for bit). */
static inline void project_on_grid(v3 V, const int *dims, v3 lo, v3 hi, int *out)
{
  const v3 V01 = v3div(v3sub(V, lo), v3sub(hi, lo));
  const v3 Vs = v3mul(V01, v3mk((float)dims[0], (float)dims[1], (float)dims[2]));
  const int c[3] = {(int)Vs.x, (int)Vs.y, (int)Vs.z};
  for (int k = 0; k < 3; ++k) { int m = c[k] < dims[k] - 1 ? c[k] : dims[k] - 1; out[k] = 0 > m ? 0 : m; }
}

static inline float get_value(const orc_cell *cell, float h) { return cell->value[orc_find_height(cell, h)]; }

/* ranges: numMC * 2 floats {lower, upper}, initialised to {FLT_MAX, -FLT_MAX} (initGrid, hostCode.cu:62-71) */
void orc_grid_ranges(const orc_cell *cells, int n, const int *dims, const float *wLo, const float *wHi, float *ranges)
{
  const size_t numMC = (size_t)dims[0] * dims[1] * dims[2];
  const v3 lo = v3mk(wLo[0], wLo[1], wLo[2]), hi = v3mk(wHi[0], wHi[1], wHi[2]);
  for (size_t i = 0; i < numMC; ++i) { ranges[2 * i] = 3.402823466e+38f; ranges[2 * i + 1] = -3.402823466e+38f; }
  for (int c = 0; c < n; ++c) {
    const orc_cell *cell = &cells[c];
    for (int i = 0; i < cell->numLayers; ++i) {
      v3 blo = v3mk(INFINITY, INFINITY, INFINITY), bhi = v3mk(-INFINITY, -INFINITY, -INFINITY);
      v3 tv[3];
      for (int k = 0; k < 3; ++k) {
        v3 bv = to_cartesian(cell->height[i], cell->lat[k], cell->lon[k]);
        blo = v3min(blo, bv); bhi = v3max(bhi, bv);
        tv[k] = to_cartesian(cell->height[i + 1], cell->lat[k], cell->lon[k]);
      }
      v3 bary = v3divs(v3add(v3add(tv[0], tv[1]), tv[2]), 3.f);
      float R = cell->height[i + 1];
      float D = R - v3length(bary);
      float off = D / R;
      for (int k = 0; k < 3; ++k) {
        tv[k] = v3add(tv[k], v3scale(tv[k], off));
        blo = v3min(blo, tv[k]); bhi = v3max(bhi, tv[k]);
      }
      float vlo = INFINITY, vhi = -INFINITY;
      float a = get_value(cell, cell->height[i]), b = get_value(cell, cell->height[i + 1]);
      vlo = fminf(vlo, a); vhi = fmaxf(vhi, a);
      vlo = fminf(vlo, b); vhi = fmaxf(vhi, b);
      int l[3], u[3];
      project_on_grid(blo, dims, lo, hi, l);
      project_on_grid(bhi, dims, lo, hi, u);
      for (int z = l[2]; z <= u[2]; ++z)
        for (int y = l[1]; y <= u[1]; ++y)
          for (int x = l[0]; x <= u[0]; ++x) {
            const size_t id = (size_t)z * dims[0] * dims[1] + (size_t)y * dims[0] + x;
            if (vlo < ranges[2 * id]) ranges[2 * id] = vlo;            /* atomicMin: `val < current` */
            if (vhi > ranges[2 * id + 1]) ranges[2 * id + 1] = vhi;    /* atomicMax: `val > current` */
          }
    }
  }
}

/* hostCode.cu:400-434; tf_relRange and tf_opacity are passed but unused there */
void orc_max_opacities(const float *ranges, long numMC, const float *lut, int size, float tfLo, float tfHi, float *maxOp)
{
  for (long m = 0; m < numMC; ++m) {
    float lower = ranges[2 * m], upper = ranges[2 * m + 1];
    if (upper < lower) { maxOp[m] = 0.f; continue; }
    lower -= tfLo; lower /= tfHi - tfLo;
    upper -= tfLo; upper /= tfHi - tfLo;
    int lo = clampi((int)(lower * (size - 1)), 0, size - 1);
    int hi = clampi((int)(upper * (size - 1)) + 1, 0, size - 1);
    float mo = 0.f;
    for (int i = lo; i <= hi; ++i) mo = fmaxf(mo, lut[4 * i + 3]);
    maxOp[m] = mo;
  }
}

/* icon_rt/deviceCode.cu:149-186 woodcockTracking on [tmin,tmax] with the given majorant; returns fminf(t,tmax) */
static float woodcock_tracking(const orc_scene *sc, v3 org, v3 dir, float tmin, float tmax, lcg *rnd, float majorant,
                               v3 *albedo, float *extinction)
{
  float t = tmin;
  while (1) {
    if (majorant <= 0.f) break;
    t -= (logf(1.f - lcg_next(rnd)) / (majorant / sc->unitDistance));
    if (t > tmax) break;
    v3 P = v3add(org, v3scale(dir, t));
    float pos[3] = {P.x, P.y, P.z};
    float value = 0.f;
    if (locate1(sc->cells, sc->numCells, pos, &value) < 0) continue;
    v4 sm = post_classify(sc->lut, sc->lutSize, sc->tfLo, sc->tfHi, sc->opacityScale, value);
    float u = lcg_next(rnd);
    if (sm.w >= u * majorant) {
      *albedo = v3mk(sm.x, sm.y, sm.z);
      *extinction = sm.w;
      break;
    }
  }
  return fminf(t, tmax);
}

/* icon_rt/deviceCode.cu:281-341 woodcockTrackingWithAccel, GRID_ACCEL_MODE: DDA.h:35-136 dda3 over the majorant grid
 * (simultaneous multi-axis step on ties, :111-132), per macrocell woodcockTracking with maxOpacities[leafID];
 * the first collision strictly inside a macrocell's (t0,t1) ends the traversal (:303-311). */
void orc_woodcock_grid_frame(const orc_scene *sc, int W, int H, const float *volLo, const float *volHi,
                             const int *dims, const float *gLo, const float *gHi, const float *maxOp,
                             uint32_t *fbPointer, float *accumBuffer)
{
  const v3 lo = v3mk(volLo[0], volLo[1], volLo[2]), hi = v3mk(volHi[0], volHi[1], volHi[2]);
  const v3 mlo = v3mk(gLo[0], gLo[1], gLo[2]), mhi = v3mk(gHi[0], gHi[1], gHi[2]);
  const v3 amb = v3mk(sc->ambientColor[0], sc->ambientColor[1], sc->ambientColor[2]);
  const v3 fdims = v3mk((float)dims[0], (float)dims[1], (float)dims[2]);
  if (sc->accumID == 0) {
    memset(fbPointer, 0, sizeof(uint32_t) * (size_t)W * H);
    memset(accumBuffer, 0, sizeof(float) * 4 * (size_t)W * H);
  }
#pragma omp parallel for schedule(dynamic, 8)
  for (int p = 0; p < W * H; ++p) {
    int x = p % W, y = p / W;
    lcg rnd;
    v3 org, dir;
    orc_scene s2 = *sc;
    s2.jitter = 1;
    pixel_ray(&s2, x, y, W, H, &rnd, &org, &dir);
    float rt0, rt1;
    if (!box_test(org, dir, 0.f, 1e10f, lo, hi, &rt0, &rt1)) continue;
    v3 color = v3mk(0, 0, 0);
    float alpha = 0.f;
    /* dda3 (works on a shifted copy of the ray; the callback tracks along the caller's ray) */
    const float ray_tmin = rt0;
    const v3 dorg = v3add(org, v3mul(v3mk(ray_tmin, ray_tmin, ray_tmin), dir));
    const float dtmax = rt1 - ray_tmin;
    const v3 rcp = v3div(v3mk(1.f, 1.f, 1.f), dir);
    const v3 l = v3mul(v3sub(mlo, dorg), rcp), h = v3mul(v3sub(mhi, dorg), rcp);
    v3 tnear = v3min(l, h);
    const v3 tfar = v3max(l, h);
    if (dir.x == 0.f) tnear.x = 3.402823466e+38f;
    if (dir.y == 0.f) tnear.y = 3.402823466e+38f;
    if (dir.z == 0.f) tnear.z = 3.402823466e+38f;
    int cell[3];
    project_on_grid(dorg, dims, mlo, mhi, cell);
    const v3 dist = v3max(v3mk(0.f, 0.f, 0.f), v3div(v3sub(tfar, tnear), fdims));
    const int step[3] = {dir.x > 0.f ? 1 : -1, dir.y > 0.f ? 1 : -1, dir.z > 0.f ? 1 : -1};
    const int stop[3] = {dir.x > 0.f ? dims[0] : -1, dir.y > 0.f ? dims[1] : -1, dir.z > 0.f ? dims[2] : -1};
    float tn[3] = {
      dir.x > 0.f ? tnear.x + (float)(cell[0] + 1) * dist.x : tnear.x + (float)(dims[0] - cell[0]) * dist.x,
      dir.y > 0.f ? tnear.y + (float)(cell[1] + 1) * dist.y : tnear.y + (float)(dims[1] - cell[1]) * dist.y,
      dir.z > 0.f ? tnear.z + (float)(cell[2] + 1) * dist.z : tnear.z + (float)(dims[2] - cell[2]) * dist.z};
    const float ds[3] = {dist.x, dist.y, dist.z};
    float t0 = 0.f;
    while (1) {
      const float tmin3 = fminf(fminf(tn[0], tn[1]), tn[2]);       /* reduce_min, vecmath.h:512-514 */
      const float t1 = fminf(tmin3, dtmax);
      const size_t leaf = (size_t)cell[2] * dims[0] * dims[1] + (size_t)cell[1] * dims[0] + cell[0];
      {
        v3 albedo = v3mk(0, 0, 0);
        float extinction = 0.f;
        const float c0 = ray_tmin + t0, c1 = ray_tmin + t1;
        const float t = woodcock_tracking(sc, org, dir, c0, c1, &rnd, maxOp[(int)leaf], &albedo, &extinction);
        if (t > c0 && t < c1) {
          color = v3scale(v3mul(albedo, amb), sc->ambientRadiance);
          alpha = extinction > 0.f ? 1.f : 0.f;
          break;
        }
      }
      int out = 0;
      for (int k = 0; k < 3 && !out; ++k)
        if (tn[k] == tmin3) {
          tn[k] += ds[k];
          cell[k] += step[k];
          if (cell[k] == stop[k]) out = 1;
        }
      if (out) break;
      t0 = t1;
    }
    v4 C = {color.x, color.y, color.z, alpha};
    write_pixel(sc, p, C, fbPointer, accumBuffer);
  }
}

/* ---- SPHERE_ACCEL_MODE: the (rad, lat, lon) shell accel and sdda ---------------------------------
 *
 * icon_rt/ShellAccel.h:55-69 projectToSphericalGrid: ((v - lower) / size) * (dims-1), float -> int by truncation,
 * UNBOUNDED (no clamp).  With the reference's dims (1,1024,1024) (hostCode.cu:654) the radial coordinate is
 * (r-lo)/size*0 = 0. */
static inline void project_spherical(v3 sph, const int *dims, const float *sLo, const float *sHi, int *out)
{
  out[0] = (int)((sph.x - sLo[0]) / (sHi[0] - sLo[0]) * (dims[0] - 1));
  out[1] = (int)((sph.y - sLo[1]) / (sHi[1] - sLo[1]) * (dims[1] - 1));
  out[2] = (int)((sph.z - sLo[2]) / (sHi[2] - sLo[2]) * (dims[2] - 1));
}

/* icon_rt/hostCode.cu:298-334 buildShell_ICON (+ :62-71 initGrid(ShellAccel)): per layer, the cell's three corners at
 * height[i] give loMC (component-wise min), at height[i+1] upMC (max); every shell cell in between takes
 * atomicMin(lower, getValue(height[i])) and atomicMax(upper, getValue(height[i+1])) -- NOT min/max of the two. */
void orc_shell_ranges(const orc_cell *cells, int n, const int *dims, const float *sLo, const float *sHi, float *ranges)
{
  const size_t numMC = (size_t)dims[0] * dims[1] * dims[2];
  for (size_t i = 0; i < numMC; ++i) { ranges[2 * i] = 3.402823466e+38f; ranges[2 * i + 1] = -3.402823466e+38f; }
  for (int c = 0; c < n; ++c) {
    const orc_cell *cell = &cells[c];
    for (int i = 0; i < cell->numLayers; ++i) {
      int lo[3] = {INT32_MAX, INT32_MAX, INT32_MAX}, up[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
      for (int k = 0; k < 3; ++k) {
        int a[3], b[3];
        project_spherical(v3mk(cell->height[i], cell->lat[k], cell->lon[k]), dims, sLo, sHi, a);
        project_spherical(v3mk(cell->height[i + 1], cell->lat[k], cell->lon[k]), dims, sLo, sHi, b);
        for (int q = 0; q < 3; ++q) { if (a[q] < lo[q]) lo[q] = a[q]; if (b[q] > up[q]) up[q] = b[q]; }
      }
      const float vlo = get_value(cell, cell->height[i]), vhi = get_value(cell, cell->height[i + 1]);
      for (int z = lo[2]; z <= up[2]; ++z)
        for (int y = lo[1]; y <= up[1]; ++y)
          for (int x = lo[0]; x <= up[0]; ++x) {
            const size_t id = (size_t)z * dims[0] * dims[1] + (size_t)y * dims[0] + x;     /* linearIndex, DDA.h:15-21 */
            if (vlo < ranges[2 * id]) ranges[2 * id] = vlo;
            if (vhi > ranges[2 * id + 1]) ranges[2 * id + 1] = vhi;
          }
    }
  }
}

static inline v3 to_spherical(v3 c)                 /* ICONGrid.h:36-42 */
{
  const float r = v3length(c);
  return v3mk(r, asinf(c.z / r), atan2f(c.y, c.x));
}

static inline int wrap_coord(int c, int d) { while (c < 0) c += d; while (c >= d) c -= d; return c; }   /* ShellAccel.h:72-81 */

/* icon_rt/deviceCode.cu:281-341 woodcockTrackingWithAccel, SPHERE_ACCEL_MODE: ShellAccel.h:83-199 sdda.
 * What sdda does as written: the lat/lon cell planes are makePlane(0, toCartesian(0,..), toCartesian(0,..)) -- three
 * points at the ORIGIN (radius 0), so the plane is (0,0,0,0) and evalPlane gives 0 for every position: tnext.y =
 * tnext.z = 0.  Each radial segment therefore yields one callback over the whole segment [lower, upper] in the cell of
 * its entry point, then the loop walks diagonally through (lat,lon) cells with t = t1 = 0 -- zero-length callbacks,
 * each costing one random number when its majorant is positive -- until the lat or lon index reaches the exit
 * point's.  Restated literally (every comparison and update of the loop), not simplified. */
void orc_woodcock_sphere_frame(const orc_scene *sc, int W, int H, const float *volLo, const float *volHi,
                               const int *dims, const float *sLo, const float *sHi, const float *maxOp,
                               uint32_t *fbPointer, float *accumBuffer)
{
  const v3 lo = v3mk(volLo[0], volLo[1], volLo[2]), hi = v3mk(volHi[0], volHi[1], volHi[2]);
  const v3 amb = v3mk(sc->ambientColor[0], sc->ambientColor[1], sc->ambientColor[2]);
  if (sc->accumID == 0) {
    memset(fbPointer, 0, sizeof(uint32_t) * (size_t)W * H);
    memset(accumBuffer, 0, sizeof(float) * 4 * (size_t)W * H);
  }
#pragma omp parallel for schedule(dynamic, 8)
  for (int p = 0; p < W * H; ++p) {
    int x = p % W, y = p / W;
    lcg rnd;
    v3 org, dir;
    orc_scene s2 = *sc;
    s2.jitter = 1;
    pixel_ray(&s2, x, y, W, H, &rnd, &org, &dir);
    float rt0, rt1;
    if (!box_test(org, dir, 0.f, 1e10f, lo, hi, &rt0, &rt1)) continue;
    v3 color = v3mk(0, 0, 0);
    float alpha = 0.f;
    int done = 0;
    const float sceneEPS = sLo[0] * 1e-6f;
    float t1s = 0.f, t2s = 0.f, t3s = 0.f, t4s = 0.f;
    const int s1 = intersect_sphere(org, dir, sHi[0], &t1s, &t4s);
    const int s2i = intersect_sphere(org, dir, sLo[0], &t2s, &t3s);
    float rl[2] = {INFINITY, INFINITY}, ru[2] = {-INFINITY, -INFINITY};
    if ((s1 || s2i) && !(t4s < rt0)) {
      if (s1 && !s2i) { rl[0] = t1s; ru[0] = t4s; }
      else if (rt0 < t2s) { rl[0] = t1s; ru[0] = t2s; rl[1] = t3s; ru[1] = t4s; }
      else { rl[0] = t3s; ru[0] = t4s; }
    }
    for (int i = 0; i < 2 && !done; ++i) {
      if (ru[i] < rl[i]) break;                                    /* box1f::empty() */
      const v3 P1 = v3add(org, v3scale(dir, rl[i] + sceneEPS)), P2 = v3add(org, v3scale(dir, ru[i] - sceneEPS));
      const v3 SP1 = to_spherical(P1), SP2 = to_spherical(P2);
      int cell[3], stop[3];
      project_spherical(SP1, dims, sLo, sHi, cell);
      const int step[3] = {SP1.x < SP2.x ? 1 : -1, SP1.y < SP2.y ? 1 : -1, SP1.z < SP2.z ? 1 : -1};
      project_spherical(SP2, dims, sLo, sHi, stop);
      for (int k = 0; k < 3; ++k) stop[k] += step[k];
      float tn[3] = {ru[i], 0.f, 0.f};                             /* evalPlane of the all-zero planes */
      float t = rl[i];
      while (1) {
        float t1 = 3.402823466e+38f;
        if (tn[0] < t1 && tn[0] >= t) t1 = tn[0];
        if (tn[1] < t1 && tn[1] >= t) t1 = tn[1];
        if (tn[2] < t1 && tn[2] >= t) t1 = tn[2];
        const size_t leaf = (size_t)wrap_coord(cell[2], dims[2]) * dims[0] * dims[1] + (size_t)wrap_coord(cell[1], dims[1]) * dims[0]
                          + wrap_coord(cell[0], dims[0]);
        {
          v3 albedo = v3mk(0, 0, 0);
          float extinction = 0.f;
          const float tt = woodcock_tracking(sc, org, dir, t, t1, &rnd, maxOp[(int)leaf], &albedo, &extinction);
          if (tt > t && tt < t1) {
            color = v3scale(v3mul(albedo, amb), sc->ambientRadiance);
            alpha = extinction > 0.f ? 1.f : 0.f;
            done = 1;
            break;
          }
        }
        const float tc = fminf(fminf(tn[0], tn[1]), tn[2]);
        if (tn[0] == tc) { cell[0] += step[0]; if (cell[0] == stop[0]) break; }
        if (tn[1] == tc) { cell[1] += step[1]; if (cell[1] == stop[1]) break; tn[1] = 0.f; }
        if (tn[2] == tc) { cell[2] += step[2]; if (cell[2] == stop[2]) break; tn[2] = 0.f; }
        t = tc;
      }
    }
    v4 C = {color.x, color.y, color.z, alpha};
    write_pixel(sc, p, C, fbPointer, accumBuffer);
  }
}
