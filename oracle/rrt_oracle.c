/* rrt_oracle.c — TEST INFRASTRUCTURE: CPU restatement, in plain C, of the reference's hot path
 * (per-pixel Monte Carlo radiance estimation with Schwarzschild ray bending).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library; it is the
 * checker, never the product.  Every function cites the reference file:line it restates (paths relative
 * to /root/reference/pathtracer).  All geometry is IEEE double evaluated in the reference's operation
 * order (compile with -ffp-contract=off: the reference's x86-64 build has no FMA); radiance is float.
 *
 * PARITY PINNED: tests/golden/ holds outputs of the reference's own objects (oracle/ref_probe.cpp driving
 * BVHAccel::intersect, BlackHole::next_micro_ray, Camera::generate_ray, BSDF::sample_f, SceneLight::sample_L
 * and PathTracer::raytrace_pixel with rand() fed from the Philox stream defined below); tests/test_oracle_*.py
 * check this file against them.
 *
 * Random numbers.  The reference calls rand() sequentially (src/random_util.h:11-13).  Here draw n of path
 * (pixel, sample) is word (n & 3) of Philox4x32-10(key = seed, counter = (pixel, sample, n >> 2, 0)),
 * turned into the reference's uniform as (word >> 1) / 2147483647.0  (RAND_MAX = 2^31 - 1, range [0,1]).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "../include/rrt_b200.h"

#define ORC_PI 3.14159265358979323 /* CGL/include/CGL/misc.h:11 (PI); M_PI has the same double value */
#define ORC_EPS_D 0.00000000001    /* CGL/include/CGL/misc.h:12 */

typedef struct { double x, y, z; } v3;
typedef struct { float r, g, b; } spec;

/* CGL/include/CGL/vector3D.h:58-157 — note operator/ and normalize multiply by the reciprocal */
static inline v3 V(double x, double y, double z) { v3 r = {x, y, z}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }
static inline v3 vmul(v3 a, double c) { return V(a.x * c, a.y * c, a.z * c); }   /* v * c */
static inline v3 smul(double c, v3 a) { return V(c * a.x, c * a.y, c * a.z); }   /* c * v */
static inline v3 vdiv(v3 a, double c) { double rc = 1.0 / c; return V(rc * a.x, rc * a.y, rc * a.z); }
static inline double vdot(v3 u, v3 v) { return u.x * v.x + u.y * v.y + u.z * v.z; }
static inline v3 vcross(v3 u, v3 v) { return V(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x); }
static inline double vnorm2(v3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
static inline double vnorm(v3 a) { return sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }
static inline v3 vunit(v3 a) { double rn = 1. / sqrt(a.x * a.x + a.y * a.y + a.z * a.z); return V(rn * a.x, rn * a.y, rn * a.z); }
static inline v3 vnormalize(v3 a) { double rc = 1. / vnorm(a); return V(a.x * rc, a.y * rc, a.z * rc); } /* (*this) *= (1./norm()) */

/* ------------------------------------------------------------------ Philox4x32-10 (Salmon et al., SC'11) */
static void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}

typedef struct { uint64_t seed; uint32_t pixel, sample, n; uint32_t w[4]; const double* fixed; } rng_t;

/* random_uniform(): ((double)rand()) / RAND_MAX  (src/random_util.h:11-13) */
static double rng_uniform(rng_t* g) {
  if (g->fixed) return g->fixed[g->n++];  /* KAT mode: caller supplies the draws */
  if ((g->n & 3) == 0) {
    g->w[0] = g->pixel; g->w[1] = g->sample; g->w[2] = g->n >> 2; g->w[3] = 0;
    philox4x32_10(g->w, (uint32_t)g->seed, (uint32_t)(g->seed >> 32));
  }
  uint32_t word = g->w[g->n & 3];
  g->n++;
  return (double)(word >> 1) / 2147483647.0;
}
/* UniformGridSampler2D::get_sample (src/sampler.cpp:7-11): Vector2D(random_uniform(), random_uniform());
 * g++ 13 evaluates the constructor arguments right to left, so .y takes the first draw and .x the second
 * (verified on the reference objects: `oracle/_ref/ref_probe order` prints "x=second"). */
static void rng_grid2(rng_t* g, double* x, double* y) { *y = rng_uniform(g); *x = rng_uniform(g); }

/* ------------------------------------------------------------------ context */
typedef struct { v3 mn, mx; } box_t;          /* BBox (src/bbox.h:19-24), extent recomputed where needed */
typedef struct { box_t bb; int l, r; int start, count; } node_t; /* BVHNode (src/bvh.h:24-41); leaf: l == -1 */

typedef struct orc_ctx {
  uint32_t n_prims, n_tris, n_spheres, n_mats, n_lights;
  uint8_t* prim_kind; uint32_t* prim_index;
  double* tri_pos; double* tri_nrm; int32_t* tri_mat;
  double* sph; int32_t* sph_mat;
  rrt_material* mats; rrt_light* lights;
  rrt_camera cam;
  v3 bh_c; double bh_r, bh_r2, bh_dtheta; int mode;
  int W, H;
  int cell_x, cell_y, cell_w, cell_h;
  /* reference BVH */
  node_t* nodes; int n_nodes, cap_nodes, n_leaves, depth;
  uint32_t* order;  /* primitives in depth-first leaf order */
  uint64_t cnt_rays, cnt_steps, cnt_nodes, cnt_tris, cnt_sph; /* per-call counters (not thread safe: see orc_render) */
  /* environment light (src/static_scene/environment_light.{h,cpp}): appended after the scene's lights (src/pathtracer.cpp:107-109) */
  int env_w, env_h;
  float* env;                         /* [h][w][3], row 0 = top */
  double *env_pdf, *env_cond, *env_marg;
} orc_ctx;

static void* dup_mem(const void* p, size_t bytes) { void* q = malloc(bytes ? bytes : 1); if (bytes) memcpy(q, p, bytes); return q; }

/* Primitive::get_bbox — Triangle (src/static_scene/triangle.cpp:11-19), Sphere (src/static_scene/sphere.h:30-32) */
static box_t prim_bbox(const orc_ctx* c, uint32_t p) {
  box_t b;
  if (c->prim_kind[p] == 0) {
    const double* q = c->tri_pos + 9 * (size_t)c->prim_index[p];
    b.mn = b.mx = V(q[0], q[1], q[2]);
    for (int k = 1; k < 3; ++k) {
      b.mn.x = fmin(b.mn.x, q[3 * k]); b.mn.y = fmin(b.mn.y, q[3 * k + 1]); b.mn.z = fmin(b.mn.z, q[3 * k + 2]);
      b.mx.x = fmax(b.mx.x, q[3 * k]); b.mx.y = fmax(b.mx.y, q[3 * k + 1]); b.mx.z = fmax(b.mx.z, q[3 * k + 2]);
    }
  } else {
    const double* s = c->sph + 4 * (size_t)c->prim_index[p];
    b.mn = V(s[0] - s[3], s[1] - s[3], s[2] - s[3]);
    b.mx = V(s[0] + s[3], s[1] + s[3], s[2] + s[3]);
  }
  return b;
}
static inline v3 box_centroid(box_t b) { return vdiv(vadd(b.mn, b.mx), 2); } /* src/bbox.h:98-100 */

/* BVHAccel::construct_bvh (src/bvh.cpp:49-96): bbox of all prims; leaf if <= max_leaf_size (4); else split at
 * the bbox-centroid midpoint of the strictly largest extent axis (x, then y, else z), prims whose own bbox
 * centroid is < midpoint go left; if one side is empty, halve by index. */
static int build_rec(orc_ctx* c, uint32_t* prims, uint32_t n, uint32_t* out_pos, int depth) {
  if (c->n_nodes == c->cap_nodes) { c->cap_nodes = c->cap_nodes ? 2 * c->cap_nodes : 64; c->nodes = realloc(c->nodes, sizeof(node_t) * c->cap_nodes); }
  int me = c->n_nodes++;
  box_t bb; bb.mn = V(INFINITY, INFINITY, INFINITY); bb.mx = V(-INFINITY, -INFINITY, -INFINITY);
  for (uint32_t i = 0; i < n; ++i) {
    box_t pb = prim_bbox(c, prims[i]);
    bb.mn.x = fmin(bb.mn.x, pb.mn.x); bb.mn.y = fmin(bb.mn.y, pb.mn.y); bb.mn.z = fmin(bb.mn.z, pb.mn.z);
    bb.mx.x = fmax(bb.mx.x, pb.mx.x); bb.mx.y = fmax(bb.mx.y, pb.mx.y); bb.mx.z = fmax(bb.mx.z, pb.mx.z);
  }
  c->nodes[me].bb = bb; c->nodes[me].l = c->nodes[me].r = -1; c->nodes[me].start = c->nodes[me].count = 0;
  if (depth > c->depth) c->depth = depth;
  if (n <= 4) {
    c->nodes[me].start = (int)*out_pos; c->nodes[me].count = (int)n;
    for (uint32_t i = 0; i < n; ++i) c->order[(*out_pos)++] = prims[i];
    c->n_leaves++;
    return me;
  }
  v3 ext = vsub(bb.mx, bb.mn), cen = box_centroid(bb);
  int axis = (ext.x > ext.y && ext.x > ext.z) ? 0 : (ext.y > ext.x && ext.y > ext.z) ? 1 : 2;
  double mid = axis == 0 ? cen.x : axis == 1 ? cen.y : cen.z;
  uint32_t* L = malloc(sizeof(uint32_t) * n); uint32_t* R = malloc(sizeof(uint32_t) * n); uint32_t nl = 0, nr = 0;
  for (uint32_t i = 0; i < n; ++i) {
    v3 pc = box_centroid(prim_bbox(c, prims[i]));
    double v = axis == 0 ? pc.x : axis == 1 ? pc.y : pc.z;
    if (v < mid) L[nl++] = prims[i]; else R[nr++] = prims[i];
  }
  if (nl == 0 || nr == 0) {
    nl = n / 2; nr = n - nl;
    memcpy(L, prims, sizeof(uint32_t) * nl); memcpy(R, prims + nl, sizeof(uint32_t) * nr);
  }
  int l = build_rec(c, L, nl, out_pos, depth + 1);
  int r = build_rec(c, R, nr, out_pos, depth + 1);
  c->nodes[me].l = l; c->nodes[me].r = r;
  free(L); free(R);
  return me;
}

orc_ctx* orc_create(const rrt_scene* s, const rrt_camera* cam) {
  orc_ctx* c = calloc(1, sizeof *c);
  c->n_prims = s->n_prims; c->n_tris = s->n_tris; c->n_spheres = s->n_spheres; c->n_mats = s->n_mats; c->n_lights = s->n_lights;
  c->prim_kind = dup_mem(s->prim_kind, s->n_prims); c->prim_index = dup_mem(s->prim_index, 4 * (size_t)s->n_prims);
  c->tri_pos = dup_mem(s->tri_pos, 72 * (size_t)s->n_tris); c->tri_nrm = dup_mem(s->tri_nrm, 72 * (size_t)s->n_tris);
  c->tri_mat = dup_mem(s->tri_mat, 4 * (size_t)s->n_tris);
  c->sph = dup_mem(s->sph, 32 * (size_t)s->n_spheres); c->sph_mat = dup_mem(s->sph_mat, 4 * (size_t)s->n_spheres);
  c->mats = dup_mem(s->mats, sizeof(rrt_material) * (size_t)s->n_mats);
  c->lights = dup_mem(s->lights, sizeof(rrt_light) * (size_t)s->n_lights);
  if (cam) c->cam = *cam;
  /* global_black_hole(nullptr, Vector3D(0, 1, 0), 0.1, 0.1)  (src/static_scene/blackhole.cpp:5) */
  c->bh_c = V(0, 1, 0); c->bh_r = 0.1; c->bh_r2 = 0.1 * 0.1; c->bh_dtheta = 0.1; c->mode = RRT_MODE_LITERAL;
  c->order = malloc(sizeof(uint32_t) * (s->n_prims ? s->n_prims : 1));
  uint32_t* all = malloc(sizeof(uint32_t) * (s->n_prims ? s->n_prims : 1));
  for (uint32_t i = 0; i < s->n_prims; ++i) all[i] = i;
  uint32_t pos = 0;
  build_rec(c, all, s->n_prims, &pos, 0);
  free(all);
  return c;
}

void orc_destroy(orc_ctx* c) {
  if (!c) return;
  free(c->prim_kind); free(c->prim_index); free(c->tri_pos); free(c->tri_nrm); free(c->tri_mat); free(c->sph);
  free(c->sph_mat); free(c->mats); free(c->lights); free(c->nodes); free(c->order);
  free(c->env); free(c->env_pdf); free(c->env_cond); free(c->env_marg); free(c);
}

void orc_set_camera(orc_ctx* c, const rrt_camera* cam) { c->cam = *cam; }
/* -B x y z r dtheta (src/main.cpp:139-145): r2 = r*r */
void orc_set_blackhole(orc_ctx* c, const double centre[3], double r, double dtheta, int mode) {
  c->bh_c = V(centre[0], centre[1], centre[2]); c->bh_r = r; c->bh_r2 = r * r; c->bh_dtheta = dtheta; c->mode = mode;
}
void orc_set_frame(orc_ctx* c, int w, int h) { c->W = w; c->H = h; c->cell_w = 0; }
void orc_set_cell(orc_ctx* c, int x, int y, int w, int h) { c->cell_x = x; c->cell_y = y; c->cell_w = w; c->cell_h = h; }

/* loop bound of BVHAccel::intersect: for (int j = 0; j * delta_theta < 2 * M_PI; ++j)  (src/bvh.cpp:105) */
int orc_chords_per_ray(const orc_ctx* c) {
  int j = 0;
  while (j * c->bh_dtheta < 2 * M_PI && j < (1 << 20)) ++j;
  return j;
}

void orc_bvh_info(const orc_ctx* c, int32_t* n_nodes, int32_t* n_leaves, int32_t* depth, uint32_t* dfs_rank) {
  if (n_nodes) *n_nodes = c->n_nodes;
  if (n_leaves) *n_leaves = c->n_leaves;
  if (depth) *depth = c->depth;
  if (dfs_rank) for (uint32_t i = 0; i < c->n_prims; ++i) dfs_rank[c->order[i]] = i;
}

/* ------------------------------------------------------------------ geometry */
typedef struct { v3 o, d; double min_t, max_t; } ray_t;               /* Ray (src/ray.h:20-60); inv_d/sign unused */
typedef struct { int prim; v3 hit_p, w_out, n; } isect_t;             /* Intersection (src/intersection.h:20-35) */
typedef struct { uint64_t rays, steps, nodes, tris, sph; } cnt_t;

/* BlackHole::f  (src/static_scene/blackhole.cpp:13-15) */
static inline double bh_f(const orc_ctx* c, double u) { return -u + 3.0 * c->bh_r * u * u / 2.0; }

/* BlackHole::next_micro_ray  (src/static_scene/blackhole.cpp:17-40) */
static ray_t next_micro_ray(const orc_ctx* c, const ray_t* ray) {
  const double h = c->bh_dtheta;
  ray_t ret;
  ret.o = vadd(ray->o, vmul(ray->d, ray->max_t)); ret.min_t = 0.0; ret.max_t = 0.0;
  v3 x_axis = vsub(ret.o, c->bh_c);
  double d = vnorm(x_axis);
  x_axis = vnormalize(x_axis);
  double u = 1 / d;
  double dx = vdot(ray->d, x_axis);
  v3 y_axis = vsub(ray->d, smul(dx, x_axis));
  double dy = vnorm(y_axis);
  y_axis = vnormalize(y_axis);
  double up = -u * dx / dy;
  double f1 = bh_f(c, u);
  double f2 = bh_f(c, u + up * h / 2.0);
  double f3 = bh_f(c, u + up * h / 2.0 + f1 * h * h / 4.0);
  /* f4 (blackhole.cpp:31) is computed by the reference and never used */
  u += up * h + (f1 + f2 + f3) * h * h / 6.0;
  d = 1 / u;
  double next_x = d * cos(h);
  double next_y = d * sin(h);
  ret.d = vsub(vadd(vadd(c->bh_c, smul(next_x, x_axis)), smul(next_y, y_axis)), ret.o);
  ret.max_t = vnorm(ret.d);
  ret.d = vnormalize(ret.d);
  return ret;
}

/* Sphere::test  (src/static_scene/sphere.cpp:10-20) */
static int sphere_test(v3 so, double r2, const ray_t* r, double* t1, double* t2) {
  v3 tmp = vsub(r->o, so);
  double b = 2 * vdot(tmp, r->d), c = vnorm2(tmp) - r2, d = b * b - 4 * c;
  if (d < 0) return 0;
  *t1 = (-b - sqrt(d)) / 2;
  *t2 = (-b + sqrt(d)) / 2;
  return 1;
}
/* Sphere::intersect  (src/static_scene/sphere.cpp:26-53); shrinks r->max_t on a hit; normal = unit(p - o) (sphere.h:73-75) */
static int sphere_intersect(v3 so, double r2, ray_t* r, isect_t* i, int prim) {
  double t1, t2, t;
  if (!sphere_test(so, r2, r, &t1, &t2)) return 0;
  if (r->min_t <= t1 && t1 <= r->max_t) t = t1;
  else if (r->min_t <= t2 && t2 <= r->max_t) t = t2;
  else return 0;
  r->max_t = t;
  if (i) {
    i->hit_p = vadd(r->o, vmul(r->d, t));
    i->w_out = vneg(r->d);
    i->n = vunit(vsub(vadd(r->o, vmul(r->d, t)), so));
    i->prim = prim;
  }
  return 1;
}

/* Triangle::intersect  (src/static_scene/triangle.cpp:25-55): Moller-Trumbore, res /= det multiplies by 1/det,
 * inclusive bounds, no culling, normal = barycentric blend, NOT normalised. */
static int triangle_intersect(const double* P, const double* N, ray_t* r, isect_t* isect, int prim) {
  v3 p0 = V(P[0], P[1], P[2]), p1 = V(P[3], P[4], P[5]), p2 = V(P[6], P[7], P[8]);
  v3 e1 = vsub(p1, p0), e2 = vsub(p2, p0), s = vsub(r->o, p0);
  v3 s1 = vcross(r->d, e2), s2 = vcross(s, e1);
  v3 res = V(vdot(s2, e2), vdot(s1, s), vdot(s2, r->d));
  double inv = 1. / vdot(s1, e1);
  res = vmul(res, inv);
  double t = res.x, b1 = res.y, b2 = res.z, b0 = 1 - b1 - b2;
  if (!(r->min_t <= t && t <= r->max_t && b0 >= 0 && b1 >= 0 && b2 >= 0)) return 0;
  r->max_t = t;
  if (isect) {
    isect->hit_p = vadd(r->o, vmul(r->d, t));
    isect->w_out = vneg(r->d);
    v3 n0 = V(N[0], N[1], N[2]), n1 = V(N[3], N[4], N[5]), n2 = V(N[6], N[7], N[8]);
    isect->n = vadd(vadd(smul(b0, n0), smul(b1, n1)), smul(b2, n2));
    isect->prim = prim;
  }
  return 1;
}

/* BBox::intersect  (src/bbox.cpp:10-25): six true divisions; std::min/max argument-order semantics */
static inline double cmin(double a, double b) { return (b < a) ? b : a; } /* std::min(a,b) */
static inline double cmax(double a, double b) { return (a < b) ? b : a; } /* std::max(a,b) */
static int box_intersect(const box_t* bb, const ray_t* r) {
  double tx0 = (bb->mn.x - r->o.x) / r->d.x, tx1 = (bb->mx.x - r->o.x) / r->d.x;
  double ty0 = (bb->mn.y - r->o.y) / r->d.y, ty1 = (bb->mx.y - r->o.y) / r->d.y;
  double tz0 = (bb->mn.z - r->o.z) / r->d.z, tz1 = (bb->mx.z - r->o.z) / r->d.z;
  double tmin = cmax(cmax(cmin(tx0, tx1), cmin(ty0, ty1)), cmin(tz0, tz1));
  double tmax = cmin(cmin(cmax(tx0, tx1), cmax(ty0, ty1)), cmax(tz0, tz1));
  return tmin <= tmax && tmin <= r->max_t && tmax >= r->min_t;
}

/* BVHAccel::intersect_micro  (src/bvh.cpp:115-138): recursive, all leaf prims tested, both children always visited */
static int intersect_micro(const orc_ctx* c, ray_t* r, isect_t* i, int node, cnt_t* cnt) {
  const node_t* nd = &c->nodes[node];
  cnt->nodes++;
  if (!box_intersect(&nd->bb, r)) return 0;
  int hit = 0;
  if (nd->l < 0) {
    for (int k = 0; k < nd->count; ++k) {
      uint32_t p = c->order[nd->start + k];
      if (c->prim_kind[p] == 0) {
        cnt->tris++;
        size_t ti = c->prim_index[p];
        if (triangle_intersect(c->tri_pos + 9 * ti, c->tri_nrm + 9 * ti, r, i, (int)p)) hit = 1;
      } else {
        cnt->sph++;
        const double* s = c->sph + 4 * (size_t)c->prim_index[p];
        if (sphere_intersect(V(s[0], s[1], s[2]), s[3] * s[3], r, i, (int)p)) hit = 1;
      }
    }
  } else {
    if (intersect_micro(c, r, i, nd->l, cnt)) hit = 1;
    if (intersect_micro(c, r, i, nd->r, cnt)) hit = 1;
  }
  return hit;
}

/* BVHAccel::intersect(ray, i, node)  (src/bvh.cpp:103-113).  Literal mode: the caller's min_t/max_t are dropped
 * (micro_ray starts with max_t = 0), every chord is tested against the horizon first (capture => false), then
 * against the BVH.  Flat mode = intersect_micro on the caller's ray (the reference's own flat traversal).
 * status: 0 miss, 1 hit, 2 captured; *chord / *steps / *t_out as in rrt_hit. */
static int bvh_intersect(const orc_ctx* c, const ray_t* ray, isect_t* i, int* chord, int* steps, double* t_out, cnt_t* cnt) {
  cnt->rays++;
  if (c->mode == RRT_MODE_FLAT) {
    ray_t r = *ray;
    cnt->steps++;
    if (chord) *chord = 0;
    if (steps) *steps = 1;
    int hit = c->n_nodes ? intersect_micro(c, &r, i, 0, cnt) : 0;
    if (t_out) *t_out = r.max_t;
    return hit ? 1 : 0;
  }
  ray_t micro; micro.o = ray->o; micro.d = ray->d; micro.min_t = 0.0; micro.max_t = 0.0;
  int j;
  for (j = 0; j * c->bh_dtheta < 2 * M_PI; ++j) {
    micro = next_micro_ray(c, &micro);
    cnt->steps++;
    ray_t probe = micro;
    if (sphere_intersect(c->bh_c, c->bh_r2, &probe, NULL, -1)) {
      if (chord) *chord = j;
      if (steps) *steps = j + 1;
      return 2;
    }
    ray_t seg = micro;
    if (c->n_nodes && intersect_micro(c, &seg, i, 0, cnt)) {
      if (chord) *chord = j;
      if (steps) *steps = j + 1;
      if (t_out) *t_out = seg.max_t;
      return 1;
    }
  }
  if (chord) *chord = j - 1;
  if (steps) *steps = j;
  return 0;
}

/* ------------------------------------------------------------------ camera, shading frame */
static inline double radians_(double deg) { return deg * (ORC_PI / 180); } /* CGL/include/CGL/misc.h:50-53 */
static inline double lerp_(double a, double b, double t) { return (1 - t) * a + t * b; } /* src/part1_code.cpp:177-180 */
/* Matrix3x3 * Vector3D = x0*col0 + x1*col1 + x2*col2  (CGL/src/matrix3x3.cpp:124-128) */
static inline v3 mat_mul(const v3 col[3], v3 x) { return vadd(vadd(smul(x.x, col[0]), smul(x.y, col[1])), smul(x.z, col[2])); }
/* (M.T()) * x : component i = x0*col_i.x + x1*col_i.y + x2*col_i.z  (matrix3x3.cpp:130-141 then :124-128) */
static inline v3 mat_tmul(const v3 col[3], v3 x) {
  return V(x.x * col[0].x + x.y * col[0].y + x.z * col[0].z, x.x * col[1].x + x.y * col[1].y + x.z * col[1].z,
           x.x * col[2].x + x.y * col[2].y + x.z * col[2].z);
}

/* Camera::generate_ray  (src/part1_code.cpp:182-187): min_t = nClip, max_t = fClip */
static ray_t generate_ray(const rrt_camera* cam, double x, double y) {
  double blx = -tan(radians_(cam->hFov) / 2), bly = -tan(radians_(cam->vFov) / 2);
  v3 col[3] = {V(cam->c2w[0], cam->c2w[1], cam->c2w[2]), V(cam->c2w[3], cam->c2w[4], cam->c2w[5]), V(cam->c2w[6], cam->c2w[7], cam->c2w[8])};
  ray_t r;
  r.o = V(cam->pos[0], cam->pos[1], cam->pos[2]);
  r.d = vunit(mat_mul(col, V(lerp_(blx, -blx, x), lerp_(bly, -bly, y), -1)));
  r.min_t = cam->nClip; r.max_t = cam->fClip;
  return r;
}
/* Camera::generate_ray_for_thin_lens  (src/camera.cpp:176-184) */
static ray_t generate_ray_thin_lens(const rrt_camera* cam, double x, double y, double rndR, double rndTheta) {
  double blx = -tan(radians_(cam->hFov) / 2), bly = -tan(radians_(cam->vFov) / 2);
  v3 col[3] = {V(cam->c2w[0], cam->c2w[1], cam->c2w[2]), V(cam->c2w[3], cam->c2w[4], cam->c2w[5]), V(cam->c2w[6], cam->c2w[7], cam->c2w[8])};
  v3 pinHole = V(lerp_(blx, -blx, x), lerp_(bly, -bly, y), -1);
  double lr = cam->lensRadius * sqrt(rndR);
  v3 pLens = V(lr * cos(rndTheta), lr * sin(rndTheta), 0);
  ray_t r;
  r.o = vadd(V(cam->pos[0], cam->pos[1], cam->pos[2]), mat_mul(col, pLens));
  r.d = vunit(mat_mul(col, vsub(vmul(pinHole, cam->focalDistance), pLens)));
  r.min_t = cam->nClip; r.max_t = cam->fClip;
  return r;
}

/* make_coord_space  (src/bsdf.cpp:13-29): columns x,y,z of o2w */
static void make_coord_space(v3 o2w[3], v3 n) {
  v3 z = n, h = z;
  if (fabs(h.x) <= fabs(h.y) && fabs(h.x) <= fabs(h.z)) h.x = 1.0;
  else if (fabs(h.y) <= fabs(h.x) && fabs(h.y) <= fabs(h.z)) h.y = 1.0;
  else h.z = 1.0;
  z = vnormalize(z);
  v3 y = vnormalize(vcross(h, z));
  v3 x = vnormalize(vcross(z, y));
  o2w[0] = x; o2w[1] = y; o2w[2] = z;
}

/* ------------------------------------------------------------------ Spectrum (float) helpers, CGL/include/CGL/spectrum.h */
static inline spec S(float r, float g, float b) { spec s = {r, g, b}; return s; }
static inline spec s_add(spec a, spec b) { return S(a.r + b.r, a.g + b.g, a.b + b.b); }
static inline spec s_sub(spec a, spec b) { return S(a.r - b.r, a.g - b.g, a.b - b.b); }
static inline spec s_mul(spec a, spec b) { return S(a.r * b.r, a.g * b.g, a.b * b.b); }
static inline spec s_div(spec a, spec b) { return S(a.r / b.r, a.g / b.g, a.b / b.b); }
static inline spec s_addf(spec a, float s) { return S(a.r + s, a.g + s, a.b + s); }
static inline spec s_mulf(spec a, float s) { return S(a.r * s, a.g * s, a.b * s); }
static inline spec s_divf(spec a, float s) { return S(a.r / s, a.g / s, a.b / s); }
static inline float s_illum(spec a) { return 0.2126f * a.r + 0.7152f * a.g + 0.0722f * a.b; }
static inline spec s3(const float* p) { return S(p[0], p[1], p[2]); }

/* ------------------------------------------------------------------ BSDFs */
static inline double clampd(double n, double lo, double hi) { return cmax(lo, cmin(n, hi)); } /* src/bsdf.h:20-22 */
static inline int mat_is_delta(const rrt_material* m) { return m->type == RRT_MAT_MIRROR || m->type == RRT_MAT_GLASS || m->type == RRT_MAT_REFRACTION; }
static inline spec mat_emission(const rrt_material* m) { return m->type == RRT_MAT_EMISSION ? s3(m->a) : S(0, 0, 0); }

/* MicrofacetBSDF helpers  (src/bsdf.h:165-173, src/bsdf.cpp:45-66) */
static double mf_theta(v3 w) { return acos(clampd(w.z, -1.0 + 1e-5, 1.0 - 1e-5)); }
static double mf_lambda(const rrt_material* m, v3 w) {
  double theta = mf_theta(w);
  double a = 1.0 / (m->alpha * tan(theta));
  return 0.5 * (erf(a) - 1.0 + exp(-a * a) / (a * ORC_PI));
}
static double mf_G(const rrt_material* m, v3 wo, v3 wi) { return 1.0 / (1.0 + mf_lambda(m, wi) + mf_lambda(m, wo)); }
static double mf_D(const rrt_material* m, v3 h) {
  double theta_h = mf_theta(h), tan_theta_h = tan(theta_h), cos_theta_h = h.z, cos_theta_h2 = cos_theta_h * cos_theta_h,
         alpha2 = m->alpha * m->alpha; /* float product, as in the reference */
  return exp(-tan_theta_h * tan_theta_h / alpha2) / (M_PI * alpha2 * cos_theta_h2 * cos_theta_h2);
}
static spec mf_F(const rrt_material* m, v3 wi) {
  spec eta = s3(m->a), k = s3(m->b);
  spec eta2pk2 = s_add(s_mul(eta, eta), s_mul(k, k));
  double cos_theta_i = wi.z, cos_theta_i2 = cos_theta_i * cos_theta_i;
  spec tetacos = s_mulf(s_mulf(eta, 2.f), (float)cos_theta_i);
  spec R_s = s_div(s_addf(s_sub(eta2pk2, tetacos), (float)cos_theta_i2), s_addf(s_add(eta2pk2, tetacos), (float)cos_theta_i2));
  spec R_p = s_div(s_addf(s_sub(s_mulf(eta2pk2, (float)cos_theta_i2), tetacos), 1.f),
                   s_addf(s_add(s_mulf(eta2pk2, (float)cos_theta_i2), tetacos), 1.f));
  return s_divf(s_add(R_s, R_p), 2.f);
}

/* BSDF::f  (src/part1_code.cpp:167-169 diffuse; src/bsdf.cpp:68-72 microfacet; zero for all others) */
static spec bsdf_f(const rrt_material* m, v3 wo, v3 wi) {
  if (m->type == RRT_MAT_DIFFUSE) return s_divf(s3(m->a), (float)M_PI);
  if (m->type == RRT_MAT_MICROFACET) {
    if (wo.z <= 0 || wi.z <= 0) return S(0, 0, 0);
    spec F = mf_F(m, wi);
    return s_divf(s_mulf(s_mulf(F, (float)mf_G(m, wo, wi)), (float)mf_D(m, vunit(vadd(wo, wi)))), (float)(4 * wo.z * wi.z));
  }
  return S(0, 0, 0);
}

/* BSDF::refract  (src/bsdf.cpp:146-159); ior is a float parameter: 1/ior is a float division */
static int bsdf_refract(v3 wo, v3* wi, float ior) {
  double eta = (wo.z > 0) ? (double)(1 / ior) : (double)ior;
  double wi_z2 = 1 - eta * eta * (1 - wo.z * wo.z);
  if (wi_z2 < 0) return 0;
  *wi = V(-eta * wo.x, -eta * wo.y, sqrt(wi_z2));
  if (wo.z > 0) wi->z = -wi->z;
  return 1;
}

/* BSDF::sample_f for every class  (src/part1_code.cpp:171-173; src/bsdf.cpp:37-41, 74-96, 104-106, 114-140, 167-171;
 * src/sampler.cpp:47-56).  Draws come from g in the reference's order. */
static spec bsdf_sample_f(const rrt_material* m, v3 wo, v3* wi, float* pdf, rng_t* g) {
  switch (m->type) {
    case RRT_MAT_DIFFUSE:
    case RRT_MAT_EMISSION: {
      double Xi1 = rng_uniform(g), Xi2 = rng_uniform(g);
      double r = sqrt(Xi1), theta = 2. * ORC_PI * Xi2;
      *pdf = (float)(sqrt(1 - Xi1) / ORC_PI);
      *wi = V(r * cos(theta), r * sin(theta), sqrt(1 - Xi1));
      return m->type == RRT_MAT_DIFFUSE ? bsdf_f(m, wo, *wi) : S(0, 0, 0);
    }
    case RRT_MAT_MIRROR:
      *wi = V(-wo.x, -wo.y, wo.z); *pdf = 1.0f;
      return s_divf(s3(m->a), (float)fabs(wi->z));
    case RRT_MAT_GLASS: {
      if (bsdf_refract(wo, wi, m->ior)) {
        double R0 = (double)((1 - m->ior) / (1 + m->ior));
        R0 *= R0;
        double t = (1 - fabs(wi->z)), t2 = t * t, t4 = t2 * t2, R = R0 + (1 - R0) * t4 * t;
        if (rng_uniform(g) < R) {
          *wi = V(-wo.x, -wo.y, wo.z); *pdf = (float)R;
          return s_divf(s_mulf(s3(m->a), (float)R), (float)fabs(wi->z));
        } else {
          double eta = (wo.z > 0) ? (double)(1 / m->ior) : (double)m->ior;
          *pdf = (float)(1 - R);
          return s_divf(s_mulf(s3(m->b), (float)(1 - R)), (float)(fabs(wi->z) * eta * eta));
        }
      }
      *wi = V(-wo.x, -wo.y, wo.z); *pdf = 1.0f;
      return s_divf(s3(m->a), (float)fabs(wi->z));
    }
    case RRT_MAT_MICROFACET: {
      double ux, uy; rng_grid2(g, &ux, &uy);
      double alpha2 = m->alpha * m->alpha, /* float product */
             theta_h = atan(sqrt(-alpha2 * log(1 - ux))), phi_h = 2 * M_PI * uy, sin_theta_h = sin(theta_h),
             cos_theta_h = cos(theta_h), tan_theta_h = tan(theta_h),
             p_theta = 2 * sin_theta_h * exp(-tan_theta_h * tan_theta_h / alpha2) / (alpha2 * cos_theta_h * cos_theta_h * cos_theta_h),
             p_phi = 0.5 / M_PI;
      v3 h = V(sin_theta_h * cos(phi_h), sin_theta_h * sin(phi_h), cos_theta_h);
      *wi = vsub(smul(2 * vdot(wo, h), h), wo);
      if (wi->z <= 0) { *pdf = 0; return S(0, 0, 0); }
      *pdf = (float)(p_theta * p_phi / (sin_theta_h * 4 * vdot(*wi, h)));
      return bsdf_f(m, wo, *wi);
    }
    default: /* RefractionBSDF: the reference returns Spectrum() and leaves pdf unwritten (src/bsdf.cpp:104-106); pdf = 0 ends the path */
      *wi = V(0, 0, 1); *pdf = 0; return S(0, 0, 0);
  }
}

/* ------------------------------------------------------------------ lights: SceneLight::sample_L  (src/static_scene/light.cpp:17-92) */
/* ---- EnvironmentLight ---- */
/* init (src/static_scene/environment_light.cpp:21-49).  The reference accumulates into an uninitialised marginal_y
 * (new double[h]); zero is what a fresh heap gives it and what is restated here. */
void orc_set_envmap(orc_ctx* c, int w, int h, const float* rgb) {
  free(c->env); free(c->env_pdf); free(c->env_cond); free(c->env_marg);
  c->env = NULL; c->env_pdf = c->env_cond = c->env_marg = NULL; c->env_w = c->env_h = 0;
  if (!rgb || w <= 0 || h <= 0) return;
  c->env_w = w; c->env_h = h;
  c->env = (float*)dup_mem(rgb, sizeof(float) * 3 * (size_t)w * h);
  c->env_pdf = (double*)malloc(sizeof(double) * (size_t)w * h);
  c->env_cond = (double*)malloc(sizeof(double) * (size_t)w * h);
  c->env_marg = (double*)calloc((size_t)h, sizeof(double));
  double sum = 0;
  for (int j = 0; j < h; ++j)
    for (int i = 0; i < w; ++i) {
      c->env_pdf[(size_t)w * j + i] = s_illum(s3(rgb + 3 * ((size_t)w * j + i))) * sin(ORC_PI * (j + .5) / h);
      sum += c->env_pdf[(size_t)w * j + i];
    }
  for (int j = 0; j < h; ++j) {
    for (int i = 0; i < w; ++i) c->env_marg[j] += (c->env_pdf[(size_t)w * j + i] /= sum);
    for (int i = 0; i < w; ++i) {
      c->env_cond[(size_t)w * j + i] = c->env_pdf[(size_t)w * j + i] / c->env_marg[j];
      if (i > 0) c->env_cond[(size_t)w * j + i] += c->env_cond[(size_t)w * j + i - 1];
    }
    if (j > 0) c->env_marg[j] += c->env_marg[j - 1];
  }
}

/* bilerp (src/static_scene/environment_light.cpp:99-115) */
static spec env_bilerp(const orc_ctx* c, double x, double y) {
  const long w = c->env_w, h = c->env_h;
  long right = lround(x), left, v = lround(y);
  double u1 = right - x + .5, v1;
  if (right == 0 || right == w) { left = w - 1; right = 0; } else left = right - 1;
  if (v == 0) { v1 = v = 1; } else if (v == h) { v = h - 1; v1 = 0; } else v1 = v - y + .5;
  size_t bottom = (size_t)w * (size_t)v, top = bottom - (size_t)w;
  double u0 = 1 - u1;
  spec tl = s3(c->env + 3 * (top + left)), tr = s3(c->env + 3 * (top + right));
  spec bl = s3(c->env + 3 * (bottom + left)), br = s3(c->env + 3 * (bottom + right));
  return s_add(s_mulf(s_add(s_mulf(tl, (float)u1), s_mulf(tr, (float)u0)), (float)v1),
               s_mulf(s_add(s_mulf(bl, (float)u1), s_mulf(br, (float)u0)), (float)(1 - v1)));
}

/* sample_dir = bilerp(theta_phi_to_xy(dir_to_theta_phi(r.d)))  (:76-80, :60-65, :145-147) */
static spec env_sample_dir(const orc_ctx* c, v3 d) {
  v3 u = vunit(d);
  double theta = acos(u.y), phi = atan2(-u.z, u.x) + ORC_PI;
  double x = phi / 2. / ORC_PI * c->env_w, y = theta / ORC_PI * c->env_h;
  return env_bilerp(c, x, y);
}

/* first index whose value is greater than key (std::upper_bound) */
static int upper_bound_d(const double* a, int n, double key) {
  int lo = 0, hi = n;
  while (lo < hi) { int mid = lo + (hi - lo) / 2; if (!(key < a[mid])) lo = mid + 1; else hi = mid; }
  return lo;
}

/* sample_L, ENV_HEMI == 0 (:118-133).  A draw of exactly 1.0 would index one past the tables in the reference
 * (upper_bound returns h or w); clamped here, as on the device. */
static spec env_sample_L(const orc_ctx* c, v3* wi, float* dist, float* pdf, rng_t* g) {
  const int w = c->env_w, h = c->env_h;
  *dist = INFINITY;
  double sx, sy; rng_grid2(g, &sx, &sy);
  int y = upper_bound_d(c->env_marg, h, sy); if (y > h - 1) y = h - 1;
  const double* cond = c->env_cond + (size_t)w * y;
  int x = upper_bound_d(cond, w, sx); if (x > w - 1) x = w - 1;
  double phi = (double)x / w * 2.0 * ORC_PI, theta = (double)y / h * ORC_PI;     /* xy_to_theta_phi (:67-74) */
  double dy = cos(theta), dx = cos(phi - ORC_PI) * sin(theta), dz = -sin(phi - ORC_PI) * sin(theta);   /* theta_phi_to_dir (:82-91) */
  *wi = V(dx, dy, dz);
  *pdf = (float)(c->env_pdf[(size_t)w * y + x] * w * h / (2 * ORC_PI * ORC_PI * sin(theta)));
  return env_bilerp(c, (double)x, (double)y);
}

static spec light_sample_L(const rrt_light* l, v3 p, v3* wi, float* dist, float* pdf, rng_t* g) {
  switch (l->type) {
    case RRT_LIGHT_DIRECTIONAL:
      *wi = V(l->direction[0], l->direction[1], l->direction[2]); *dist = INFINITY; *pdf = 1.0f;
      return s3(l->radiance);
    case RRT_LIGHT_HEMISPHERE: { /* UniformHemisphereSampler3D (src/sampler.cpp:15-29): float sinf/cosf */
      double Xi1 = rng_uniform(g), Xi2 = rng_uniform(g);
      double theta = acos(Xi1), phi = 2.0 * ORC_PI * Xi2;
      double xs = sinf(theta) * cosf(phi), ys = sinf(theta) * sinf(phi), zs = cosf(theta);
      v3 col[3] = {V(1, 0, 0), V(0, 0, -1), V(0, 1, 0)};
      *wi = mat_mul(col, V(xs, ys, zs)); *dist = INFINITY; *pdf = (float)(1.0 / (2.0 * ORC_PI));
      return s3(l->radiance);
    }
    case RRT_LIGHT_POINT: {
      v3 d = vsub(V(l->position[0], l->position[1], l->position[2]), p);
      *wi = vunit(d); *dist = (float)vnorm(d); *pdf = 1.0f;
      return s3(l->radiance);
    }
    case RRT_LIGHT_AREA: {
      double sx, sy; rng_grid2(g, &sx, &sy);
      sx = sx - 0.5f; sy = sy - 0.5f;
      v3 pos = V(l->position[0], l->position[1], l->position[2]), dx = V(l->dim_x[0], l->dim_x[1], l->dim_x[2]),
         dy = V(l->dim_y[0], l->dim_y[1], l->dim_y[2]), dir = V(l->direction[0], l->direction[1], l->direction[2]);
      v3 d = vsub(vadd(vadd(pos, smul(sx, dx)), smul(sy, dy)), p);
      float sqDist = (float)vnorm2(d);
      float dst = sqrtf(sqDist);
      *wi = vdiv(d, dst);
      float cosTheta = (float)vdot(*wi, dir);
      *dist = dst;
      *pdf = sqDist / (l->area * fabsf(cosTheta));
      return cosTheta < 0 ? s3(l->radiance) : S(0, 0, 0);
    }
    default:
      *wi = V(0, 0, 0); *dist = 0; *pdf = 0; return S(0, 0, 0);
  }
}

/* ------------------------------------------------------------------ estimator  (src/part1_code.cpp:15-163) */
typedef struct {
  const orc_ctx* c; const rrt_params* p; rng_t* g; cnt_t* cnt;
} est_t;

static const rrt_material* prim_material(const orc_ctx* c, int prim) {
  uint32_t k = c->prim_index[prim];
  return &c->mats[c->prim_kind[prim] == 0 ? c->tri_mat[k] : c->sph_mat[k]];
}

/* UniformHemisphereSampler3D::get_sample  (src/sampler.cpp:15-29) */
static v3 hemisphere_sample(rng_t* g) {
  double Xi1 = rng_uniform(g), Xi2 = rng_uniform(g);
  double theta = acos(Xi1), phi = 2.0 * ORC_PI * Xi2;
  return V(sinf(theta) * cosf(phi), sinf(theta) * sinf(phi), cosf(theta));
}

/* estimate_direct_lighting_hemisphere  (src/part1_code.cpp:15-31) */
static spec direct_hemisphere(est_t* e, const isect_t* isect) {
  const orc_ctx* c = e->c;
  v3 o2w[3]; make_coord_space(o2w, isect->n);
  v3 w_out = mat_tmul(o2w, isect->w_out);
  const rrt_material* m = prim_material(c, isect->prim);
  int num_samples = (int)(c->n_lights + (c->env ? 1u : 0u)) * e->p->ns_area_light;   /* scene->lights.size() counts the environment light */
  spec L_out = S(0, 0, 0);
  for (int i = 0; i < num_samples; ++i) {
    v3 w_in = hemisphere_sample(e->g);
    v3 wi_world = mat_mul(o2w, w_in);
    ray_t r2; r2.o = vadd(isect->hit_p, smul(ORC_EPS_D, wi_world)); r2.d = wi_world; r2.min_t = 0.0; r2.max_t = INFINITY;
    isect_t i2;
    if (bvh_intersect(c, &r2, &i2, NULL, NULL, NULL, e->cnt) == 1)
      L_out = s_add(L_out, s_mulf(s_mul(mat_emission(prim_material(c, i2.prim)), bsdf_f(m, w_out, w_in)), (float)w_in.z));
  }
  return s_divf(s_mulf(s_mulf(L_out, 2.f), (float)M_PI), (float)num_samples);
}

/* estimate_direct_lighting_importance  (src/part1_code.cpp:33-57) */
static spec direct_importance(est_t* e, const isect_t* isect) {
  const orc_ctx* c = e->c;
  v3 o2w[3]; make_coord_space(o2w, isect->n);
  v3 w_out = mat_tmul(o2w, isect->w_out);
  const rrt_material* m = prim_material(c, isect->prim);
  spec L_out = S(0, 0, 0);
  int total = 0;
  const uint32_t n_all = c->n_lights + (c->env ? 1u : 0u);
  for (uint32_t li = 0; li < n_all; ++li) {
    const rrt_light* l = li < c->n_lights ? &c->lights[li] : NULL;   /* NULL: the environment light, last in the list */
    int ns = (l && l->is_delta) ? 1 : e->p->ns_area_light;
    total += ns;
    for (int i = 0; i < ns; ++i) {
      v3 wi_world; float dist, pdf;
      spec sample = l ? light_sample_L(l, isect->hit_p, &wi_world, &dist, &pdf, e->g) : env_sample_L(c, &wi_world, &dist, &pdf, e->g);
      if (l && l->type == RRT_LIGHT_NULL) continue; /* stub lights return zero radiance */
      v3 w_in = mat_tmul(o2w, wi_world);
      if (w_in.z < 0) continue;
      ray_t sr; sr.o = vadd(isect->hit_p, smul(ORC_EPS_D, wi_world)); sr.d = wi_world; sr.min_t = 0.0; sr.max_t = dist;
      if (bvh_intersect(c, &sr, NULL, NULL, NULL, NULL, e->cnt) != 1)
        L_out = s_add(L_out, s_divf(s_mulf(s_mul(sample, bsdf_f(m, w_out, w_in)), (float)w_in.z), pdf));
    }
  }
  return s_divf(L_out, (float)total);
}

static spec one_bounce(est_t* e, const isect_t* isect) {
  return e->p->direct_hemisphere_sample ? direct_hemisphere(e, isect) : direct_importance(e, isect);
}

/* at_least_one_bounce_radiance  (src/part1_code.cpp:69-101); depth counts down from max_ray_depth */
static spec at_least_one_bounce(est_t* e, int depth, const isect_t* isect) {
  const orc_ctx* c = e->c;
  v3 o2w[3]; make_coord_space(o2w, isect->n);
  v3 w_out = mat_tmul(o2w, isect->w_out);
  const rrt_material* m = prim_material(c, isect->prim);
  spec L_out = S(0, 0, 0);
  if (!mat_is_delta(m)) L_out = s_add(L_out, one_bounce(e, isect));
  const double prob = 0.7;
  if (depth == e->p->max_ray_depth || (depth > 1 && rng_uniform(e->g) < prob)) {
    v3 w_in; float pdf;
    spec sample = bsdf_sample_f(m, w_out, &w_in, &pdf, e->g);
    if (pdf == 0.0f) return L_out;
    v3 wi_world = mat_mul(o2w, w_in);
    ray_t r2; r2.o = vadd(isect->hit_p, smul(ORC_EPS_D, wi_world)); r2.d = wi_world; r2.min_t = 0.0; r2.max_t = INFINITY;
    isect_t i2;
    if (bvh_intersect(c, &r2, &i2, NULL, NULL, NULL, e->cnt) == 1) {
      spec L = at_least_one_bounce(e, depth - 1, &i2);
      if (mat_is_delta(m)) L = s_add(L, mat_emission(prim_material(c, i2.prim)));
      L_out = s_add(L_out, s_divf(s_divf(s_mulf(s_mul(L, sample), (float)fabs(w_in.z)), pdf), (float)prob));
    }
  }
  return L_out;
}

/* est_radiance_global_illumination  (src/part1_code.cpp:103-123), ILLUM == 2 */
static spec est_radiance(est_t* e, const ray_t* r) {
  isect_t isect;
  if (bvh_intersect(e->c, r, &isect, NULL, NULL, NULL, e->cnt) != 1)   /* a miss OR a capture; the ray's own, unbent direction */
    return e->c->env ? env_sample_dir(e->c, r->d) : S(0, 0, 0);
  spec Le = mat_emission(prim_material(e->c, isect.prim));
  if (e->p->max_ray_depth == 0) return Le;
  if (e->p->max_ray_depth == 1) return s_add(Le, one_bounce(e, &isect));
  return s_add(Le, at_least_one_bounce(e, e->p->max_ray_depth, &isect));
}

/* raytrace_pixel  (src/part1_code.cpp:125-163), ADAPTIVE == 1 */
static spec raytrace_pixel(const orc_ctx* c, const rrt_params* p, int x, int y, int* count, cnt_t* cnt, uint64_t* paths) {
  spec ret = S(0, 0, 0);
  int i; double s1 = 0.0, s2 = 0.0;
  for (i = 0; i < p->ns_aa; ++i) {
    rng_t g; memset(&g, 0, sizeof g); g.seed = p->seed; g.pixel = (uint32_t)(x + y * c->W); g.sample = (uint32_t)i;
    double sx = x, sy = y;
    if (p->ns_aa == 1) { sx += 0.5; sy += 0.5; }
    else { double jx, jy; rng_grid2(&g, &jx, &jy); sx += jx; sy += jy; }
    ray_t r;
    if (p->thin_lens) {
      double lx, ly; rng_grid2(&g, &lx, &ly);
      r = generate_ray_thin_lens(&c->cam, sx / c->W, sy / c->H, lx, ly * 2 * M_PI);
    } else r = generate_ray(&c->cam, sx / c->W, sy / c->H);
    est_t e = {c, p, &g, cnt};
    spec s = est_radiance(&e, &r);
    (*paths)++;
    ret = s_add(ret, s);
    double illum = s_illum(s);
    s1 += illum; s2 += illum * illum;
    if ((i + 1) % p->samples_per_batch == 0) {
      double avg = s1 / (i + 1), sd = sqrt((s2 - avg * s1) / i);
      if (1.96 * sd / sqrt(i + 1) <= p->max_tolerance * avg) { ++i; break; }
    }
  }
  *count = i;
  return s_divf(ret, (float)i);
}

static int g_shard_tile = 32, g_shard_rank = 0, g_shard_world = 1;
void orc_set_tile_shard(orc_ctx* c, int tile, int rank, int world) { (void)c; g_shard_tile = tile; g_shard_rank = rank; g_shard_world = world; }

/* tile loop of PathTracer (src/pathtracer.cpp:243-255, 549-581) flattened to a pixel loop; pixels outside the cell
 * or owned by another shard stay 0 */
int orc_render(orc_ctx* c, const rrt_params* p, float* hdr_rgb, int32_t* counts, rrt_stats* st) {
  if (!c->W || !c->H || p->samples_per_batch <= 0) return RRT_E_STATE;
  int x0 = 0, y0 = 0, x1 = c->W, y1 = c->H;
  if (c->cell_w > 0) { x0 = c->cell_x; y0 = c->cell_y; x1 = x0 + c->cell_w; y1 = y0 + c->cell_h; if (x1 > c->W) x1 = c->W; if (y1 > c->H) y1 = c->H; }
  if (hdr_rgb) memset(hdr_rgb, 0, sizeof(float) * 3 * (size_t)c->W * c->H);
  if (counts) memset(counts, 0, sizeof(int32_t) * (size_t)c->W * c->H);
  /* owner of tile (tx, ty) = (tx + ty * stride) % world, stride = the smallest integer >= round(sqrt(world)) coprime with world
   * (the product's rule, include/rrt_b200.h rrt_set_tile_shard) */
  int ntx = 1;
  while ((ntx + 1) * (ntx + 1) <= g_shard_world + ntx) ++ntx;
  for (;; ++ntx) { int a = ntx, b = g_shard_world; while (b) { int t = a % b; a = b; b = t; } if (a == 1) break; }
  cnt_t tot = {0, 0, 0, 0, 0}; uint64_t paths = 0;
  double t0 = 0, t1 = 0;
#ifdef _OPENMP
  t0 = omp_get_wtime();
#endif
#pragma omp parallel
  {
    cnt_t cnt = {0, 0, 0, 0, 0}; uint64_t mypaths = 0;
#pragma omp for schedule(dynamic, 1) nowait
    for (int y = y0; y < y1; ++y)
      for (int x = x0; x < x1; ++x) {
        int owner = ((x / g_shard_tile) + (y / g_shard_tile) * ntx) % g_shard_world;
        if (owner != g_shard_rank) continue;
        int n; spec s = raytrace_pixel(c, p, x, y, &n, &cnt, &mypaths);
        size_t idx = (size_t)x + (size_t)y * c->W;
        if (hdr_rgb) { hdr_rgb[3 * idx] = s.r; hdr_rgb[3 * idx + 1] = s.g; hdr_rgb[3 * idx + 2] = s.b; }
        if (counts) counts[idx] = n;
      }
#pragma omp critical
    { tot.rays += cnt.rays; tot.steps += cnt.steps; tot.nodes += cnt.nodes; tot.tris += cnt.tris; tot.sph += cnt.sph; paths += mypaths; }
  }
#ifdef _OPENMP
  t1 = omp_get_wtime();
#endif
  if (st) {
    memset(st, 0, sizeof *st);
    st->seconds = t1 - t0; st->paths = paths; st->rays = tot.rays; st->geodesic_steps = tot.steps;
    st->node_visits = tot.nodes; st->tri_tests = tot.tris; st->sphere_tests = tot.sph;
  }
  return 0;
}

/* ------------------------------------------------------------------ parity hooks (same signatures as the C-ABI's rrt_* hooks) */
/* seg_out [n*nsteps*8]: chord origin, unit direction, length, captured-by-horizon flag */
void orc_geodesic(const orc_ctx* c, int64_t n, const double* o, const double* d, int nsteps, double* seg_out) {
  for (int64_t i = 0; i < n; ++i) {
    ray_t micro; micro.o = V(o[3 * i], o[3 * i + 1], o[3 * i + 2]); micro.d = V(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
    micro.min_t = 0.0; micro.max_t = 0.0;
    for (int j = 0; j < nsteps; ++j) {
      micro = next_micro_ray(c, &micro);
      double* s = seg_out + ((size_t)i * nsteps + j) * 8;
      s[0] = micro.o.x; s[1] = micro.o.y; s[2] = micro.o.z; s[3] = micro.d.x; s[4] = micro.d.y; s[5] = micro.d.z; s[6] = micro.max_t;
      ray_t probe = micro;
      s[7] = sphere_intersect(c->bh_c, c->bh_r2, &probe, NULL, -1) ? 1.0 : 0.0;
    }
  }
}

void orc_trace_rays(const orc_ctx* c, int64_t n, const double* o, const double* d, const double* min_t, const double* max_t,
                    int any_hit, rrt_hit* out, rrt_stats* st) {
  cnt_t tot = {0, 0, 0, 0, 0};
#pragma omp parallel
  {
    cnt_t cnt = {0, 0, 0, 0, 0};
#pragma omp for schedule(dynamic, 64) nowait
    for (int64_t i = 0; i < n; ++i) {
      ray_t r; r.o = V(o[3 * i], o[3 * i + 1], o[3 * i + 2]); r.d = V(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
      r.min_t = min_t ? min_t[i] : 0.0; r.max_t = max_t ? max_t[i] : INFINITY;
      isect_t is; memset(&is, 0, sizeof is); is.prim = -1;
      rrt_hit* h = &out[i]; memset(h, 0, sizeof *h);
      h->status = bvh_intersect(c, &r, any_hit ? NULL : &is, &h->chord, &h->steps, &h->t, &cnt);
      h->prim = -1;
      if (h->status == 1 && !any_hit) {
        h->prim = is.prim;
        h->hit_p[0] = is.hit_p.x; h->hit_p[1] = is.hit_p.y; h->hit_p[2] = is.hit_p.z;
        h->w_out[0] = is.w_out.x; h->w_out[1] = is.w_out.y; h->w_out[2] = is.w_out.z;
        h->n[0] = is.n.x; h->n[1] = is.n.y; h->n[2] = is.n.z;
      } else if (h->status != 1) h->t = 0.0;
    }
#pragma omp critical
    { tot.rays += cnt.rays; tot.steps += cnt.steps; tot.nodes += cnt.nodes; tot.tris += cnt.tris; tot.sph += cnt.sph; }
  }
  if (st) {
    memset(st, 0, sizeof *st);
    st->rays = tot.rays; st->geodesic_steps = tot.steps; st->node_visits = tot.nodes; st->tri_tests = tot.tris; st->sphere_tests = tot.sph;
  }
}

void orc_generate_rays(const orc_ctx* c, int64_t n, const double* xy, double* o, double* d) {
  for (int64_t i = 0; i < n; ++i) {
    ray_t r = generate_ray(&c->cam, xy[2 * i], xy[2 * i + 1]);
    o[3 * i] = r.o.x; o[3 * i + 1] = r.o.y; o[3 * i + 2] = r.o.z; d[3 * i] = r.d.x; d[3 * i + 1] = r.d.y; d[3 * i + 2] = r.d.z;
  }
}

/* Camera::generate_ray_for_thin_lens (reference src/camera.cpp:176-184); lens [n*2] = (rndR, rndTheta) */
void orc_generate_rays_thin_lens(const orc_ctx* c, int64_t n, const double* xy, const double* lens, double* o, double* d) {
  for (int64_t i = 0; i < n; ++i) {
    ray_t r = generate_ray_thin_lens(&c->cam, xy[2 * i], xy[2 * i + 1], lens[2 * i], lens[2 * i + 1]);
    o[3 * i] = r.o.x; o[3 * i + 1] = r.o.y; o[3 * i + 2] = r.o.z; d[3 * i] = r.d.x; d[3 * i + 1] = r.d.y; d[3 * i + 2] = r.d.z;
  }
}

/* u [n*4]: the draws the reference would take from rand()/RAND_MAX, in draw order */
int orc_bsdf_sample(const orc_ctx* c, int mat, int64_t n, const double* wo, const double* u, float* f, double* wi, float* pdf) {
  if (mat < 0 || mat >= (int)c->n_mats) return RRT_E_ARG;
  for (int64_t i = 0; i < n; ++i) {
    rng_t g; memset(&g, 0, sizeof g); g.fixed = u + 4 * i;
    v3 w; float p = 0;
    spec s = bsdf_sample_f(&c->mats[mat], V(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]), &w, &p, &g);
    f[3 * i] = s.r; f[3 * i + 1] = s.g; f[3 * i + 2] = s.b; wi[3 * i] = w.x; wi[3 * i + 1] = w.y; wi[3 * i + 2] = w.z; pdf[i] = p;
  }
  return 0;
}
int orc_bsdf_eval(const orc_ctx* c, int mat, int64_t n, const double* wo, const double* wi, float* f) {
  if (mat < 0 || mat >= (int)c->n_mats) return RRT_E_ARG;
  for (int64_t i = 0; i < n; ++i) {
    spec s = bsdf_f(&c->mats[mat], V(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]), V(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]));
    f[3 * i] = s.r; f[3 * i + 1] = s.g; f[3 * i + 2] = s.b;
  }
  return 0;
}
int orc_light_sample(const orc_ctx* c, int light, int64_t n, const double* p, const double* u, float* radiance, double* wi,
                     float* dist, float* pdf) {
  const int is_env = c->env && light == (int)c->n_lights;   /* the environment light is the last light */
  if (light < 0 || (light >= (int)c->n_lights && !is_env)) return RRT_E_ARG;
  for (int64_t i = 0; i < n; ++i) {
    rng_t g; memset(&g, 0, sizeof g); g.fixed = u + 2 * i;
    v3 w; float ds = 0, pd = 0;
    spec s = is_env ? env_sample_L(c, &w, &ds, &pd, &g) : light_sample_L(&c->lights[light], V(p[3 * i], p[3 * i + 1], p[3 * i + 2]), &w, &ds, &pd, &g);
    radiance[3 * i] = s.r; radiance[3 * i + 1] = s.g; radiance[3 * i + 2] = s.b;
    wi[3 * i] = w.x; wi[3 * i + 1] = w.y; wi[3 * i + 2] = w.z; dist[i] = ds; pdf[i] = pd;
  }
  return 0;
}

/* EnvironmentLight::sample_dir for n directions d [n*3] -> radiance [n*3] */
int orc_env_lookup(const orc_ctx* c, int64_t n, const double* d, float* radiance) {
  if (!c->env) return RRT_E_STATE;
  for (int64_t i = 0; i < n; ++i) {
    spec s = env_sample_dir(c, V(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
    radiance[3 * i] = s.r; radiance[3 * i + 1] = s.g; radiance[3 * i + 2] = s.b;
  }
  return 0;
}

void orc_philox(uint64_t seed, int64_t n, const uint32_t* counters, uint32_t* out) {
  for (int64_t i = 0; i < n; ++i) {
    uint32_t c[4] = {counters[4 * i], counters[4 * i + 1], counters[4 * i + 2], counters[4 * i + 3]};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    memcpy(out + 4 * i, c, 16);
  }
}

/* HDRImageBuffer::toColor + ImageBuffer::update_pixel  (src/image.h:183-198, 53-62): gamma 2.2, level 1;
 * the reference's clamp(0.f, 1.f, c) has its arguments in (x, lo, hi) order, i.e. max(1, min(0, c))... evaluated
 * literally below with the reference's clamp definition (src/bsdf.h:20-22 is the only clamp in scope: double). */
void orc_tonemap(int64_t n, const float* hdr, uint32_t* rgba) {
  float gamma = 2.2f, level = 1.0f, one_over_gamma = 1.0f / gamma;
  float exposure = (float)sqrt(pow(2, level));
  for (int64_t i = 0; i < n; ++i) {
    float c[3];
    for (int k = 0; k < 3; ++k) c[k] = powf(hdr[3 * i + k] * exposure, one_over_gamma);
    uint32_t p = 0;
    for (int k = 0; k < 3; ++k) {
      /* clamp(n = 0, lower = 1, upper = c) = max(1, min(0, c)) per the literal argument order?  No: see DESIGN.md —
       * ImageBuffer::update_pixel's clamp resolves to CGL::clamp(T x, T lo, T hi) of CGL/misc.h with x = 0, lo = 1, hi = c
       * being a template on float: it returns min(max(x, lo), hi)-style; the observable result is min(1, c) for c >= 0. */
      float v = c[k] < 1.f ? c[k] : 1.f;
      if (!(v >= 0.f)) v = 0.f;
      p |= ((uint32_t)(v * 255)) << (8 * k);
    }
    rgba[i] = p | 0xFF000000u;
  }
}
