// rrt_trace.cuh — device side of BVHAccel::intersect: the Schwarzschild chord stepper, the horizon test and
// the BVH traversal of one chord.  Reference: src/bvh.cpp:99-138, src/static_scene/blackhole.cpp:13-40,
// src/static_scene/sphere.cpp:10-53, src/static_scene/triangle.cpp:25-55, src/bbox.cpp:10-25.
//
// Exactness contract (DESIGN.md "Numerics"): chord recurrence, horizon test and primitive tests are IEEE
// double in the reference's operation order (bit-identical results); interior-node slab tests are FP32 and
// CONSERVATIVE (never reject a box the reference accepts), so which leaves get tested may be a superset
// of the reference's, and the primitive tests alone decide the hit.  Equal-t ties go to the primitive
// with the higher reference depth-first rank (the reference keeps the last one it visits).
#pragma once
#include <type_traits>
#include "rrt_math.cuh"
#include "../../include/rrt_b200.h"

#ifndef RRT_TOP_AABB
#define RRT_TOP_AABB 0            /* A/B: 1 = short chords test the tree-top record with their bounding box (traverse()).  Bit-identical, and not a gain on
                                     B200 (profiles/r02_r_ab_top_aabb.jsonl): bunny in-box 165.5 -> 164.5 ms, bunny 1080p 75.4 -> 76.2, C5 51.3 -> 52.0 with ten times
                                     the triangle tests (380 M against 39 M): what the cheaper test saves in the record it loses in leaves it lets through */
#endif
#ifndef RRT_TOP_AABB_MAXLEN
#define RRT_TOP_AABB_MAXLEN 0.5f  /* ... "short": up to this length; longer chords (escaped rays coming back through the scene) keep the slab test */
#endif
#ifndef RRT_STACK
#define RRT_STACK 40
#define RRT_SPILL 136   /* wide traversal: entries beyond the shared-memory stack go to local memory (never seen in practice) */
#endif  /* traversal stack entries per thread (shared memory); tree depth bound */

struct DevScene {
  const float4* nodes;        // 4 float4 per interior node
  const double2* leaf;        // 5 double2 per leaf record
  const double* tri_nrm;      // 9 per triangle: vertex normals
  const uint8_t* prim_kind;   // [n_prims]
  const uint32_t* prim_index; // [n_prims]
  const int32_t* prim_mat;    // [n_prims]
  const double* sph;          // 4 per sphere
  const rrt_material* mats;
  const rrt_light* lights;
  int32_t n_prims, n_mats, n_lights, root;
  float root_lo[3], root_hi[3];   // root box, rounded outward
  double ctr[3], rad;             // sphere around the root box (far-chord origin shift, boxctx_make)
  // environment light (rrt_set_envmap): map [h][w][3] row 0 = top, and the reference's sampling tables
  const float* env; const double* env_pdf; const double* env_cond; const double* env_marg;
  int env_w, env_h;
  const float* top;               // binary trees: the top of the tree collapsed into ONE 8-wide record (64 floats, layout of `wide`; child codes
                                  // are binary node indices / leaf codes), or nullptr: see traverse()
  const float* wide;              // wide nodes (rrt_bvh.h): 8 * width floats each
  int32_t wroot;                  // child code of the wide root
  int width;                      // 2: `nodes`; 4 / 8: `wide`
  // clearance grid (rrt_bvh.h): lower bound on the distance from any point of a cell to any primitive
  const float* clearance;
  int gnx, gny, gnz;
  double glo[3], ginv[3];
};

struct DevHole {            // global_black_hole (reference src/static_scene/blackhole.cpp:5) + per-frame constants
  double cx, cy, cz, r, r2, h, cos_h, sin_h;
  int32_t mode, nchords;
  // FP32 stepper (RRT_OPT_STEPPER = 1, fast_step below): the same constants rounded once on the host
  float f_cx, f_cy, f_cz, f_r2, f_r15 /* 1.5 r */, f_h, f_h2 /* h/2 */, f_h24 /* h^2/4 */, f_h26 /* h^2/6 */;
  float f_ch, f_sh, f_k /* 1 - cos h */, f_cot /* cos h / sin h */, f_tan2 /* tan(h/2) */;
};

struct Counters { unsigned long long rays, steps, nodes, tris, sph, paths; };

struct Hit {
  int status;   // 0 miss, 1 hit, 2 captured
  int prim, chord, steps;
  double t;
  d3 p, w_out, n;
};

// one chord (or, in flat mode, the whole ray)
struct Seg { d3 o, d; double tmin, tmax; };

// best hit so far inside a chord
struct Best { double t, b1, b2; int prim; int rank; int kind; };

/* BlackHole::f (reference src/static_scene/blackhole.cpp:13-15) */
__device__ __forceinline__ double bh_f(double r, double u) { return -u + 3.0 * r * u * u / 2.0; }

/* Two of the stepper's six divisions have a divisor whose correctly rounded reciprocal is already at hand, and for those the IEEE quotient
 * RN(a / b) can be had from FMAs instead of the ~14 FP64-pipe slots of a division (MUFU seed + Newton steps + range check):
 *   y = RN(1 / b) exactly;  q0 = RN(a y);  r0 = RN(a - b q0);  q1 = RN(q0 + r0 y);  r1 = a - b q1 (exact);  q = RN(q1 + r1 y).
 * q1 is within 1 ulp of a / b (|a y - a / b| <= |a / b| 2^-53, one rounding, one correction); with q1 faithful, r1 exact and y the correctly
 * rounded reciprocal, RN(q1 + r1 y) = RN(a / b) (Markstein 1990, Theorem on the final division step; Cornea, Harrison, Tang, "Scientific
 * computing on Itanium-based systems", Thm 8.1): q1 + r1 y differs from a / b by |r1| |y - 1 / b| < 2^-53 ulp, while a quotient of two doubles is
 * never closer than 2^-53 ulp to a rounding boundary without lying on it.  Operands outside [1e-140, 1e140] (and inf / NaN / 0) take the
 * division itself.  For the constant divisor 6 three steps suffice: a / 6 is at least ulp / 6 away from every rounding boundary (6 M = a has no
 * solution with M a midpoint: 3 x odd needs 55 bits), q0 + r0 y is within 1.5 x 2^-53 ulp of a / 6.
 * tools/div_check.cu compares both against `/` on the device for 2^36 operand pairs incl. adversarial significands: 0 mismatches
 * (profiles/r02_e_div_check.txt); the chord goldens (bit-exact against the reference's own chords) run through this code. */
#ifndef RRT_FAST_DIV
#define RRT_FAST_DIV 0   /* exact (0 mismatches in 2.7e11 pairs) but NOT faster on B200: tools/ilp_probe 46.5 -> 43.3 G chords/s at 6 blocks per SM, frames
                          unchanged (C2 32.0 / 32.2 ms, C5 51.1 / 50.9): the two divisions are not what the stepper waits for.  A/B record. */
#endif
__device__ __forceinline__ bool div_safe(double x) {   /* 1e-140 < |x| < 1e140, finite, non-zero: biased exponent in [558, 1488] */
  const unsigned e = ((unsigned)__double2hiint(x) >> 20) & 0x7ffu;
  return e - 558u <= 930u;
}
/* RN(a / b) given rb == RN(1 / b) */
__device__ __forceinline__ double div_known_rcp(double a, double b, double rb) {
#if RRT_FAST_DIV
  const double q0 = a * rb;
  if (div_safe(b) && div_safe(q0)) {
    const double r0 = __fma_rn(-b, q0, a);
    const double q1 = __fma_rn(r0, rb, q0);
    const double r1 = __fma_rn(-b, q1, a);
    return __fma_rn(r1, rb, q1);
  }
#endif
  return a / b;
}
/* RN(a / 6) */
__device__ __forceinline__ double div_by_6(double a) {
#if RRT_FAST_DIV
  const double y = 0.16666666666666666;   /* RN(1 / 6) */
  if (div_safe(a)) {
    const double q0 = a * y;
    const double r0 = __fma_rn(-6.0, q0, a);
    return __fma_rn(r0, y, q0);
  }
#endif
  return a / 6.0;
}

/* BlackHole::next_micro_ray (reference src/static_scene/blackhole.cpp:17-40).  In: previous chord (o, d, len).
 * Out: next chord.  The conserved plane (x_axis, y_axis) is re-derived from the chord each call, exactly as the
 * reference does; nothing but (o, d, len) is carried. */
__device__ __forceinline__ void next_micro_ray(const DevHole& bh, d3& o, d3& d, double& len) {
  const double h = bh.h;
  d3 c = D3(bh.cx, bh.cy, bh.cz);
  d3 p = vadd(o, vmul(d, len));
  d3 x_axis = vsub(p, c);
  double dist = vnorm(x_axis);
  double u = 1.0 / dist;              /* both `1. / norm()` in normalize() and `u = 1 / d` */
  x_axis = D3(x_axis.x * u, x_axis.y * u, x_axis.z * u);
  double dx = vdot(d, x_axis);
  d3 y_axis = vsub(d, smul(dx, x_axis));
  double dy = vnorm(y_axis);
  double rdy = 1.0 / dy;
  y_axis = D3(y_axis.x * rdy, y_axis.y * rdy, y_axis.z * rdy);
  double up = div_known_rcp(-u * dx, dy, rdy);   /* -u * dx / dy */
  double f1 = bh_f(bh.r, u);
  double f2 = bh_f(bh.r, u + up * h / 2.0);
  double f3 = bh_f(bh.r, u + up * h / 2.0 + f1 * h * h / 4.0);
  u += up * h + div_by_6((f1 + f2 + f3) * h * h);   /* ... * h * h / 6.0 */
  double dn = 1.0 / u;
  double next_x = dn * bh.cos_h;
  double next_y = dn * bh.sin_h;
  d3 q = vsub(vadd(vadd(c, smul(next_x, x_axis)), smul(next_y, y_axis)), p);
  double l = vnorm(q);
  double rl = 1.0 / l;
  o = p; d = D3(q.x * rl, q.y * rl, q.z * rl); len = l;
}

/* Sphere::test + the range logic of Sphere::intersect (reference src/static_scene/sphere.cpp:10-53) for the horizon:
 * true when the chord [0, len] crosses the sphere of radius^2 r2 about c.
 * (Measured and dropped: skipping the roots when |o - c| (1 - 1e-6) > len + 2 r, which holds for almost every chord
 * and removes ~13 of a step's ~175 FP64 instructions, made C2 0.6 ms per frame SLOWER (37.7 vs 37.05 ms): this
 * arithmetic is independent of the traversal set-up that follows and fills its latency gaps.  DESIGN.md section 6.) */
__device__ __forceinline__ bool horizon_hit(const DevHole& bh, d3 o, d3 d, double len) {
  d3 tmp = vsub(o, D3(bh.cx, bh.cy, bh.cz));
  double b = 2 * vdot(tmp, d), c = vnorm2(tmp) - bh.r2, disc = b * b - 4 * c;
  if (disc < 0) return false;
  double sq = sqrt(disc);
  double t1 = (-b - sq) / 2, t2 = (-b + sq) / 2;
  return (0.0 <= t1 && t1 <= len) || (0.0 <= t2 && t2 <= len);
}

/* ---------------------------------------------------------------- FP32 state-carrying stepper (RRT_OPT_STEPPER = 1)
 * The reference re-derives the orbital frame and u' = du/dphi from the previous chord every step
 * (src/static_scene/blackhole.cpp:20-29).  Written out in the orbital plane that recurrence is EXACTLY
 *     G    = h u' + (f1 + f2 + f3) h^2 / 6          f1..f3 as in the reference, f(u) = -u + 1.5 r u^2
 *     u+   = u + G                                   (signed: u+ <= 0 is a ray that left; the reference carries on)
 *     x^'  = sign(u+) (cos h x^ + sin h y^)           new radial unit vector   (p' = c + x^' / |u+|)
 *     y^'  = -sin h x^ + cos h y^                     new tangential unit vector: the chord's tangential part at p'
 *                                                     is sin h / u > 0 along it whatever the sign of u+
 *     u''  = (u+ cos h - u) / sin h = G cot h - u tan(h/2)     what -u dx/dy of the NEW chord evaluates to
 *     u    = |u+|
 * (derivation: DESIGN.md section 6.2; with u+ < 0 the point lands on the far side of the hole, the frame changes
 * handedness and the ray falls into the reference's two-point cycle, reproduced here because images depend on it).
 * So a ray is (x^, y^, u, u') — the conserved orbital plane and the scalar state — in 8 registers; nothing is
 * re-derived from positions, no division or square root touches the carried state, and the chord is formed from
 * small terms:  q = p' - p = x^ A + y^ B,  A = -(u (1 - cos h) + G) / (u u+),  B = sin h / u+.
 * Per chord: ~70 FP32 instructions + 3 MUFU instead of ~175 FP64 instructions.  Accuracy against the reference's
 * doubles (tests/test_gpu_parity.py::test_fp32_stepper_*): chord end points and directions ~1e-6 relative after 63
 * chords — well inside BASELINE.json's 1e-4 — but NOT bit-identical, so primitive ids can differ for rays that graze
 * an edge; that is why this is an option and the FP64 stepper is the default. */
struct FastRay { float xx, xy, xz, yx, yy, yz, u, up; };

__device__ __forceinline__ float fast_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float fast_rsqrt(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

/* first chord: the reference's own derivation from (o, d) with len = 0 (src/static_scene/blackhole.cpp:18-29), once per
 * ray.  In double because near-radial rays cancel in d - dx x^, but without double divisions or square roots: the few
 * lanes of a warp that are being refilled execute this alone (ncu: ~3 active threads per instruction), so its length is
 * paid in full.  1/sqrt(.) = FP32 MUFU seed (rel. 2^-22) + one Newton step in double (rel. < 1e-13). */
__device__ __forceinline__ double fast_rsqrt_d(double a) {
  const double y = (double)fast_rsqrt((float)a);
  return y * (1.5 - 0.5 * a * y * y);
}
__device__ __forceinline__ void fast_init(const DevHole& bh, const d3& o, const d3& d, FastRay& fr) {
  d3 x = vsub(o, D3(bh.cx, bh.cy, bh.cz));
  const double u = fast_rsqrt_d(vnorm2(x));
  x = D3(x.x * u, x.y * u, x.z * u);
  const double dx = vdot(d, x);
  const d3 y = vsub(d, smul(dx, x));
  const double rdy = fast_rsqrt_d(vnorm2(y));
  fr.xx = (float)x.x; fr.xy = (float)x.y; fr.xz = (float)x.z;
  fr.yx = (float)(y.x * rdy); fr.yy = (float)(y.y * rdy); fr.yz = (float)(y.z * rdy);
  fr.u = (float)u; fr.up = (float)(-u * dx * rdy);
}

/* one chord: returns origin, unit direction and length of the chord from the current point to the next one, whether
 * it crosses the horizon (Sphere::intersect on the hole, src/static_scene/sphere.cpp:10-53, in plane coordinates:
 * distance of the chord's line from the centre = rho |B| / len, foot at t = -rho A / len), and advances the state */
__device__ __forceinline__ bool fast_step(const DevHole& bh, FastRay& fr, float& ox, float& oy, float& oz, float& dx, float& dy, float& dz, float& len) {
  const float u = fr.u, up = fr.up;
  const float rho = fast_rcp(u);
  ox = __fmaf_rn(rho, fr.xx, bh.f_cx); oy = __fmaf_rn(rho, fr.xy, bh.f_cy); oz = __fmaf_rn(rho, fr.xz, bh.f_cz);
  const float f1 = __fmaf_rn(bh.f_r15 * u, u, -u);
  const float v2 = __fmaf_rn(up, bh.f_h2, u);
  const float f2 = __fmaf_rn(bh.f_r15 * v2, v2, -v2);
  const float v3 = __fmaf_rn(f1, bh.f_h24, v2);
  const float f3 = __fmaf_rn(bh.f_r15 * v3, v3, -v3);
  const float G = __fmaf_rn(up, bh.f_h, (f1 + f2 + f3) * bh.f_h26);
  const float un = u + G;
  const float rhon = fast_rcp(un);
  const float A = -__fmaf_rn(u, bh.f_k, G) * rho * rhon;
  const float B = bh.f_sh * rhon;
  const float qx = __fmaf_rn(A, fr.xx, B * fr.yx), qy = __fmaf_rn(A, fr.xy, B * fr.yy), qz = __fmaf_rn(A, fr.xz, B * fr.yz);
  const float l2 = __fmaf_rn(qx, qx, __fmaf_rn(qy, qy, qz * qz));
  const float rl = fast_rsqrt(l2);
  len = l2 * rl;
  dx = qx * rl; dy = qy * rl; dz = qz * rl;
  /* horizon */
  const float perp = rho * B * rl;
  const float disc = __fmaf_rn(-perp, perp, bh.f_r2);
  bool captured = false;
  if (disc >= 0.f) {
    const float sq = sqrtf(disc), tc = -rho * A * rl;
    const float t1 = tc - sq, t2 = tc + sq;
    captured = (0.f <= t1 && t1 <= len) || (0.f <= t2 && t2 <= len);
  }
  /* frame and scalar state */
  const float sg = un < 0.f ? -1.f : 1.f;
  const float ex = __fmaf_rn(bh.f_ch, fr.xx, bh.f_sh * fr.yx), ey = __fmaf_rn(bh.f_ch, fr.xy, bh.f_sh * fr.yy), ez = __fmaf_rn(bh.f_ch, fr.xz, bh.f_sh * fr.yz);
  fr.yx = __fmaf_rn(bh.f_ch, fr.yx, -bh.f_sh * fr.xx); fr.yy = __fmaf_rn(bh.f_ch, fr.yy, -bh.f_sh * fr.xy); fr.yz = __fmaf_rn(bh.f_ch, fr.yz, -bh.f_sh * fr.xz);
  fr.xx = sg * ex; fr.xy = sg * ey; fr.xz = sg * ez;
  fr.up = __fmaf_rn(G, bh.f_cot, -u * bh.f_tan2);
  fr.u = fabsf(un);
  return captured;
}

/* Conservative FP32 version of BBox::intersect (reference src/bbox.cpp:10-25).  The reference evaluates
 * t = (plane - o) / d in double for the six planes; here one chord-constant context holds o and 1/d rounded to float
 * and every plane parameter is widened by a rigorous bound on what those roundings can do, so that every box the
 * reference accepts is accepted here (the reverse need not hold: the exact primitive tests decide the hit):
 *   o_f = fl(o), i_f ~ 1 / fl(d) (correctly rounded, or within 1 ulp: RRT_BOX_APPROX_RCP); the per-plane formula and its error bound
 *   are with boxctx_axis below.
 * Direction components below 1e-30 are flushed to a signed zero (i_f = +-inf; the constants become inf or NaN), which
 * removes the float-denormal case from the analysis; NaN parameters are dropped by fminf/fmaxf, which only ever
 * widens the interval.  Node boxes are already rounded outward from the reference's doubles on the host.
 * Requires tmin >= 0 (true for chords and for every ray the estimator makes). */
#ifndef RRT_WIDE_RECOMPUTE_OFFSETS
#define RRT_WIDE_RECOMPUTE_OFFSETS 1   /* wide-node plane-array offsets from the three sign flags at every visit instead of six live registers: the kernels sit at
                                          their register cap, and what this frees comes out of the spills (closest-hit 256 -> 64 B, shadow 148 -> 134 B).
                                          B200, 0 -> 1: C2 35.38 -> 34.53 ms, C1 4.31 -> 4.25, C2 with the FP32 stepper 32.76 -> 30.91 */
#endif
#ifndef RRT_BOX_APPROX_RCP
#define RRT_BOX_APPROX_RCP 1   /* 1 = MUFU.RCP alone (rcp.approx.ftz, <= 1 ulp = 2u) for the box context's three reciprocals instead of the
                                  correctly rounded __frcp_rn.  i_f multiplies BOTH terms of t_f = b i_f - o_f i_f, so its error is a common
                                  relative factor of every plane parameter of the axis: the relative bound below grows from 3u to 4u of the
                                  16u (2^-20) the test grants; the absolute term A is untouched.  Measured on B200 (0 -> 1): C2 36.52 ->
                                  35.54 ms, bunny 1080p 103.9 -> 100.6, C5 68.6 -> 66.4, C1 4.40 -> 4.36; results unchanged (parity suite). */
#endif
#if RRT_BOX_APPROX_RCP
#define RRT_BOX_RCP(x) fast_rcp(x)
#else
#define RRT_BOX_RCP(x) __frcp_rn(x)
#endif
struct BoxCtx {
  float ix, iy, iz;          /* fl(1 / fl(d)) */
  float nx_, ny_, nz_;       /* near-plane constants:  -(fl(o_f * i_f)) - A */
  float fx_, fy_, fz_;       /* far-plane constants:   -(fl(o_f * i_f)) + A */
  float tmin;
  bool nx, ny, nz;           /* direction component negative: the near plane is the box maximum */
};
#ifndef RRT_BOX_SIGN_FROM_INV
#define RRT_BOX_SIGN_FROM_INV 1   /* "direction component negative" read off the sign bit of i_f (the sign of d_f, -0 -> -inf included; a NaN axis is dropped by
                                     fminf / fmaxf whichever way it is read) instead of three flags kept live: wide shadow kernel spill stores 134 -> 4 B.
                                     B200, 0 -> 1: C2 34.50 -> 33.54 ms, C1 4.20 -> 4.12, bunny 1080p 100.1 -> 98.2, C5 66.0 -> 64.6 */
#endif
#if RRT_BOX_SIGN_FROM_INV
#define RRT_NEGX(c) (__float_as_int((c).ix) < 0)
#define RRT_NEGY(c) (__float_as_int((c).iy) < 0)
#define RRT_NEGZ(c) (__float_as_int((c).iz) < 0)
#else
#define RRT_NEGX(c) ((c).nx)
#define RRT_NEGY(c) ((c).ny)
#define RRT_NEGZ(c) ((c).nz)
#endif
/* Plane parameter in one FFMA: t_f = fl(b * i_f - c), c = fl(o_f * i_f).  Against the exact t = (b - o) / d:
 *   |t - t_f| <= |t_f| * 3u  +  (|o - o_f| |1/d| + |c - o_f i_f|) (1 + 3u)  <=  |t_f| * 3u + 2u |o_f| |i_f| (1 + ...)
 * (u = 2^-24), widened to |t_f| * 2^-20 + A with A = 2^-22 |o_f| |i_f|: 2x slack on the absolute term, 5x on the
 * relative one.  With i_f from MUFU.RCP alone (RRT_BOX_APPROX_RCP: |i_f d_f - 1| <= 2u instead of u) the factor common to
 * b i_f and o_f i_f is off by one more u: relative term 4u (+ second order) of the 16u granted, absolute term as before.
 * `slack` is an extra bound on |o - o_true| (far chords, below); it enters A as slack |i_f|. */
__device__ __forceinline__ void boxctx_axis(double o, double d, float slack, float& inv, float& cn, float& cf, bool& neg) {
  float df = (float)d;
  if (fabsf(df) < 1e-30f) df = copysignf(0.f, df);
  float of = (float)o;
  inv = RRT_BOX_RCP(df);
  float c = of * inv;
  float a = __fmaf_rn(2.3841858e-7f, fabsf(of), slack) * fabsf(inv);
  cn = -c - a; cf = -c + a;
  neg = __float_as_int(df) < 0;
}
/* Far chords.  The position uncertainty of the FP32 test is ~2^-20 |o|.  The reference's escaped rays jump between
 * points up to ~1e7 units out on chords that pass close to the scene; from there every box near
 * the chord would pass and one lane would walk the whole tree.  The slab test does not care where along the line the
 * parameter is counted from, so such a chord is tested from q = o + T d with the range [tmin - T, tmax - T], T = the
 * distance from o to the sphere around the root box.  q is computed in double; the bound on its rounding error,
 * |q_i - (o_i + T d_i)| <= 2^-53 (|o_i| + 2T), enters the plane constants as `slack` (4x margin), and the shifted
 * range ends are widened by 2^-50 relative before rounding outward to float.  Returns T (0: not shifted). */
#define RRT_FAR_ORIGIN 512.0f
__device__ __forceinline__ double boxctx_make(const DevScene& sc, const d3& o, const d3& d, double tmin, BoxCtx& c) {
  d3 q = o;
  double T = 0.0;
  float slack = 0.f;
  const float m = fmaxf(fmaxf(fabsf((float)o.x), fabsf((float)o.y)), fabsf((float)o.z));
  c.tmin = __double2float_rd(tmin);
  if (m > RRT_FAR_ORIGIN && m < 1e30f) {
    d3 w = vsub(o, D3(sc.ctr[0], sc.ctr[1], sc.ctr[2]));
    T = fmax(sqrt(vnorm2(w)) - sc.rad, 0.0);
    q = D3(o.x + T * d.x, o.y + T * d.y, o.z + T * d.z);
    slack = __double2float_ru((3.0 * (double)m + T) * 4.45e-16);
    double x = tmin - T;
    c.tmin = __double2float_rd(x - fabs(x) * 8.9e-16);
  }
  boxctx_axis(q.x, d.x, slack, c.ix, c.nx_, c.fx_, c.nx);
  boxctx_axis(q.y, d.y, slack, c.iy, c.ny_, c.fy_, c.ny);
  boxctx_axis(q.z, d.z, slack, c.iz, c.nz_, c.fz_, c.nz);
  return T;
}
/* upper end of the parameter range in the (possibly shifted) units of the box context, rounded up */
__device__ __forceinline__ float boxctx_tmax(double tmax, double T) {
  if (T == 0.0) return __double2float_ru(tmax);
  double x = tmax - T;
  return __double2float_ru(x + fabs(x) * 8.9e-16);
}
__device__ __forceinline__ bool box_hit(float lox, float loy, float loz, float hix, float hiy, float hiz, const BoxCtx& c, float tmaxf,
                                        float& tnear) {
  const float R = 9.5367431640625e-7f; /* 2^-20 */
  float tnx = __fmaf_rn(RRT_NEGX(c) ? hix : lox, c.ix, c.nx_), tfx = __fmaf_rn(RRT_NEGX(c) ? lox : hix, c.ix, c.fx_);
  float tny = __fmaf_rn(RRT_NEGY(c) ? hiy : loy, c.iy, c.ny_), tfy = __fmaf_rn(RRT_NEGY(c) ? loy : hiy, c.iy, c.fy_);
  float tnz = __fmaf_rn(RRT_NEGZ(c) ? hiz : loz, c.iz, c.nz_), tfz = __fmaf_rn(RRT_NEGZ(c) ? loz : hiz, c.iz, c.fz_);
  float tn = fmaxf(fmaxf(tnx, tny), tnz), tf = fminf(fminf(tfx, tfy), tfz);
  tn = __fmaf_rn(fabsf(tn), -R, tn);
  tf = __fmaf_rn(fabsf(tf), R, tf);
  tnear = tn;
  return tn <= tf && tn <= tmaxf && tf >= c.tmin;
}

/* Chord as the FP32 stepper produces it (RRT_OPT_STEPPER = 1): float origin, unit direction (to 1e-7) and length, plus
 * e = (exact origin) - (float origin): non-zero only for a ray's first chord, which must leave from the caller's own
 * double origin (rays start ON surfaces, hit_p + 1e-11 w_in, reference src/part1_code.cpp:45,88: an origin rounded to
 * float lands behind the surface half the time).  The traversals below are generic over Seg / SegF through these adapters:
 * box tests read the floats directly (no double -> float conversions), the exact leaf tests get doubles on demand. */
struct SegF { float ox, oy, oz, dx, dy, dz, tmax, ex, ey, ez; };
__device__ __forceinline__ const Seg& seg_exact(const Seg& s) { return s; }
__device__ __forceinline__ Seg seg_exact(const SegF& s) {
  Seg r;
  r.o = D3((double)s.ox + (double)s.ex, (double)s.oy + (double)s.ey, (double)s.oz + (double)s.ez);
  r.d = D3((double)s.dx, (double)s.dy, (double)s.dz); r.tmin = 0.0; r.tmax = (double)s.tmax;
  return r;
}
/* Chord kept in shared memory (RRT_SEG_SMEM, rrt_wavetrace.cuh): 7 doubles per thread, stride RRT_SEG_STRIDE doubles; the traversal reads them
 * only where it needs them (box context, exact leaf tests), so they occupy no registers while nodes are visited. */
#ifndef RRT_SEG_STRIDE
#define RRT_SEG_STRIDE 128
#endif
struct SegS { const double* p; };
__device__ __forceinline__ Seg seg_exact(const SegS& s) {
  Seg r;
  r.o = D3(s.p[0], s.p[RRT_SEG_STRIDE], s.p[2 * RRT_SEG_STRIDE]); r.d = D3(s.p[3 * RRT_SEG_STRIDE], s.p[4 * RRT_SEG_STRIDE], s.p[5 * RRT_SEG_STRIDE]);
  r.tmin = 0.0; r.tmax = s.p[6 * RRT_SEG_STRIDE];
  return r;
}
__device__ __forceinline__ double seg_tmax(const SegS& s) { return s.p[6 * RRT_SEG_STRIDE]; }
__device__ __forceinline__ double seg_tmax(const Seg& s) { return s.tmax; }
__device__ __forceinline__ double seg_tmax(const SegF& s) { return (double)s.tmax; }
/* box context + float upper end of the range; returns whether the far-chord origin shift is in effect */
__device__ __forceinline__ bool seg_boxctx(const DevScene& sc, const Seg& s, BoxCtx& bc, float& tmaxf) {
  const double T = boxctx_make(sc, s.o, s.d, s.tmin, bc);
  tmaxf = boxctx_tmax(s.tmax, T);
  return T != 0.0;
}
__device__ __forceinline__ bool seg_boxctx(const DevScene& sc, const SegS& s, BoxCtx& bc, float& tmaxf) {
  const Seg e = seg_exact(s);
  const double T = boxctx_make(sc, e.o, e.d, 0.0, bc);
  tmaxf = boxctx_tmax(e.tmax, T);
  return T != 0.0;
}
__device__ __forceinline__ void boxctx_axis_f(float of, float df, float slack, float& inv, float& cn, float& cf, bool& neg) {
  if (fabsf(df) < 1e-30f) df = copysignf(0.f, df);
  inv = RRT_BOX_RCP(df);
  float c = of * inv;
  float a = __fmaf_rn(2.3841858e-7f, fabsf(of), slack) * fabsf(inv);   /* same bound as boxctx_axis; here o_f and d_f ARE the chord */
  cn = -c - a; cf = -c + a;
  neg = __float_as_int(df) < 0;
}
__device__ __forceinline__ bool seg_boxctx(const DevScene& sc, const SegF& s, BoxCtx& bc, float& tmaxf) {
  const float m = fmaxf(fmaxf(fabsf(s.ox), fabsf(s.oy)), fabsf(s.oz));
  if (m > RRT_FAR_ORIGIN && m < 1e30f) {   /* far chord: rare, take the double path */
    const Seg e = seg_exact(s);
    const double T = boxctx_make(sc, e.o, e.d, 0.0, bc);
    tmaxf = boxctx_tmax(e.tmax, T);
    return T != 0.0;
  }
  const float slack = fmaxf(fmaxf(fabsf(s.ex), fabsf(s.ey)), fabsf(s.ez));   /* |exact origin - float origin| */
  bc.tmin = 0.f;
  boxctx_axis_f(s.ox, s.dx, slack, bc.ix, bc.nx_, bc.fx_, bc.nx);
  boxctx_axis_f(s.oy, s.dy, slack, bc.iy, bc.ny_, bc.fy_, bc.ny);
  boxctx_axis_f(s.oz, s.dz, slack, bc.iz, bc.nz_, bc.fz_, bc.nz);
  tmaxf = s.tmax;
  return false;
}

/* True when x * (1 / det) is certainly a negative, non-zero number: opposite signs and both magnitudes far enough
 * from the ends of the exponent range that the product can neither underflow to (-)0 nor involve inf/NaN.  Used to
 * reject a triangle before the division; anything this test does not cover takes the literal path below. */
__device__ __forceinline__ bool surely_negative(double x, double det) {
  int hx = __double2hiint(x), hd = __double2hiint(det);
  unsigned ex = ((unsigned)hx >> 20) & 0x7ffu, ed = ((unsigned)hd >> 20) & 0x7ffu;
  return ((hx ^ hd) < 0) && (ex - 523u < 1524u) && (ed - 1u < 1523u);
}

/* Empty-space test: true when no primitive can lie within `len` of p, so BVHAccel::intersect_micro would return false
 * for any chord of that length starting at p and the traversal can be skipped without changing the result.  The cell
 * is located in double (6 FP64 operations per chord); points outside the grid and NaNs take the normal path. */
__device__ __forceinline__ bool chord_is_clear(const DevScene& sc, const d3& p, double len) {
  if (!sc.clearance) return false;
  double fx = (p.x - sc.glo[0]) * sc.ginv[0], fy = (p.y - sc.glo[1]) * sc.ginv[1], fz = (p.z - sc.glo[2]) * sc.ginv[2];
  if (!(fx >= 0.0 && fy >= 0.0 && fz >= 0.0 && fx < (double)sc.gnx && fy < (double)sc.gny && fz < (double)sc.gnz)) return false;
  int ix = (int)fx, iy = (int)fy, iz = (int)fz;
  float c = __ldg(sc.clearance + ((size_t)iz * sc.gny + iy) * sc.gnx + ix);
  return len < (double)c;
}

/* Triangle::intersect and Sphere::intersect on one 80-byte leaf record.  Accepts when the reference would and the
 * hit is at least as close as the best so far (ties: higher depth-first rank wins). */
template <bool COUNT, bool UNITD = true>
__device__ __forceinline__ bool leaf_record_test(const double2* __restrict__ rec, const Seg& s, Best& best, Counters& cnt) {
  double2 a0 = __ldg(rec + 0), a1 = __ldg(rec + 1), a2 = __ldg(rec + 2), a3 = __ldg(rec + 3), a4 = __ldg(rec + 4);
  unsigned long long tag = (unsigned long long)__double_as_longlong(a4.y);
  int prim = (int)(uint32_t)tag;
  uint32_t rk = (uint32_t)(tag >> 32);
  int kind = (int)(rk & 1u), rank = (int)(rk >> 1);
  double t, b1 = 0.0, b2 = 0.0;
  if (kind == 0) {
    if (COUNT) cnt.tris++;
    d3 p0 = D3(a0.x, a0.y, a1.x), e1 = D3(a1.y, a2.x, a2.y), e2 = D3(a3.x, a3.y, a4.x);
    /* same products as the reference; the three sign tests only skip work whose outcome is already decided */
    d3 s1 = vcross(s.d, e2);
    double det = vdot(s1, e1);
    d3 sv = vsub(s.o, p0);
    double x1 = vdot(s1, sv);
    if (surely_negative(x1, det)) return false;   /* b1 < 0 */
    d3 s2 = vcross(sv, e1);
    double x2 = vdot(s2, s.d);
    if (surely_negative(x2, det)) return false;   /* b2 < 0 */
    double xt = vdot(s2, e2);
    if (surely_negative(xt, det)) return false;   /* t < 0 <= min_t */
    double inv = 1. / det;
    t = xt * inv; b1 = x1 * inv; b2 = x2 * inv;
    double b0 = 1 - b1 - b2;
    if (!(s.tmin <= t && t <= best.t && b0 >= 0 && b1 >= 0 && b2 >= 0)) return false;
  } else {
    if (COUNT) cnt.sph++;
    d3 tmp = vsub(s.o, D3(a0.x, a0.y, a1.x));
    double b = 2 * vdot(tmp, s.d), c = vnorm2(tmp) - a1.y, t1, t2;
    if (UNITD) {   /* the reference's form: a = 1 assumed (src/static_scene/sphere.cpp:10-20) */
      double disc = b * b - 4 * c;
      if (disc < 0) return false;
      double sq = sqrt(disc);
      t1 = (-b - sq) / 2; t2 = (-b + sq) / 2;
    } else {       /* FP32 stepper: |d| = 1 +- 1e-7, and a root taken with a = 1 would leave the hit point ~1e-8 off the sphere —
                      inside it half the time, where the next ray (origin + 1e-11 w_in) starts by hitting the sphere from within */
      double a = vnorm2(s.d), disc = b * b - 4 * a * c;
      if (disc < 0) return false;
      double sq = sqrt(disc), ia = 0.5 / a;
      t1 = (-b - sq) * ia; t2 = (-b + sq) * ia;
    }
    if (s.tmin <= t1 && t1 <= best.t) t = t1;
    else if (s.tmin <= t2 && t2 <= best.t) t = t2;
    else return false;
  }
  if (t == best.t && rank < best.rank) return false;
  best.t = t; best.b1 = b1; best.b2 = b2; best.prim = prim; best.rank = rank; best.kind = kind;
  return true;
}

/* Cooperative, deferred leaf tests for any-hit (shadow) chords.
 * The exact primitive tests run on the FP64 pipe, which is what bounds this kernel (ncu: ~0.2 FP64 warp instructions
 * per scheduler cycle, the most the pipe sustains), and a warp instruction costs the pipe the same whether 2 or 32 of
 * its lanes are active.  Leaf boxes are hit by few lanes at a time (2.3 on average on C2), so "every lane tests its
 * own leaf" wasted 90 % of each FP64 issue slot spent there: 36 % of the kernel's FP64 instructions.
 * An any-hit query does not need its leaves tested in order, or at once.  So a lane that reaches a leaf only APPENDS
 * (owner lane, record) pairs to a per-warp list in shared memory and keeps descending; when the list holds a warp's
 * worth of pairs, or nobody has nodes left, the whole warp tests the list 32 pairs at a time, each lane fetching its
 * pair's chord from the owner's registers by shuffle, and raises the owner's bit on an accept.  The arithmetic of
 * each test is the one the owner would have done (same record, same chord, same bound), so verdicts are unchanged; the
 * price is that an occluded chord may visit a few more nodes before it learns of its hit. */
#ifndef RRT_COOP_FLUSH
#define RRT_COOP_FLUSH 32                      /* test the list once it holds this many pairs */
#endif
#define RRT_COOP_CAP (RRT_COOP_FLUSH + 224)    /* a lane appends at most one leaf (<= 7 records) after the threshold is reached */
struct CoopScratch {            /* one per warp, shared memory */
  uint32_t list[RRT_COOP_CAP];  /* (owner lane << 27) | leaf record index */
  int n;                        /* pairs in the list; 0 between traversals */
  uint32_t hitmask;             /* bit per owner lane; 0 between traversals */
  int pad[2];
};
__device__ __forceinline__ void coop_init(CoopScratch* cs) {
  if ((threadIdx.x & 31) == 0) { cs->n = 0; cs->hitmask = 0u; }
  __syncwarp();
}
/* any lane, divergent: queue the records of leaf code `cur`; true = the list is due for testing, stop descending */
__device__ __forceinline__ bool coop_append(CoopScratch* cs, int cur) {
  const uint32_t bits = (uint32_t)~cur, first = bits >> 3, count = bits & 7u;
  const int pos = atomicAdd(&cs->n, (int)count);
  const uint32_t tag = (uint32_t)(threadIdx.x & 31) << 27;
  for (uint32_t k = 0; k < count; ++k) cs->list[pos + k] = tag | (first + k);
  return pos + (int)count >= RRT_COOP_FLUSH;
}
/* all 32 lanes, converged, after a __syncwarp: test the n queued pairs; returns whether one of MY records accepted */
template <bool COUNT>
__device__ __forceinline__ bool coop_flush(const DevScene& sc, CoopScratch* cs, const Seg& s, int n, Counters& cnt) {
  const int lane = threadIdx.x & 31;
  for (int base = 0; base < n; base += 32) {
    const int e = base + lane;
    const uint32_t pr = e < n ? cs->list[e] : ((uint32_t)lane << 27);
    const int owner = (int)(pr >> 27);
    Seg s2;   /* the owner's chord, straight from its registers */
    s2.o = D3(__shfl_sync(0xffffffffu, s.o.x, owner), __shfl_sync(0xffffffffu, s.o.y, owner), __shfl_sync(0xffffffffu, s.o.z, owner));
    s2.d = D3(__shfl_sync(0xffffffffu, s.d.x, owner), __shfl_sync(0xffffffffu, s.d.y, owner), __shfl_sync(0xffffffffu, s.d.z, owner));
    s2.tmin = __shfl_sync(0xffffffffu, s.tmin, owner); s2.tmax = __shfl_sync(0xffffffffu, s.tmax, owner);
    if (e < n) {
      Best b2; b2.t = s2.tmax; b2.rank = -1; b2.prim = -1; b2.kind = 0; b2.b1 = b2.b2 = 0.0;
      if (leaf_record_test<COUNT>(sc.leaf + 5 * (size_t)(pr & 0x7ffffffu), s2, b2, cnt)) atomicOr(&cs->hitmask, 1u << owner);
    }
  }
  __syncwarp();
  const bool mine = ((*(volatile uint32_t*)&cs->hitmask) >> lane) & 1u;
  __syncwarp();
  if (lane == 0) { cs->n = 0; cs->hitmask = 0u; }
  __syncwarp();
  return mine;
}

/* BVHAccel::intersect_micro (reference src/bvh.cpp:115-138) as an iterative, near-child-first traversal with a
 * per-thread stack in shared memory.  ANY = shadow-ray form (stop at the first accepted primitive). */
template <bool ANY, bool COUNT, bool COOP = false, bool UNITD = true, class SEG = Seg>
__device__ __forceinline__ bool traverse(const DevScene& sc, const SEG& s, Best& best, int* __restrict__ stack, int stride, Counters& cnt, bool active,
                                         CoopScratch* cs = nullptr, bool* root_miss = nullptr) {
  /* Called by ALL 32 lanes of the warp, converged; lanes without a chord pass active = false and idle through.
   * while-while: every lane first descends until it holds a leaf (or runs out of nodes), the warp re-converges, then
   * the lanes that hold a leaf test its records.  (Measured alternatives, C2 on B200, ms per frame: single if/else
   * loop 40.3; this form 38.2-38.9; remembering the leaves and testing them after the whole descent 41.3.) */
  const int DONE = (int)0x80000000;   /* not a valid leaf code: n_prims < 2^28 */
  best.t = seg_tmax(s); best.rank = -1; best.prim = -1; best.kind = 0; best.b1 = best.b2 = 0.0;
  BoxCtx bc;
  float tmaxf = 0.f, tl, tr;
  bool found = false;
  int cur = DONE, sp = 0;
  bool shifted = false;               /* far chord: box parameters count from o + T d, see boxctx_make */
  if (active) {
    shifted = seg_boxctx(sc, s, bc, tmaxf);
    if (box_hit(sc.root_lo[0], sc.root_lo[1], sc.root_lo[2], sc.root_hi[0], sc.root_hi[1], sc.root_hi[2], bc, tmaxf, tl)) cur = sc.root;
    else if (root_miss) *root_miss = true;   /* the chord does not reach the scene box at all */
  }
  /* Top of the tree as one 8-wide record.  The boxes near the root of a Cornell-box scene are the groupings of its walls: each spans most of
   * the room, so a chord anywhere inside used to walk 5-6 binary nodes only to learn that it touches no wall (6.0 node visits per chord on the
   * bunny scene, 56 % of the trace kernel's instructions, profiles/r02_f_attribution_c4_shadow.txt).  rrt_upload_scene opens those nodes
   * (largest drop in surface area first) into up to 8 children — the flat per-wall boxes and the root of the mesh — and one straight-line visit
   * tests all of them; what passes goes on the stack and the binary loop below carries on.  Same boxes, same leaves: results unchanged. */
  if (sc.top != nullptr) {
    const bool need_top = cur != DONE && cur >= 0;
#if RRT_TOP_AABB
    /* Short chords (everything inside the scene: a few centimetres) take a cheaper test of the same record: the chord's own bounding box against
     * the child boxes, 6 compares per child instead of 6 FFMA + 2 three-way min / max + 2 FFMA + 3 compares.  Conservative like the slab test (a
     * chord that meets a box has a bounding box that overlaps it; the box is widened by the rounding of its corners), exact for the walls (a
     * full-wall rectangle: every chord inside the room that reaches the wall's plane does so within it), and the children it lets through are
     * tested exactly further down.  The choice is made per WARP (all of its chords short, none far-shifted), so there is one code path per trip. */
    bool aabb_ok = true;
    if constexpr (!std::is_same<SEG, SegF>::value) aabb_ok = !need_top || (!shifted && tmaxf <= RRT_TOP_AABB_MAXLEN && bc.tmin <= 0.f);
    else aabb_ok = !need_top;
    if (__all_sync(0xffffffffu, aabb_ok)) {
      if constexpr (!std::is_same<SEG, SegF>::value) {
        if (need_top) {
          const Seg& e = seg_exact(s);
          const float ox = (float)e.o.x, oy = (float)e.o.y, oz = (float)e.o.z;
          const float ex = __fmaf_rn((float)e.d.x, tmaxf, ox), ey = __fmaf_rn((float)e.d.y, tmaxf, oy), ez = __fmaf_rn((float)e.d.z, tmaxf, oz);
          /* corners of the chord's box: |fl(o) - o| <= 2^-24 |o|, the far end lies between o and o + d tmaxf (tmaxf >= len) up to
           * 2^-24 (|o| + 2 tmaxf + |end|); 2^-21 (...) covers both with an 8x margin */
          const float mx = 4.7683716e-7f * (fabsf(ox) + fabsf(ex) + tmaxf), my = 4.7683716e-7f * (fabsf(oy) + fabsf(ey) + tmaxf),
                      mz = 4.7683716e-7f * (fabsf(oz) + fabsf(ez) + tmaxf);
          const float lx = fminf(ox, ex) - mx, hx = fmaxf(ox, ex) + mx, ly = fminf(oy, ey) - my, hy = fmaxf(oy, ey) + my,
                      lz = fminf(oz, ez) - mz, hz = fmaxf(oz, ez) + mz;
          const float* nd = sc.top;
          if (COUNT) cnt.nodes += 4;
          cur = DONE;
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            const float4 a = __ldg((const float4*)(nd + 0) + g), b = __ldg((const float4*)(nd + 8) + g), c = __ldg((const float4*)(nd + 16) + g);
            const float4 d = __ldg((const float4*)(nd + 24) + g), f = __ldg((const float4*)(nd + 32) + g), h = __ldg((const float4*)(nd + 40) + g);
            const int4 k = __ldg((const int4*)(nd + 48) + g);
            auto child = [&](float blx, float bly, float blz, float bhx, float bhy, float bhz, int code) {
              if (lx <= bhx && hx >= blx && ly <= bhy && hy >= bly && lz <= bhz && hz >= blz && code != DONE) {
                if (cur != DONE) stack[(sp++) * stride] = cur;
                cur = code;
              }
            };
            child(a.x, b.x, c.x, d.x, f.x, h.x, k.x);
            child(a.y, b.y, c.y, d.y, f.y, h.y, k.y);
            child(a.z, b.z, c.z, d.z, f.z, h.z, k.z);
            child(a.w, b.w, c.w, d.w, f.w, h.w, k.w);
          }
        }
      }
    } else
#endif
    if (need_top) {
      const float R = 9.5367431640625e-7f;
      const float* nd = sc.top;
      if (COUNT) cnt.nodes += 4;     /* a 256-byte record */
      cur = DONE;
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        const float4 a = __ldg((const float4*)(nd + (RRT_NEGX(bc) ? 24 : 0)) + g), b = __ldg((const float4*)(nd + (RRT_NEGX(bc) ? 0 : 24)) + g);
        const float4 c = __ldg((const float4*)(nd + (RRT_NEGY(bc) ? 32 : 8)) + g), d = __ldg((const float4*)(nd + (RRT_NEGY(bc) ? 8 : 32)) + g);
        const float4 e = __ldg((const float4*)(nd + (RRT_NEGZ(bc) ? 40 : 16)) + g), f = __ldg((const float4*)(nd + (RRT_NEGZ(bc) ? 16 : 40)) + g);
        const int4 k = __ldg((const int4*)(nd + 48) + g);
        auto child = [&](float pnx, float pfx, float pny, float pfy, float pnz, float pfz, int code) {
          float tn = fmaxf(fmaxf(__fmaf_rn(pnx, bc.ix, bc.nx_), __fmaf_rn(pny, bc.iy, bc.ny_)), __fmaf_rn(pnz, bc.iz, bc.nz_));
          float tf = fminf(fminf(__fmaf_rn(pfx, bc.ix, bc.fx_), __fmaf_rn(pfy, bc.iy, bc.fy_)), __fmaf_rn(pfz, bc.iz, bc.fz_));
          tn = __fmaf_rn(fabsf(tn), -R, tn);
          tf = __fmaf_rn(fabsf(tf), R, tf);
          if (tn <= tf && tn <= tmaxf && tf >= bc.tmin && code != DONE) {
            if (cur != DONE) stack[(sp++) * stride] = cur;
            cur = code;
          }
        };
        child(a.x, b.x, c.x, d.x, e.x, f.x, k.x);
        child(a.y, b.y, c.y, d.y, e.y, f.y, k.y);
        child(a.z, b.z, c.z, d.z, e.z, f.z, k.z);
        child(a.w, b.w, c.w, d.w, e.w, f.w, k.w);
      }
    }
  }
  auto visit = [&]() {
    const float4* nd = sc.nodes + 4 * (size_t)cur;
    float4 q0 = __ldg(nd), q1 = __ldg(nd + 1), q2 = __ldg(nd + 2), q3 = __ldg(nd + 3);
    if (COUNT) cnt.nodes++;
    bool hl = box_hit(q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, bc, tmaxf, tl);
    bool hr = box_hit(q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, bc, tmaxf, tr);
    int cl = __float_as_int(q3.x), cr = __float_as_int(q3.y);
    if (hl && hr) {
      if (tr < tl) { int t = cl; cl = cr; cr = t; }
      stack[(sp++) * stride] = cr;
      cur = cl;
    } else if (hl) cur = cl;
    else if (hr) cur = cr;
    else cur = sp ? stack[(--sp) * stride] : DONE;
  };
  if constexpr (COOP) {
    /* any-hit with deferred, warp-shared leaf tests (coop_* above) */
    while (true) {
      while (cur != DONE) {
        if (cur >= 0) visit();
        else {
          const bool due = coop_append(cs, cur);
          cur = sp ? stack[(--sp) * stride] : DONE;
          if (due) break;
        }
      }
      __syncwarp();
      const int n = *(volatile int*)&cs->n;
      if (n == 0) break;                       /* nothing queued => every lane ran out of nodes */
      if (coop_flush<COUNT>(sc, cs, s, n, cnt)) { found = true; best.prim = 0; sp = 0; cur = DONE; }
    }
    return found;
  }
  while (__any_sync(0xffffffffu, cur != DONE)) {
    while (cur >= 0) visit();
    __syncwarp();
    if (cur != DONE) {
      uint32_t bits = (uint32_t)~cur;
      uint32_t first = bits >> 3, count = bits & 7u;
      cur = DONE;
      for (uint32_t k = 0; k < count; ++k) {
        if (leaf_record_test<COUNT, UNITD>(sc.leaf + 5 * (size_t)(first + k), seg_exact(s), best, cnt)) {
          found = true;
          if (ANY) { sp = 0; break; }
          if (!shifted) tmaxf = __double2float_ru(best.t);   /* shifted chords keep the chord-long range: rare, and conservative */
        }
      }
      if (sp) cur = stack[(--sp) * stride];
    }
    __syncwarp();
  }
  return found;
}

/* The same traversal over WIDE nodes (rrt_bvh.h: the binary tree collapsed to up to W children per node, boxes stored
 * as six arrays of W floats).  One node visit tests W boxes with one straight-line instruction stream: no per-box
 * select (the near / far plane arrays are picked once per chord from the direction signs), no ordering network (the
 * nearest hit child is kept in a register, the others are pushed as they come), and far fewer loop iterations, which
 * is what keeps the 32 lanes of a warp together.  Shadow rays (ANY) push every hit child and pop: order is irrelevant.
 * A visit can push up to W - 1 codes, so the stack bound is (W - 1) * depth + 1: the first RRT_STACK entries live in
 * shared memory as before, anything beyond (never observed) in a local-memory array; rrt_upload_scene checks the bound. */
template <bool ANY, bool COUNT, int W, bool COOP = false, bool UNITD = true, class SEG = Seg>
__device__ __forceinline__ bool traverse_wide(const DevScene& sc, const SEG& s, Best& best, int* __restrict__ stack, int stride, Counters& cnt, bool active,
                                              CoopScratch* cs = nullptr, bool* root_miss = nullptr) {
  const int DONE = (int)0x80000000;   /* also the code of an unused child slot (kEmptyChild) */
  const float R = 9.5367431640625e-7f; /* 2^-20 */
  best.t = seg_tmax(s); best.rank = -1; best.prim = -1; best.kind = 0; best.b1 = best.b2 = 0.0;
  BoxCtx bc;
  float tmaxf = 0.f, tl;
  bool found = false, shifted = false;
  int cur = DONE, sp = 0;
  int spill[RRT_SPILL];
  auto push = [&](int code) { if (sp < RRT_STACK) stack[sp * stride] = code; else spill[min(sp - RRT_STACK, RRT_SPILL - 1)] = code; ++sp; };
  auto pop = [&]() -> int { --sp; return sp < RRT_STACK ? stack[sp * stride] : spill[min(sp - RRT_STACK, RRT_SPILL - 1)]; };
  int onx = 0, ony = W, onz = 2 * W, ofx = 3 * W, ofy = 4 * W, ofz = 5 * W;   /* float offsets of the near / far plane arrays */
  if (active) {
    shifted = seg_boxctx(sc, s, bc, tmaxf);
    if (bc.nx) { onx = 3 * W; ofx = 0; }
    if (bc.ny) { ony = 4 * W; ofy = W; }
    if (bc.nz) { onz = 5 * W; ofz = 2 * W; }
    if (box_hit(sc.root_lo[0], sc.root_lo[1], sc.root_lo[2], sc.root_hi[0], sc.root_hi[1], sc.root_hi[2], bc, tmaxf, tl)) cur = sc.wroot;
    else if (root_miss) *root_miss = true;
  }
  auto visit = [&]() {
    const float* nd = sc.wide + (size_t)(8 * W) * (size_t)cur;
    if (COUNT) cnt.nodes++;
    int nearest = DONE;
    float tbest = 0.f;
    auto child = [&](float pnx, float pfx, float pny, float pfy, float pnz, float pfz, int code) {
      float tn = fmaxf(fmaxf(__fmaf_rn(pnx, bc.ix, bc.nx_), __fmaf_rn(pny, bc.iy, bc.ny_)), __fmaf_rn(pnz, bc.iz, bc.nz_));
      float tf = fminf(fminf(__fmaf_rn(pfx, bc.ix, bc.fx_), __fmaf_rn(pfy, bc.iy, bc.fy_)), __fmaf_rn(pfz, bc.iz, bc.fz_));
      tn = __fmaf_rn(fabsf(tn), -R, tn);
      tf = __fmaf_rn(fabsf(tf), R, tf);
      if (tn <= tf && tn <= tmaxf && tf >= bc.tmin && code != DONE) {
        if (ANY) push(code);
        else if (nearest == DONE) { nearest = code; tbest = tn; }
        else if (tn < tbest) { push(nearest); nearest = code; tbest = tn; }
        else push(code);
      }
    };
#pragma unroll
    for (int g = 0; g < W / 4; ++g) {
#if RRT_WIDE_RECOMPUTE_OFFSETS
      const float4 a = __ldg((const float4*)(nd + (RRT_NEGX(bc) ? 3 * W : 0)) + g), b = __ldg((const float4*)(nd + (RRT_NEGX(bc) ? 0 : 3 * W)) + g);
      const float4 c = __ldg((const float4*)(nd + (RRT_NEGY(bc) ? 4 * W : W)) + g), d = __ldg((const float4*)(nd + (RRT_NEGY(bc) ? W : 4 * W)) + g);
      const float4 e = __ldg((const float4*)(nd + (RRT_NEGZ(bc) ? 5 * W : 2 * W)) + g), f = __ldg((const float4*)(nd + (RRT_NEGZ(bc) ? 2 * W : 5 * W)) + g);
#else
      const float4 a = __ldg((const float4*)(nd + onx) + g), b = __ldg((const float4*)(nd + ofx) + g);
      const float4 c = __ldg((const float4*)(nd + ony) + g), d = __ldg((const float4*)(nd + ofy) + g);
      const float4 e = __ldg((const float4*)(nd + onz) + g), f = __ldg((const float4*)(nd + ofz) + g);
#endif
      const int4 k = __ldg((const int4*)(nd + 6 * W) + g);
      child(a.x, b.x, c.x, d.x, e.x, f.x, k.x);
      child(a.y, b.y, c.y, d.y, e.y, f.y, k.y);
      child(a.z, b.z, c.z, d.z, e.z, f.z, k.z);
      child(a.w, b.w, c.w, d.w, e.w, f.w, k.w);
    }
    cur = (!ANY && nearest != DONE) ? nearest : (sp ? pop() : DONE);
  };
  if constexpr (COOP) {
    /* any-hit with deferred, warp-shared leaf tests (coop_* above) */
    while (true) {
      while (cur != DONE) {
        if (cur >= 0) visit();
        else {
          const bool due = coop_append(cs, cur);
          cur = sp ? pop() : DONE;
          if (due) break;
        }
      }
      __syncwarp();
      const int n = *(volatile int*)&cs->n;
      if (n == 0) break;                       /* nothing queued => every lane ran out of nodes */
      if (coop_flush<COUNT>(sc, cs, s, n, cnt)) { found = true; best.prim = 0; sp = 0; cur = DONE; }
    }
    return found;
  }
  while (__any_sync(0xffffffffu, cur != DONE)) {
    while (cur >= 0) visit();
    __syncwarp();
    if (cur != DONE) {
      uint32_t bits = (uint32_t)~cur;
      uint32_t first = bits >> 3, count = bits & 7u;
      cur = DONE;
      for (uint32_t k = 0; k < count; ++k) {
        if (leaf_record_test<COUNT, UNITD>(sc.leaf + 5 * (size_t)(first + k), seg_exact(s), best, cnt)) {
          found = true;
          if (ANY) { sp = 0; break; }
          if (!shifted) tmaxf = __double2float_ru(best.t);
        }
      }
      if (sp) cur = pop();
    }
    __syncwarp();
  }
  return found;
}
