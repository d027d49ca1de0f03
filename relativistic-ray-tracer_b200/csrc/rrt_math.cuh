// rrt_math.cuh — double/float vector helpers, Philox, for the sm_100a kernels.
//
// The geometry side of the hot path is IEEE double evaluated in exactly the operation order of the
// reference (x86-64 build without FMA, reference CMakeLists.txt:67-75): this translation unit is compiled
// with -fmad=false so that nvcc never contracts a*b+c, and every helper below spells out the order the
// reference's CGL::Vector3D operators use (reference CGL/include/CGL/vector3D.h:58-157 — note that
// operator/ and normalize() multiply by the reciprocal).  Double +,-,*,/ and sqrt are correctly rounded on
// the device, so chords, hit parameters and barycentrics come out bit-identical to the CPU path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define RRT_PI 3.14159265358979323 /* reference CGL/include/CGL/misc.h:11; same double as M_PI */
#define RRT_EPS_D 0.00000000001    /* reference CGL/include/CGL/misc.h:12 */
#define RRT_HD __host__ __device__ __forceinline__

struct d3 { double x, y, z; };
struct f3 { float r, g, b; };

RRT_HD d3 D3(double x, double y, double z) { d3 r; r.x = x; r.y = y; r.z = z; return r; }
RRT_HD d3 vadd(d3 a, d3 b) { return D3(a.x + b.x, a.y + b.y, a.z + b.z); }
RRT_HD d3 vsub(d3 a, d3 b) { return D3(a.x - b.x, a.y - b.y, a.z - b.z); }
RRT_HD d3 vneg(d3 a) { return D3(-a.x, -a.y, -a.z); }
RRT_HD d3 vmul(d3 a, double c) { return D3(a.x * c, a.y * c, a.z * c); }  /* v * c */
RRT_HD d3 smul(double c, d3 a) { return D3(c * a.x, c * a.y, c * a.z); }  /* c * v */
RRT_HD d3 vdiv(d3 a, double c) { double rc = 1.0 / c; return D3(rc * a.x, rc * a.y, rc * a.z); }
RRT_HD double vdot(d3 u, d3 v) { return u.x * v.x + u.y * v.y + u.z * v.z; }
RRT_HD d3 vcross(d3 u, d3 v) { return D3(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x); }
RRT_HD double vnorm2(d3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
RRT_HD double vnorm(d3 a) { return sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }
/* Vector3D::unit(): rNorm = 1/sqrt(..); (rNorm*x, ..) */
RRT_HD d3 vunit(d3 a) { double rn = 1. / sqrt(a.x * a.x + a.y * a.y + a.z * a.z); return D3(rn * a.x, rn * a.y, rn * a.z); }
/* Vector3D::normalize(): (*this) *= 1/norm() -> (x*rc, ..) */
RRT_HD d3 vnormalize(d3 a) { double rc = 1. / vnorm(a); return D3(a.x * rc, a.y * rc, a.z * rc); }

/* std::min / std::max with the reference's argument-order NaN behaviour (reference src/bbox.cpp:10-25) */
RRT_HD double cmin(double a, double b) { return (b < a) ? b : a; }
RRT_HD double cmax(double a, double b) { return (a < b) ? b : a; }

/* Spectrum (3 x float), reference CGL/include/CGL/spectrum.h */
RRT_HD f3 F3(float r, float g, float b) { f3 s; s.r = r; s.g = g; s.b = b; return s; }
RRT_HD f3 s_add(f3 a, f3 b) { return F3(a.r + b.r, a.g + b.g, a.b + b.b); }
RRT_HD f3 s_sub(f3 a, f3 b) { return F3(a.r - b.r, a.g - b.g, a.b - b.b); }
RRT_HD f3 s_mul(f3 a, f3 b) { return F3(a.r * b.r, a.g * b.g, a.b * b.b); }
RRT_HD f3 s_div(f3 a, f3 b) { return F3(a.r / b.r, a.g / b.g, a.b / b.b); }
RRT_HD f3 s_addf(f3 a, float s) { return F3(a.r + s, a.g + s, a.b + s); }
RRT_HD f3 s_mulf(f3 a, float s) { return F3(a.r * s, a.g * s, a.b * s); }
RRT_HD f3 s_divf(f3 a, float s) { return F3(a.r / s, a.g / s, a.b / s); }
RRT_HD float s_illum(f3 a) { return 0.2126f * a.r + 0.7152f * a.g + 0.0722f * a.b; }
RRT_HD f3 s3(const float* p) { return F3(p[0], p[1], p[2]); }

/* Matrix3x3 * Vector3D = x0*col0 + x1*col1 + x2*col2 (reference CGL/src/matrix3x3.cpp:124-128) */
RRT_HD d3 mat_mul(const d3 col[3], d3 x) { return vadd(vadd(smul(x.x, col[0]), smul(x.y, col[1])), smul(x.z, col[2])); }
/* M.T() * x (reference CGL/src/matrix3x3.cpp:130-141) */
RRT_HD d3 mat_tmul(const d3 col[3], d3 x) {
  return D3(x.x * col[0].x + x.y * col[0].y + x.z * col[0].z, x.x * col[1].x + x.y * col[1].y + x.z * col[1].z,
            x.x * col[2].x + x.y * col[2].y + x.z * col[2].z);
}

/* ---------------------------------------------------------------- Philox4x32-10 (Salmon et al., SC'11)
 * Draw n of path (pixel, sample) is word (n & 3) of philox(key = seed, counter = (pixel, sample, n >> 2, 0)),
 * mapped to the reference's rand()/RAND_MAX uniform as (word >> 1) / 2147483647.0 (range [0, 1] inclusive,
 * reference src/random_util.h:11-13).  The path carries only n; the block is recomputed when n crosses it. */
RRT_HD void philox4x32_10(uint32_t& c0, uint32_t& c1, uint32_t& c2, uint32_t& c3, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}

struct Rng {
  uint32_t pixel, sample, n, k0, k1;
  uint32_t blk;          /* block index currently held in w (0xFFFFFFFF = none) */
  uint32_t w0, w1, w2, w3;
  const double* fixed;   /* parity hooks: the caller supplies the uniforms in draw order */
};
RRT_HD Rng rng_make(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t n) {
  Rng g; g.pixel = pixel; g.sample = sample; g.n = n; g.k0 = (uint32_t)seed; g.k1 = (uint32_t)(seed >> 32);
  g.blk = 0xFFFFFFFFu; g.w0 = g.w1 = g.w2 = g.w3 = 0; g.fixed = nullptr; return g;
}
RRT_HD double rng_uniform(Rng& g) {
  if (g.fixed) return g.fixed[g.n++];
  uint32_t b = g.n >> 2;
  if (b != g.blk) {
    g.w0 = g.pixel; g.w1 = g.sample; g.w2 = b; g.w3 = 0;
    philox4x32_10(g.w0, g.w1, g.w2, g.w3, g.k0, g.k1);
    g.blk = b;
  }
  uint32_t k = g.n & 3, word = k == 0 ? g.w0 : k == 1 ? g.w1 : k == 2 ? g.w2 : g.w3;
  g.n++;
  return (double)(word >> 1) / 2147483647.0;
}
/* UniformGridSampler2D::get_sample (reference src/sampler.cpp:7-11): Vector2D(random_uniform(), random_uniform());
 * the reference build evaluates the constructor arguments right to left, so y takes the first draw. */
RRT_HD void rng_grid2(Rng& g, double& x, double& y) { y = rng_uniform(g); x = rng_uniform(g); }
