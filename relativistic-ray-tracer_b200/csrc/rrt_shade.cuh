// rrt_shade.cuh — device side of the estimator's local work: camera rays, the shading frame, BSDF::f / sample_f,
// SceneLight::sample_L.  Reference: src/part1_code.cpp:167-187, src/camera.cpp:176-184, src/bsdf.cpp:13-171,
// src/bsdf.h:20-50,165-173, src/sampler.cpp:7-56, src/static_scene/light.cpp:17-115.
//
// Geometry is double, radiance float, in the reference's operation order.  The transcendental functions
// (sin cos tan acos atan exp log erf) are CUDA's double-precision libm: correctly rounded to within 1-2 ulp,
// not bit-identical to glibc, which is why BSDF/light parity is stated to 1e-12 relative and not bitwise.
#pragma once
#include "rrt_math.cuh"
#include "../../include/rrt_b200.h"

__device__ __forceinline__ double radians_(double deg) { return deg * (RRT_PI / 180); }           /* reference CGL/include/CGL/misc.h:50-53 */
__device__ __forceinline__ double lerp_(double a, double b, double t) { return (1 - t) * a + t * b; } /* reference src/part1_code.cpp:177-180 */

struct DevCamera {  /* rrt_camera with tan(fov/2) formed once on the host would differ from the reference's per-call tan(): keep the angles */
  double pos[3], c2w[9], hFov, vFov, nClip, fClip, lensRadius, focalDistance;
};

/* Camera::generate_ray (reference src/part1_code.cpp:182-187) */
__device__ __forceinline__ void generate_ray(const DevCamera& cam, double x, double y, d3& o, d3& d) {
  double blx = -tan(radians_(cam.hFov) / 2), bly = -tan(radians_(cam.vFov) / 2);
  d3 col[3] = {D3(cam.c2w[0], cam.c2w[1], cam.c2w[2]), D3(cam.c2w[3], cam.c2w[4], cam.c2w[5]), D3(cam.c2w[6], cam.c2w[7], cam.c2w[8])};
  o = D3(cam.pos[0], cam.pos[1], cam.pos[2]);
  d = vunit(mat_mul(col, D3(lerp_(blx, -blx, x), lerp_(bly, -bly, y), -1)));
}
/* Camera::generate_ray_for_thin_lens (reference src/camera.cpp:176-184) */
__device__ __forceinline__ void generate_ray_thin_lens(const DevCamera& cam, double x, double y, double rndR, double rndTheta, d3& o, d3& d) {
  double blx = -tan(radians_(cam.hFov) / 2), bly = -tan(radians_(cam.vFov) / 2);
  d3 col[3] = {D3(cam.c2w[0], cam.c2w[1], cam.c2w[2]), D3(cam.c2w[3], cam.c2w[4], cam.c2w[5]), D3(cam.c2w[6], cam.c2w[7], cam.c2w[8])};
  d3 pinHole = D3(lerp_(blx, -blx, x), lerp_(bly, -bly, y), -1);
  double lr = cam.lensRadius * sqrt(rndR);
  d3 pLens = D3(lr * cos(rndTheta), lr * sin(rndTheta), 0);
  o = vadd(D3(cam.pos[0], cam.pos[1], cam.pos[2]), mat_mul(col, pLens));
  d = vunit(mat_mul(col, vsub(vmul(pinHole, cam.focalDistance), pLens)));
}

/* make_coord_space (reference src/bsdf.cpp:13-29): columns x, y, z of o2w */
__device__ __forceinline__ void make_coord_space(d3 o2w[3], d3 n) {
  d3 z = n, h = z;
  if (fabs(h.x) <= fabs(h.y) && fabs(h.x) <= fabs(h.z)) h.x = 1.0;
  else if (fabs(h.y) <= fabs(h.x) && fabs(h.y) <= fabs(h.z)) h.y = 1.0;
  else h.z = 1.0;
  z = vnormalize(z);
  d3 y = vnormalize(vcross(h, z));
  d3 x = vnormalize(vcross(z, y));
  o2w[0] = x; o2w[1] = y; o2w[2] = z;
}

__device__ __forceinline__ double clampd(double n, double lo, double hi) { return cmax(lo, cmin(n, hi)); } /* reference src/bsdf.h:20-22 */
__device__ __forceinline__ bool mat_is_delta(const rrt_material& m) { return m.type == RRT_MAT_MIRROR || m.type == RRT_MAT_GLASS || m.type == RRT_MAT_REFRACTION; }
__device__ __forceinline__ f3 mat_emission(const rrt_material& m) { return m.type == RRT_MAT_EMISSION ? s3(m.a) : F3(0, 0, 0); }

/* MicrofacetBSDF (reference src/bsdf.h:165-173, src/bsdf.cpp:45-72) */
__device__ __forceinline__ double mf_theta(d3 w) { return acos(clampd(w.z, -1.0 + 1e-5, 1.0 - 1e-5)); }
__device__ __forceinline__ double mf_lambda(const rrt_material& m, d3 w) {
  double theta = mf_theta(w);
  double a = 1.0 / (m.alpha * tan(theta));
  return 0.5 * (erf(a) - 1.0 + exp(-a * a) / (a * RRT_PI));
}
__device__ __forceinline__ double mf_G(const rrt_material& m, d3 wo, d3 wi) { return 1.0 / (1.0 + mf_lambda(m, wi) + mf_lambda(m, wo)); }
__device__ __forceinline__ double mf_D(const rrt_material& m, d3 h) {   /* Beckmann distribution of the half vector, reference src/bsdf.cpp:49-56 */
  double ang = mf_theta(h), tg = tan(ang), cz = h.z, cz2 = cz * cz,
         alpha2 = m.alpha * m.alpha; /* float product, as in the reference */
  return exp(-tg * tg / alpha2) / (RRT_PI * alpha2 * cz2 * cz2);
}
__device__ __forceinline__ f3 mf_F(const rrt_material& m, d3 wi) {       /* conductor Fresnel term per channel, reference src/bsdf.cpp:58-66 */
  f3 eta = s3(m.a), k = s3(m.b);
  f3 n2k2 = s_add(s_mul(eta, eta), s_mul(k, k));                         /* |eta + i k|^2 */
  double cz = wi.z, cz2 = cz * cz;
  f3 two_n_cos = s_mulf(s_mulf(eta, 2.f), (float)cz);
  f3 perp = s_div(s_addf(s_sub(n2k2, two_n_cos), (float)cz2), s_addf(s_add(n2k2, two_n_cos), (float)cz2));
  f3 para = s_div(s_addf(s_sub(s_mulf(n2k2, (float)cz2), two_n_cos), 1.f),
                  s_addf(s_add(s_mulf(n2k2, (float)cz2), two_n_cos), 1.f));
  return s_divf(s_add(perp, para), 2.f);
}

/* BSDF::f (reference src/part1_code.cpp:167-169 diffuse; src/bsdf.cpp:68-72 microfacet; zero for every other class) */
__device__ __forceinline__ f3 bsdf_f(const rrt_material& m, d3 wo, d3 wi) {
  if (m.type == RRT_MAT_DIFFUSE) return s_divf(s3(m.a), (float)RRT_PI);
  if (m.type == RRT_MAT_MICROFACET) {
    if (wo.z <= 0 || wi.z <= 0) return F3(0, 0, 0);
    f3 F = mf_F(m, wi);
    return s_divf(s_mulf(s_mulf(F, (float)mf_G(m, wo, wi)), (float)mf_D(m, vunit(vadd(wo, wi)))), (float)(4 * wo.z * wi.z));
  }
  return F3(0, 0, 0);
}

/* BSDF::refract (reference src/bsdf.cpp:146-159) */
__device__ __forceinline__ bool bsdf_refract(d3 wo, d3& wi, float ior) {
  double eta = (wo.z > 0) ? (double)(1 / ior) : (double)ior;
  double wi_z2 = 1 - eta * eta * (1 - wo.z * wo.z);
  if (wi_z2 < 0) return false;
  wi = D3(-eta * wo.x, -eta * wo.y, sqrt(wi_z2));
  if (wo.z > 0) wi.z = -wi.z;
  return true;
}

/* BSDF::sample_f for every class (reference src/part1_code.cpp:171-173; src/bsdf.cpp:37-41,74-96,104-106,114-140,167-171;
 * cosine hemisphere sampler src/sampler.cpp:47-56).  Draws are taken from g in the reference's order. */
__device__ __forceinline__ f3 bsdf_sample_f(const rrt_material& m, d3 wo, d3& wi, float& pdf, Rng& g) {
  switch (m.type) {
    case RRT_MAT_DIFFUSE:
    case RRT_MAT_EMISSION: {
      double Xi1 = rng_uniform(g), Xi2 = rng_uniform(g);
      double r = sqrt(Xi1), theta = 2. * RRT_PI * Xi2;
      pdf = (float)(sqrt(1 - Xi1) / RRT_PI);
      wi = D3(r * cos(theta), r * sin(theta), sqrt(1 - Xi1));
      return m.type == RRT_MAT_DIFFUSE ? bsdf_f(m, wo, wi) : F3(0, 0, 0);
    }
    case RRT_MAT_MIRROR:
      wi = D3(-wo.x, -wo.y, wo.z); pdf = 1.0f;
      return s_divf(s3(m.a), (float)fabs(wi.z));
    case RRT_MAT_GLASS: {
      if (bsdf_refract(wo, wi, m.ior)) {
        double R0 = (double)((1 - m.ior) / (1 + m.ior));
        R0 *= R0;
        double t = (1 - fabs(wi.z)), t2 = t * t, t4 = t2 * t2, R = R0 + (1 - R0) * t4 * t;
        if (rng_uniform(g) < R) {
          wi = D3(-wo.x, -wo.y, wo.z); pdf = (float)R;
          return s_divf(s_mulf(s3(m.a), (float)R), (float)fabs(wi.z));
        } else {
          double eta = (wo.z > 0) ? (double)(1 / m.ior) : (double)m.ior;
          pdf = (float)(1 - R);
          return s_divf(s_mulf(s3(m.b), (float)(1 - R)), (float)(fabs(wi.z) * eta * eta));
        }
      }
      wi = D3(-wo.x, -wo.y, wo.z); pdf = 1.0f;
      return s_divf(s3(m.a), (float)fabs(wi.z));
    }
    case RRT_MAT_MICROFACET: {
      double ux, uy; rng_grid2(g, ux, uy);
      double alpha2 = m.alpha * m.alpha, /* float product */
             ang = atan(sqrt(-alpha2 * log(1 - ux))), azi = 2 * RRT_PI * uy, sn = sin(ang),
             cs = cos(ang), tg = tan(ang),
             pdf_polar = 2 * sn * exp(-tg * tg / alpha2) / (alpha2 * cs * cs * cs),
             pdf_azimuth = 0.5 / RRT_PI;
      d3 h = D3(sn * cos(azi), sn * sin(azi), cs);
      wi = vsub(smul(2 * vdot(wo, h), h), wo);
      if (wi.z <= 0) { pdf = 0; return F3(0, 0, 0); }
      pdf = (float)(pdf_polar * pdf_azimuth / (sn * 4 * vdot(wi, h)));
      return bsdf_f(m, wo, wi);
    }
    default: /* RefractionBSDF: returns Spectrum() and leaves pdf unwritten (reference src/bsdf.cpp:104-106); pdf = 0 ends the path */
      wi = D3(0, 0, 1); pdf = 0; return F3(0, 0, 0);
  }
}

/* UniformHemisphereSampler3D::get_sample (reference src/sampler.cpp:15-29): float sinf/cosf of double angles */
__device__ __forceinline__ d3 hemisphere_sample(Rng& g) {
  double Xi1 = rng_uniform(g), Xi2 = rng_uniform(g);
  double theta = acos(Xi1), phi = 2.0 * RRT_PI * Xi2;
  float st = sinf((float)theta), ct = cosf((float)theta), sp = sinf((float)phi), cp = cosf((float)phi);
  return D3(st * cp, st * sp, ct);
}

/* SceneLight::sample_L (reference src/static_scene/light.cpp:17-92; Spot/Sphere/Mesh stubs :61-115 return zero) */
__device__ __forceinline__ f3 light_sample_L(const rrt_light& l, d3 p, d3& wi, float& dist, float& pdf, Rng& g) {
  switch (l.type) {
    case RRT_LIGHT_DIRECTIONAL:
      wi = D3(l.direction[0], l.direction[1], l.direction[2]); dist = INFINITY; pdf = 1.0f;
      return s3(l.radiance);
    case RRT_LIGHT_HEMISPHERE: {
      d3 s = hemisphere_sample(g);
      d3 col[3] = {D3(1, 0, 0), D3(0, 0, -1), D3(0, 1, 0)};
      wi = mat_mul(col, s); dist = INFINITY; pdf = (float)(1.0 / (2.0 * RRT_PI));
      return s3(l.radiance);
    }
    case RRT_LIGHT_POINT: {
      d3 d = vsub(D3(l.position[0], l.position[1], l.position[2]), p);
      wi = vunit(d); dist = (float)vnorm(d); pdf = 1.0f;
      return s3(l.radiance);
    }
    case RRT_LIGHT_AREA: {
      double sx, sy; rng_grid2(g, sx, sy);
      sx = sx - 0.5f; sy = sy - 0.5f;
      d3 pos = D3(l.position[0], l.position[1], l.position[2]), dx = D3(l.dim_x[0], l.dim_x[1], l.dim_x[2]),
         dy = D3(l.dim_y[0], l.dim_y[1], l.dim_y[2]), dir = D3(l.direction[0], l.direction[1], l.direction[2]);
      d3 d = vsub(vadd(vadd(pos, smul(sx, dx)), smul(sy, dy)), p);
      float sqDist = (float)vnorm2(d);
      float dst = sqrtf(sqDist);
      wi = vdiv(d, dst);
      float cosTheta = (float)vdot(wi, dir);
      dist = dst;
      pdf = sqDist / (l.area * fabsf(cosTheta));
      return cosTheta < 0 ? s3(l.radiance) : F3(0, 0, 0);
    }
    default:
      wi = D3(0, 0, 0); dist = 0; pdf = 0; return F3(0, 0, 0);
  }
}

/* ---- EnvironmentLight (reference src/static_scene/environment_light.cpp), appended after the scene's lights ---- */
/* bilerp (:99-115): the four texels around xy with the reference's wrap / clamp rules, weights narrowed to float */
__device__ __forceinline__ f3 env_bilerp(const DevScene& sc, double x, double y) {
  const long w = sc.env_w, h = sc.env_h;
  long right = lround(x), left, v = lround(y);
  double u1 = right - x + .5, v1;
  if (right == 0 || right == w) { left = w - 1; right = 0; } else left = right - 1;
  if (v == 0) { v = 1; v1 = 1; } else if (v == h) { v = h - 1; v1 = 0; } else v1 = v - y + .5;
  size_t bottom = (size_t)w * (size_t)v, top = bottom - (size_t)w;
  double u0 = 1 - u1;
  f3 tl = s3(sc.env + 3 * (top + left)), tr = s3(sc.env + 3 * (top + right));
  f3 bl = s3(sc.env + 3 * (bottom + left)), br = s3(sc.env + 3 * (bottom + right));
  return s_add(s_mulf(s_add(s_mulf(tl, (float)u1), s_mulf(tr, (float)u0)), (float)v1),
               s_mulf(s_add(s_mulf(bl, (float)u1), s_mulf(br, (float)u0)), (float)(1 - v1)));
}
/* sample_dir (:60-65, :76-80, :145-147): radiance from direction d */
__device__ __forceinline__ f3 env_sample_dir(const DevScene& sc, d3 d) {
  d3 u = vunit(d);
  double theta = acos(u.y), phi = atan2(-u.z, u.x) + RRT_PI;
  return env_bilerp(sc, phi / 2. / RRT_PI * sc.env_w, theta / RRT_PI * sc.env_h);
}
__device__ __forceinline__ int env_upper_bound(const double* __restrict__ a, int n, double key) {   /* std::upper_bound */
  int lo = 0, hi = n;
  while (lo < hi) { int mid = lo + (hi - lo) / 2; if (!(key < __ldg(a + mid))) lo = mid + 1; else hi = mid; }
  return lo;
}
/* sample_L, ENV_HEMI == 0 (:118-133): row from the marginal, column from the row's conditional, direction of the
 * texel CORNER, pdf from the texel.  An index one past the table (a draw of exactly 1.0) is clamped. */
__device__ __forceinline__ f3 env_sample_L(const DevScene& sc, d3& wi, float& dist, float& pdf, Rng& g) {
  const int w = sc.env_w, h = sc.env_h;
  dist = INFINITY;
  double sx, sy; rng_grid2(g, sx, sy);
  int y = env_upper_bound(sc.env_marg, h, sy); if (y > h - 1) y = h - 1;
  int x = env_upper_bound(sc.env_cond + (size_t)w * y, w, sx); if (x > w - 1) x = w - 1;
  double phi = (double)x / w * 2.0 * RRT_PI, theta = (double)y / h * RRT_PI;
  double st = sin(theta);
  wi = D3(cos(phi - RRT_PI) * st, cos(theta), -sin(phi - RRT_PI) * st);
  pdf = (float)(__ldg(sc.env_pdf + (size_t)w * y + x) * w * h / (2 * RRT_PI * RRT_PI * st));
  return env_bilerp(sc, (double)x, (double)y);
}
