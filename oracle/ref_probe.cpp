// ref_probe.cpp — TEST INFRASTRUCTURE.  Our own driver around the UNMODIFIED reference objects
// (compiled in place from /root/reference by oracle/build_ref.sh).  It exists to pin the oracle and
// the CUDA path against the reference's real code: it dumps flattened scenes, runs the reference's
// BVHAccel::intersect / BlackHole::next_micro_ray / Camera::generate_ray / BSDF::sample_f /
// SceneLight::sample_L on caller-supplied inputs, and renders images with the reference's own
// PathTracer::raytrace_pixel.
//
// Random numbers: the reference draws everything from libc rand() (src/random_util.h:11-13).  This
// executable defines its own rand(), which the statically linked reference objects then call.  It
// returns either values from an explicit queue (KAT mode) or the per-path Philox4x32-10 stream the
// CUDA renderer uses (render mode):  draw n of path (pixel, sample) = word (n&3) of
// philox(key = seed, counter = (pixel, sample, n>>2, 0)), rand() = word >> 1  (31 bits, RAND_MAX = 2^31-1).
// With that, the reference's raytrace_pixel and the CUDA path consume identical uniforms and their
// images agree to float rounding, not just statistically.
//
// Nothing in the product links or runs this file.
#include <sstream>
#include <iostream>
#include <fstream>
#include <vector>
#include <string>
#include <thread>
#include <mutex>
#include <condition_variable>
#include <stack>
#include <map>
#include <list>
#include <set>
#include <algorithm>
#include <atomic>
#include <unordered_map>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <omp.h>
#define private public
#define protected public
#include "application.h"
#include "static_scene/blackhole.h"
#include "static_scene/triangle.h"
#include "static_scene/sphere.h"
#include "static_scene/light.h"
#include "static_scene/environment_light.h"
#include "collada/collada.h"
#undef private
#undef protected
#include "../include/rrt_b200.h"

using namespace CGL;
using namespace CGL::StaticScene;

// ---------------------------------------------------------------- rand() interposition
static inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}

struct RandState {
  int mode = 0;  // 0 queue, 1 philox stream, 2 glibc's own generator (random_r: the algorithm behind rand(), per-thread state, no lock)
  struct random_data rdata; char rstate[128]; bool rinit = false;
  std::vector<int> queue; size_t qpos = 0;
  uint64_t seed = 0; uint32_t pixel = 0, sample = 0, n = 0;
  uint64_t draws = 0;
  int grid_calls_per_sample = 1; uint32_t grid_calls = 0;   // thin lens: the second gridSampler call of a sample is the lens sample
};
static thread_local RandState g_rs;

extern "C" int rand(void) {
  RandState& s = g_rs;
  ++s.draws;
  if (s.mode == 0) {
    if (s.qpos >= s.queue.size()) { fprintf(stderr, "[ref_probe] rand queue exhausted\n"); abort(); }
    return s.queue[s.qpos++];
  }
  if (s.mode == 2) { int32_t r = 0; random_r(&s.rdata, &r); return (int)r; }
  uint32_t c[4] = {s.pixel, s.sample, s.n >> 2, 0u};
  philox4x32_10(c, (uint32_t)s.seed, (uint32_t)(s.seed >> 32));
  uint32_t w = c[s.n & 3];
  ++s.n;
  return (int)(w >> 1);
}

// grid sampler hook: PathTracer::raytrace_pixel calls gridSampler->get_sample() at the start of every sample when
// ns_aa > 1 (src/part1_code.cpp:131-135) and once more for the lens when useThinLens (:137-139): the first call
// of a sample switches streams.  (ns_aa == 1 with a lens: the only call is the lens sample, no switch.)
struct StreamSwitchSampler : public Sampler2D {
  UniformGridSampler2D inner;
  Vector2D get_sample() const {
    RandState& s = g_rs;
    if (s.grid_calls_per_sample > 0 && s.grid_calls++ % (uint32_t)s.grid_calls_per_sample == 0) { s.sample += 1; s.n = 0; }   // caller sets sample = (uint32_t)-1 before each pixel
    return inner.get_sample();
  }
};

// ---------------------------------------------------------------- small binary I/O helpers
static std::vector<double> read_doubles(const char* path) {
  FILE* f = fopen(path, "rb"); if (!f) { perror(path); exit(2); }
  fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
  std::vector<double> v(sz / 8);
  if (fread(v.data(), 8, v.size(), f) != v.size()) { perror("fread"); exit(2); }
  fclose(f); return v;
}
static void write_doubles(const char* path, const std::vector<double>& v) {
  FILE* f = fopen(path, "wb"); if (!f) { perror(path); exit(2); }
  fwrite(v.data(), 8, v.size(), f); fclose(f);
}
template <class T> static void put(FILE* f, const T* p, size_t n) {
  fwrite(p, sizeof(T), n, f);
  size_t bytes = sizeof(T) * n, pad = (8 - bytes % 8) % 8; const char z[8] = {0};
  fwrite(z, 1, pad, f);
}

// ---------------------------------------------------------------- scene context
struct Ctx {
  Application* app = nullptr;
  PathTracer* pt = nullptr;
  std::vector<Primitive*> prims;                 // reference primitive order (build_accel order)
  std::unordered_map<const Primitive*, int> prim_index;
  std::vector<BSDF*> mats;                       // distinct BSDFs in order of first appearance
  std::unordered_map<const BSDF*, int> mat_index;
  int W = 0, H = 0;
};

static void parse_bh(int argc, char** argv) {   // optional trailing: -B cx cy cz r dtheta  (src/main.cpp:139-145)
  for (int i = 0; i + 5 < argc; ++i)
    if (!strcmp(argv[i], "-B")) {
      global_black_hole.o = Vector3D(atof(argv[i + 1]), atof(argv[i + 2]), atof(argv[i + 3]));
      global_black_hole.r = atof(argv[i + 4]);
      global_black_hole.r2 = global_black_hole.r * global_black_hole.r;
      global_black_hole.delta_theta = atof(argv[i + 5]);
    }
}

static Ctx load(const char* dae, int W, int H, AppConfig cfg = AppConfig()) {
  Ctx c; c.W = W; c.H = H;
  Collada::SceneInfo* si = new Collada::SceneInfo();
  if (Collada::ColladaParser::load(dae, si) < 0) { fprintf(stderr, "cannot load %s\n", dae); exit(2); }
  // Optional: RRT_PROBE_ENVMAP=<file> gives the reference an environment map without going through an .exr file
  // (none ships with the reference): the file is int32 w, int32 h, then w*h*3 float32 (row 0 = top, as load_exr fills
  // the buffer, src/main.cpp:66-76).  EnvironmentLight's constructor writes probability_debug.png into the cwd.
  if (const char* envf = getenv("RRT_PROBE_ENVMAP")) {
    FILE* f = fopen(envf, "rb"); if (!f) { perror(envf); exit(2); }
    int32_t wh[2]; if (fread(wh, 4, 2, f) != 2) exit(2);
    std::vector<float> px((size_t)wh[0] * wh[1] * 3);
    if (fread(px.data(), 4, px.size(), f) != px.size()) exit(2);
    fclose(f);
    HDRImageBuffer* env = new HDRImageBuffer();
    env->resize(wh[0], wh[1]);
    for (size_t i = 0; i < (size_t)wh[0] * wh[1]; ++i) env->data[i] = Spectrum(px[3 * i], px[3 * i + 1], px[3 * i + 2]);
    cfg.pathtracer_envmap = env;
  }
  // same call sequence as the headless path of main(): src/main.cpp:172-188
  c.app = new Application(cfg, false);
  c.app->init();
  c.app->load(si);
  delete si;
  // Optional: RRT_PROBE_SCREENDIST_SCALE=k multiplies Camera::screenDist before the resize, so that a small test
  // frame covers the same field of view as a k-times larger one (the reference's -r only crops, src/camera.cpp:68-74).
  if (const char* k = getenv("RRT_PROBE_SCREENDIST_SCALE")) c.app->camera.screenDist *= atof(k);
  if (W && H) c.app->resize(W, H);
  // set_up_pathtracer (src/application.cpp:622-628) silences nothing; keep stdout clean for callers
  FILE* saved = stdout; stdout = fopen("/dev/null", "w");
  c.app->set_up_pathtracer();
  fclose(stdout); stdout = saved;
  c.pt = c.app->pathtracer;
  // Rebuild the BVH from a primitive vector we keep, in the order build_accel uses
  // (src/pathtracer.cpp:311-316), so that Intersection::primitive maps back to an index.
  for (SceneObject* obj : c.pt->scene->objects) {
    const std::vector<Primitive*>& op = obj->get_primitives();
    c.prims.insert(c.prims.end(), op.begin(), op.end());
  }
  delete c.pt->bvh;
  c.pt->bvh = new BVHAccel(c.prims);
  for (size_t i = 0; i < c.prims.size(); ++i) {
    c.prim_index[c.prims[i]] = (int)i;
    BSDF* b = c.prims[i]->get_bsdf();
    if (!c.mat_index.count(b)) { c.mat_index[b] = (int)c.mats.size(); c.mats.push_back(b); }
  }
  return c;
}

static rrt_material flatten_material(BSDF* b) {
  rrt_material m; memset(&m, 0, sizeof m);
  auto set3 = [](float* d, const Spectrum& s) { d[0] = s.r; d[1] = s.g; d[2] = s.b; };
  if (auto* p = dynamic_cast<DiffuseBSDF*>(b)) { m.type = RRT_MAT_DIFFUSE; set3(m.a, p->reflectance); }
  else if (auto* p = dynamic_cast<MirrorBSDF*>(b)) { m.type = RRT_MAT_MIRROR; set3(m.a, p->reflectance); }
  else if (auto* p = dynamic_cast<GlassBSDF*>(b)) { m.type = RRT_MAT_GLASS; set3(m.a, p->reflectance); set3(m.b, p->transmittance); m.ior = p->ior; }
  else if (auto* p = dynamic_cast<MicrofacetBSDF*>(b)) { m.type = RRT_MAT_MICROFACET; set3(m.a, p->eta); set3(m.b, p->k); m.alpha = p->alpha; }
  else if (auto* p = dynamic_cast<EmissionBSDF*>(b)) { m.type = RRT_MAT_EMISSION; set3(m.a, p->radiance); }
  else if (auto* p = dynamic_cast<RefractionBSDF*>(b)) { m.type = RRT_MAT_REFRACTION; set3(m.b, p->transmittance); m.ior = p->ior; }
  else { fprintf(stderr, "unknown BSDF class\n"); exit(2); }
  return m;
}

static rrt_light flatten_light(SceneLight* l) {
  rrt_light o; memset(&o, 0, sizeof o);
  auto set3f = [](float* d, const Spectrum& s) { d[0] = s.r; d[1] = s.g; d[2] = s.b; };
  auto set3 = [](double* d, const Vector3D& v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; };
  o.is_delta = l->is_delta_light() ? 1 : 0;
  if (auto* p = dynamic_cast<DirectionalLight*>(l)) { o.type = RRT_LIGHT_DIRECTIONAL; set3f(o.radiance, p->radiance); set3(o.direction, p->dirToLight); }
  else if (auto* p = dynamic_cast<InfiniteHemisphereLight*>(l)) { o.type = RRT_LIGHT_HEMISPHERE; set3f(o.radiance, p->radiance); }
  else if (auto* p = dynamic_cast<PointLight*>(l)) { o.type = RRT_LIGHT_POINT; set3f(o.radiance, p->radiance); set3(o.position, p->position); }
  else if (auto* p = dynamic_cast<AreaLight*>(l)) {
    o.type = RRT_LIGHT_AREA; set3f(o.radiance, p->radiance); set3(o.position, p->position); set3(o.direction, p->direction);
    set3(o.dim_x, p->dim_x); set3(o.dim_y, p->dim_y); o.area = p->area;
  } else { o.type = RRT_LIGHT_NULL; }
  return o;
}

// Scene file "RRTSCN01": see DESIGN.md §data formats.
static void dump_scene(const Ctx& c, const char* out) {
  std::vector<uint8_t> kind; std::vector<uint32_t> index; std::vector<double> tpos, tnrm, sph;
  std::vector<int32_t> tmat, smat;
  for (Primitive* p : c.prims) {
    int mi = c.mat_index.at(p->get_bsdf());
    if (auto* t = dynamic_cast<Triangle*>(p)) {
      kind.push_back(0); index.push_back((uint32_t)tmat.size()); tmat.push_back(mi);
      size_t v[3] = {t->v1, t->v2, t->v3};
      for (int k = 0; k < 3; ++k) { const Vector3D& q = t->mesh->positions[v[k]]; tpos.push_back(q.x); tpos.push_back(q.y); tpos.push_back(q.z); }
      for (int k = 0; k < 3; ++k) { const Vector3D& q = t->mesh->normals[v[k]]; tnrm.push_back(q.x); tnrm.push_back(q.y); tnrm.push_back(q.z); }
    } else if (auto* s = dynamic_cast<Sphere*>(p)) {
      kind.push_back(1); index.push_back((uint32_t)smat.size()); smat.push_back(mi);
      sph.push_back(s->o.x); sph.push_back(s->o.y); sph.push_back(s->o.z); sph.push_back(s->r);
    } else { fprintf(stderr, "unknown primitive class\n"); exit(2); }
  }
  std::vector<rrt_material> mats; for (BSDF* b : c.mats) mats.push_back(flatten_material(b));
  std::vector<rrt_light> lights; for (SceneLight* l : c.pt->scene->lights) lights.push_back(flatten_light(l));
  const Camera& cam = c.app->camera;
  double camv[24] = {cam.pos.x, cam.pos.y, cam.pos.z,
                     cam.c2w[0].x, cam.c2w[0].y, cam.c2w[0].z, cam.c2w[1].x, cam.c2w[1].y, cam.c2w[1].z,
                     cam.c2w[2].x, cam.c2w[2].y, cam.c2w[2].z,
                     cam.hFov, cam.vFov, cam.nClip, cam.fClip, cam.lensRadius, cam.focalDistance,
                     (double)cam.screenW, (double)cam.screenH, cam.screenDist, 0, 0, 0};
  FILE* f = fopen(out, "wb"); if (!f) { perror(out); exit(2); }
  fwrite("RRTSCN01", 1, 8, f);
  uint32_t hdr[8] = {(uint32_t)kind.size(), (uint32_t)tmat.size(), (uint32_t)smat.size(), (uint32_t)mats.size(),
                     (uint32_t)lights.size(), 0, 0, 0};
  fwrite(hdr, 4, 8, f);
  fwrite(camv, 8, 24, f);
  put(f, kind.data(), kind.size()); put(f, index.data(), index.size());
  put(f, tpos.data(), tpos.size()); put(f, tnrm.data(), tnrm.size()); put(f, tmat.data(), tmat.size());
  put(f, sph.data(), sph.size()); put(f, smat.data(), smat.size());
  put(f, mats.data(), mats.size()); put(f, lights.data(), lights.size());
  fclose(f);
  fprintf(stderr, "[ref_probe] %s: %zu prims (%zu tris, %zu spheres), %zu materials, %zu lights\n", out, kind.size(),
          tmat.size(), smat.size(), mats.size(), lights.size());
}

// ---------------------------------------------------------------- commands
// trace: rays.bin = n x [o3 d3 min_t max_t]; out = n x 16 doubles:
//   [hit(0/1), prim, hit_p3, w_out3, n3, anyhit_result(0/1), 0, 0, 0, 0]
// Calls the reference's BVHAccel::intersect(ray,&isect) and intersect(ray) (src/bvh.h:86-109).
static int cmd_trace(int argc, char** argv) {
  if (argc < 7) return 1;
  parse_bh(argc, argv);
  Ctx c = load(argv[2], atoi(argv[3]), atoi(argv[4]));
  std::vector<double> in = read_doubles(argv[5]);
  size_t n = in.size() / 8;
  std::vector<double> out(n * 16, 0.0);
#pragma omp parallel for schedule(dynamic, 64)
  for (long i = 0; i < (long)n; ++i) {
    const double* r = &in[i * 8];
    Ray ray(Vector3D(r[0], r[1], r[2]), Vector3D(r[3], r[4], r[5]), r[7]);
    ray.min_t = r[6];
    Intersection isect;
    bool hit = c.pt->bvh->intersect(ray, &isect);
    double* o = &out[i * 16];
    o[0] = hit ? 1 : 0;
    o[1] = hit ? c.prim_index.at(isect.primitive) : -1;
    if (hit) {
      o[2] = isect.hit_p.x; o[3] = isect.hit_p.y; o[4] = isect.hit_p.z;
      o[5] = isect.w_out.x; o[6] = isect.w_out.y; o[7] = isect.w_out.z;
      o[8] = isect.n.x; o[9] = isect.n.y; o[10] = isect.n.z;
    }
    Ray ray2(Vector3D(r[0], r[1], r[2]), Vector3D(r[3], r[4], r[5]), r[7]);
    ray2.min_t = r[6];
    o[11] = c.pt->bvh->intersect(ray2) ? 1 : 0;
  }
  write_doubles(argv[6], out);
  return 0;
}

// geodesic: rays.bin = n x [o3 d3]; out = n x nsteps x 8 = [o3 d3 len captured(0/1)] per chord.
// Repeats BVHAccel::intersect's chord loop (src/bvh.cpp:104-106) without the geometry, calling the
// reference's BlackHole::next_micro_ray and its horizon test.
static int cmd_geodesic(int argc, char** argv) {
  if (argc < 5) return 1;
  parse_bh(argc, argv);
  std::vector<double> in = read_doubles(argv[2]);
  int nsteps = atoi(argv[3]);
  size_t n = in.size() / 6;
  std::vector<double> out(n * nsteps * 8, 0.0);
  for (size_t i = 0; i < n; ++i) {
    const double* r = &in[i * 6];
    Ray micro(Vector3D(r[0], r[1], r[2]), Vector3D(r[3], r[4], r[5]), 0.0);
    for (int j = 0; j < nsteps; ++j) {
      micro = global_black_hole.next_micro_ray(micro);
      double* o = &out[(i * nsteps + j) * 8];
      o[0] = micro.o.x; o[1] = micro.o.y; o[2] = micro.o.z;
      o[3] = micro.d.x; o[4] = micro.d.y; o[5] = micro.d.z; o[6] = micro.max_t;
      Ray copy = micro;
      o[7] = global_black_hole.intersect(copy) ? 1 : 0;
    }
  }
  write_doubles(argv[4], out);
  return 0;
}

// camrays: xy.bin = n x [x y] normalised; out = n x [o3 d3 min_t max_t]
static int cmd_camrays(int argc, char** argv) {
  if (argc < 7) return 1;
  Ctx c = load(argv[2], atoi(argv[3]), atoi(argv[4]));
  std::vector<double> in = read_doubles(argv[5]);
  size_t n = in.size() / 2;
  std::vector<double> out(n * 8);
  for (size_t i = 0; i < n; ++i) {
    Ray r = c.app->camera.generate_ray(in[2 * i], in[2 * i + 1]);
    double* o = &out[i * 8];
    o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z; o[6] = r.min_t; o[7] = r.max_t;
  }
  write_doubles(argv[6], out);
  return 0;
}

// lensrays: in.bin = n x [x y rndR rndTheta]; out = n x [o3 d3 min_t max_t]  (Camera::generate_ray_for_thin_lens, src/camera.cpp:176-184)
static int cmd_lensrays(int argc, char** argv) {
  if (argc < 7) return 1;
  Ctx c = load(argv[2], atoi(argv[3]), atoi(argv[4]));
  std::vector<double> in = read_doubles(argv[5]);
  size_t n = in.size() / 4;
  std::vector<double> out(n * 8);
  for (size_t i = 0; i < n; ++i) {
    Ray r = c.app->camera.generate_ray_for_thin_lens(in[4 * i], in[4 * i + 1], in[4 * i + 2], in[4 * i + 3]);
    double* o = &out[i * 8];
    o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z; o[6] = r.min_t; o[7] = r.max_t;
  }
  write_doubles(argv[6], out);
  return 0;
}

// bsdf: in = n x [wo3, wi3(for f), r0, r1, r2, r3] (r* = integer rand() values as doubles);
// out = n x 12 = [f_eval3, fs3, wi3, pdf, draws, is_delta]
static int cmd_bsdf(int argc, char** argv) {
  if (argc < 8) return 1;
  Ctx c = load(argv[2], atoi(argv[3]), atoi(argv[4]));
  int mi = atoi(argv[5]);
  if (mi < 0 || mi >= (int)c.mats.size()) { fprintf(stderr, "material index out of range (%zu)\n", c.mats.size()); return 2; }
  BSDF* b = c.mats[mi];
  std::vector<double> in = read_doubles(argv[6]);
  size_t n = in.size() / 10;
  std::vector<double> out(n * 12, 0.0);
  for (size_t i = 0; i < n; ++i) {
    const double* r = &in[i * 10];
    Vector3D wo(r[0], r[1], r[2]), wi_e(r[3], r[4], r[5]);
    Spectrum fe = b->f(wo, wi_e);
    g_rs.mode = 0; g_rs.queue = {(int)r[6], (int)r[7], (int)r[8], (int)r[9]}; g_rs.qpos = 0;
    Vector3D wi; float pdf = -1.f;
    Spectrum fs = b->sample_f(wo, &wi, &pdf);
    double* o = &out[i * 12];
    o[0] = fe.r; o[1] = fe.g; o[2] = fe.b; o[3] = fs.r; o[4] = fs.g; o[5] = fs.b;
    o[6] = wi.x; o[7] = wi.y; o[8] = wi.z; o[9] = pdf; o[10] = (double)g_rs.qpos; o[11] = b->is_delta() ? 1 : 0;
  }
  write_doubles(argv[7], out);
  return 0;
}

// light: in = n x [p3, r0, r1]; out = n x 10 = [L3, wi3, dist, pdf, draws, is_delta]
static int cmd_light(int argc, char** argv) {
  if (argc < 8) return 1;
  Ctx c = load(argv[2], atoi(argv[3]), atoi(argv[4]));
  int li = atoi(argv[5]);
  if (li < 0 || li >= (int)c.pt->scene->lights.size()) { fprintf(stderr, "light index out of range\n"); return 2; }
  SceneLight* l = c.pt->scene->lights[li];
  std::vector<double> in = read_doubles(argv[6]);
  size_t n = in.size() / 5;
  std::vector<double> out(n * 10, 0.0);
  for (size_t i = 0; i < n; ++i) {
    const double* r = &in[i * 5];
    g_rs.mode = 0; g_rs.queue = {(int)r[3], (int)r[4]}; g_rs.qpos = 0;
    Vector3D wi; float dist = 0, pdf = 0;
    Spectrum L = l->sample_L(Vector3D(r[0], r[1], r[2]), &wi, &dist, &pdf);
    double* o = &out[i * 10];
    o[0] = L.r; o[1] = L.g; o[2] = L.b; o[3] = wi.x; o[4] = wi.y; o[5] = wi.z; o[6] = dist; o[7] = pdf;
    o[8] = (double)g_rs.qpos; o[9] = l->is_delta_light() ? 1 : 0;
  }
  write_doubles(argv[7], out);
  return 0;
}

// envdir: EnvironmentLight::sample_dir (src/static_scene/environment_light.cpp:145-147) for n directions.
// in = n x [d3]; out = n x [L3].  Needs RRT_PROBE_ENVMAP.
static int cmd_envdir(int argc, char** argv) {
  if (argc < 7) return 1;
  Ctx c = load(argv[2], atoi(argv[3]), atoi(argv[4]));
  EnvironmentLight* env = c.pt->scene->lights.empty() ? nullptr : dynamic_cast<EnvironmentLight*>(c.pt->scene->lights.back());
  if (!env) { fprintf(stderr, "no environment light (set RRT_PROBE_ENVMAP)\n"); return 2; }
  std::vector<double> in = read_doubles(argv[5]);
  size_t n = in.size() / 3;
  std::vector<double> out(n * 3, 0.0);
  for (size_t i = 0; i < n; ++i) {
    Spectrum L = env->sample_dir(Ray(Vector3D(0, 0, 0), Vector3D(in[3 * i], in[3 * i + 1], in[3 * i + 2])));
    out[3 * i] = L.r; out[3 * i + 1] = L.g; out[3 * i + 2] = L.b;
  }
  write_doubles(argv[6], out);
  return 0;
}

// order: which of two consecutive rand() draws becomes .x of UniformGridSampler2D::get_sample()
// (argument evaluation order is the compiler's choice, src/sampler.cpp:7-11).  Prints "x=first" or "x=second".
static int cmd_order(int, char**) {
  g_rs.mode = 0; g_rs.queue = {0, RAND_MAX}; g_rs.qpos = 0;
  UniformGridSampler2D s; Vector2D v = s.get_sample();
  printf("%s\n", v.x == 0.0 ? "x=first" : "x=second");
  return 0;
}

// bvhorder: the reference's OWN tree (BVHAccel::construct_bvh, src/bvh.cpp:49-96) walked as intersect_micro walks it (left, then right; a
// leaf's primitives in vector order, src/bvh.cpp:120-135): out = [nodes, leaves, depth, rank of primitive 0, rank of primitive 1, ...] as
// doubles, rank = position of the primitive in that depth-first leaf order.  Pins the tie-break ranks (Triangle::intersect keeps the LAST
// primitive among equal t, src/static_scene/triangle.cpp:43-45) to the reference's tree itself rather than to a restatement of its builder.
static void walk_bvh(const BVHNode* n, int depth, const std::map<const Primitive*, size_t>& idx, std::vector<double>& rank, size_t& next, size_t& nodes,
                     size_t& leaves, int& maxdepth) {
  ++nodes; maxdepth = std::max(maxdepth, depth);
  if (n->isLeaf()) {
    ++leaves;
    for (const Primitive* p : *n->prims) rank[idx.at(p)] = (double)next++;
    return;
  }
  if (n->l) walk_bvh(n->l, depth + 1, idx, rank, next, nodes, leaves, maxdepth);
  if (n->r) walk_bvh(n->r, depth + 1, idx, rank, next, nodes, leaves, maxdepth);
}
static int cmd_bvhorder(int argc, char** argv) {
  if (argc < 6) return 1;
  Ctx c = load(argv[2], atoi(argv[3]), atoi(argv[4]));
  std::map<const Primitive*, size_t> idx;
  for (size_t i = 0; i < c.prims.size(); ++i) idx[c.prims[i]] = i;
  std::vector<double> rank(c.prims.size(), -1.0);
  size_t next = 0, nodes = 0, leaves = 0; int depth = 0;
  walk_bvh(c.pt->bvh->get_root(), 0, idx, rank, next, nodes, leaves, depth);
  std::vector<double> out = {(double)nodes, (double)leaves, (double)depth};
  out.insert(out.end(), rank.begin(), rank.end());
  write_doubles(argv[5], out);
  fprintf(stderr, "[ref_probe] bvhorder: %zu nodes, %zu leaves, depth %d, %zu primitives ranked\n", nodes, leaves, depth, next);
  return next == c.prims.size() ? 0 : 2;
}

// render: the reference's own PathTracer::raytrace_pixel (src/part1_code.cpp:125-163) for every pixel
// (or rows [y0,y1) with RRT_PROBE_ROWS=y0:y1), its rand() fed by the per-path Philox stream.
// out = W*H x [r g b count] doubles (row 0 = bottom, sampleBuffer layout) followed by
// [total_rays, total_isects, total_draws, seconds].
static int cmd_render(int argc, char** argv) {
  if (argc < 12) return 1;
  parse_bh(argc, argv);
  AppConfig cfg;
  int W = atoi(argv[3]), H = atoi(argv[4]);
  cfg.pathtracer_ns_aa = atoi(argv[5]);
  cfg.pathtracer_ns_area_light = atoi(argv[6]);
  cfg.pathtracer_max_ray_depth = atoi(argv[7]);
  cfg.pathtracer_samples_per_patch = atoi(argv[8]);
  cfg.pathtracer_max_tolerance = atof(argv[9]);
  uint64_t seed = strtoull(argv[10], nullptr, 10);
  bool lens = false;   // THIN_LENS is compiled out of raytrace_tile (src/pathtracer.h:6) but raytrace_pixel still takes the flag
  for (int i = 12; i < argc; ++i) {
    if (!strcmp(argv[i], "hemi")) cfg.pathtracer_direct_hemisphere_sample = true;
    if (!strcmp(argv[i], "lens")) lens = true;
  }
  Ctx c = load(argv[2], W, H, cfg);
  delete c.pt->gridSampler;
  c.pt->gridSampler = new StreamSwitchSampler();
  int y0 = 0, y1 = H;
  if (const char* rows = getenv("RRT_PROBE_ROWS")) sscanf(rows, "%d:%d", &y0, &y1);
  const bool glibc_rand = getenv("RRT_PROBE_GLIBC_RAND") != nullptr;
  int x0 = 0, x1 = W;   /* with RRT_PROBE_COLS=x0:x1 a cell of the frame (pixels outside stay zero) */
  if (const char* cols = getenv("RRT_PROBE_COLS")) sscanf(cols, "%d:%d", &x0, &x1);
  std::vector<double> out((size_t)W * H * 4 + 4, 0.0);
  uint64_t draws = 0;
  double t0 = omp_get_wtime();
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : draws)
  for (int y = y0; y < y1; ++y) {
    for (int x = x0; x < x1; ++x) {
      RandState& s = g_rs;
      s.mode = 1; s.seed = seed; s.pixel = (uint32_t)(x + y * W); s.n = 0; s.draws = 0;
      if (glibc_rand) {   /* RRT_PROBE_GLIBC_RAND=1: an RNG that shares nothing with the product's Philox streams (statistical parity check) */
        if (!s.rinit) { memset(&s.rdata, 0, sizeof s.rdata); initstate_r((unsigned)(seed * 2654435761u + 977u * (unsigned)omp_get_thread_num() + 1u), s.rstate, sizeof s.rstate, &s.rdata); s.rinit = true; }
        s.mode = 2;
      }
      // ns_aa == 1: no jitter draw, the stream of sample 0 starts at the first light sample.
      s.sample = cfg.pathtracer_ns_aa == 1 ? 0u : (uint32_t)-1;
      s.grid_calls = 0; s.grid_calls_per_sample = cfg.pathtracer_ns_aa == 1 ? 0 : (lens ? 2 : 1);
      Spectrum sp = c.pt->raytrace_pixel(x, y, lens);
      double* o = &out[((size_t)x + (size_t)y * W) * 4];
      o[0] = sp.r; o[1] = sp.g; o[2] = sp.b; o[3] = c.pt->sampleCountBuffer[x + y * W];
      draws += s.draws;
    }
  }
  double t1 = omp_get_wtime();
  double* tail = &out[(size_t)W * H * 4];
  tail[0] = (double)c.pt->bvh->total_rays; tail[1] = (double)c.pt->bvh->total_isects; tail[2] = (double)draws; tail[3] = t1 - t0;
  write_doubles(argv[11], out);
  fprintf(stderr, "[ref_probe] render %dx%d rows %d:%d  %.2fs  rays=%llu (racy if threaded)\n", W, H, y0, y1, t1 - t0,
          (unsigned long long)c.pt->bvh->total_rays);
  return 0;
}

static int cmd_scene(int argc, char** argv) {
  if (argc < 6) return 1;
  Ctx c = load(argv[2], atoi(argv[3]), atoi(argv[4]));
  dump_scene(c, argv[5]);
  return 0;
}

int main(int argc, char** argv) {
  int rc = 1;
  if (argc >= 2) {
    std::string cmd = argv[1];
    if (cmd == "scene") rc = cmd_scene(argc, argv);
    else if (cmd == "trace") rc = cmd_trace(argc, argv);
    else if (cmd == "geodesic") rc = cmd_geodesic(argc, argv);
    else if (cmd == "camrays") rc = cmd_camrays(argc, argv);
    else if (cmd == "lensrays") rc = cmd_lensrays(argc, argv);
    else if (cmd == "bsdf") rc = cmd_bsdf(argc, argv);
    else if (cmd == "light") rc = cmd_light(argc, argv);
    else if (cmd == "envdir") rc = cmd_envdir(argc, argv);
    else if (cmd == "render") rc = cmd_render(argc, argv);
    else if (cmd == "order") rc = cmd_order(argc, argv);
    else if (cmd == "bvhorder") rc = cmd_bvhorder(argc, argv);
  }
  if (rc == 1)
    fprintf(stderr,
            "usage: ref_probe scene <dae> W H out.rrtscene\n"
            "       ref_probe bvhorder <dae> W H out.bin\n"
            "       ref_probe trace <dae> W H rays.bin out.bin [-B cx cy cz r dtheta]\n"
            "       ref_probe geodesic rays.bin nsteps out.bin [-B ...]\n"
            "       ref_probe camrays <dae> W H xy.bin out.bin\n"
            "       ref_probe lensrays <dae> W H in.bin out.bin\n"
            "       ref_probe bsdf <dae> W H mat in.bin out.bin\n"
            "       ref_probe light <dae> W H light in.bin out.bin\n"
            "       ref_probe render <dae> W H spp l m batch tol seed out.bin [hemi] [lens] [-B ...]\n"
            "       ref_probe order\n");
  return rc;
}
