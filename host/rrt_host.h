// rrt_host.h — C++ host side above the C-ABI (include/rrt_b200.h).
//
// Mirrors, for the rendering path only, the interface the reference's Application drives:
//   PathTracer   reference src/pathtracer.h:60-283   (set_scene, set_camera, set_frame_size, start_raytracing,
//                                                     render_to_file, save_image, states INIT/READY/RENDERING/DONE)
//   BVHAccel     reference src/bvh.h:44-144          (constructor from primitives, get_bbox, intersect)
//   Camera       reference src/camera.h / camera.cpp (configure, place, set_screen_size, settings files)
// and loads scenes: COLLADA files with the reference loader's semantics (host/dae_loader.cpp) or flattened RRTSCN01
// files.  All rendering happens in librrt_b200.so on the GPU; nothing here computes radiance, and there is no CPU
// fallback: without a CUDA device PathTracer reports the library's error.
#ifndef RRT_HOST_H
#define RRT_HOST_H

#include <stdint.h>

#include <atomic>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../include/rrt_b200.h"

namespace rrt_host {

struct Vector3D {
  double x = 0, y = 0, z = 0;
  Vector3D() {}
  Vector3D(double x_, double y_, double z_) : x(x_), y(y_), z(z_) {}
  double& operator[](int i) { return (&x)[i]; }
  const double& operator[](int i) const { return (&x)[i]; }
};

struct Ray {            // reference src/ray.h:20-60
  Vector3D o, d;
  double min_t = 0.0, max_t = 1e300;
  Ray() {}
  Ray(const Vector3D& o_, const Vector3D& d_) : o(o_), d(d_) {}
  Ray(const Vector3D& o_, const Vector3D& d_, double max_t_) : o(o_), d(d_), max_t(max_t_) {}
};

struct Intersection {   // reference src/intersection.h:20-35, plus the fields rrt_hit adds
  double t = 0;
  int primitive = -1;   // index in the reference primitive order
  Vector3D hit_p, w_out, n;
  int status = 0, chord = 0, steps = 0;
};

// The reference's StaticScene flattened in PathTracer::build_accel's primitive order; owns the arrays an rrt_scene
// points into.
struct FlatScene {
  std::vector<uint8_t> prim_kind;
  std::vector<uint32_t> prim_index;
  std::vector<double> tri_pos, tri_nrm;
  std::vector<int32_t> tri_mat;
  std::vector<double> sph;
  std::vector<int32_t> sph_mat;
  std::vector<rrt_material> mats;
  std::vector<rrt_light> lights;
  rrt_scene view() const;
  size_t n_prims() const { return prim_kind.size(); }
};

// Camera (reference src/camera.h): placement and field-of-view bookkeeping; ray generation happens on the device.
struct Camera {
  double hFov = 50, vFov = 35, ar = 4.0 / 3.0, nClip = 0.01, fClip = 100;
  Vector3D pos, targetPos;
  double phi = 0, theta = 0, r = 1, minR = 0, maxR = 1e300;
  Vector3D c2w[3];                   // columns
  size_t screenW = 800, screenH = 600;
  double screenDist = 1;
  double lensRadius = 0.25, focalDistance = 4.7;   // AppConfig defaults (reference src/application.h:62-63)

  void configure(double hFov_deg, double vFov_deg, double nClip_, double fClip_, size_t w, size_t h);   // camera.cpp:22-40
  void place(const Vector3D& target, double phi_, double theta_, double r_, double minR_, double maxR_); // camera.cpp:42-54
  void set_screen_size(size_t w, size_t h);                                                              // camera.cpp:68-74
  void compute_position();                                                                               // camera.cpp:96-119
  bool load_settings(const std::string& filename);                                                       // camera.cpp:154-169
  bool dump_settings(const std::string& filename) const;                                                 // camera.cpp:138-152
  rrt_camera view() const;
};

// ---- scene files
// COLLADA with the reference loader's semantics: ColladaParser::load + Application::load (camera placement included).
// Returns false and fills `error` on failure.  `screen_w/h` are the frame size in effect while loading (800x600 in the
// reference: Application::init).
bool load_dae(const std::string& filename, FlatScene* scene, Camera* camera, std::string* error, size_t screen_w = 800, size_t screen_h = 600);
bool load_rrtscene(const std::string& filename, FlatScene* scene, Camera* camera, std::string* error);
bool save_rrtscene(const std::string& filename, const FlatScene& scene, const Camera& camera, std::string* error);

// Vertex order and vertex normals of a polygon soup as the reference's half-edge mesh produces them
// (HalfedgeMesh::build + Vertex::computeNormal + StaticScene::Mesh, see host/halfedge_order.cpp).
// out_tri: 3 vertex indices per polygon; out_normals: one per vertex position.
bool halfedge_triangles_and_normals(const std::vector<std::vector<size_t>>& polygons, const std::vector<Vector3D>& positions,
                                    std::vector<size_t>* out_tri, std::vector<Vector3D>* out_normals, std::string* error);

// ---- environment maps for -e: scan-line OpenEXR (NONE / RLE / ZIPS / ZIP / PIZ; HALF / FLOAT / UINT) or PFM -> [h][w][3] floats, row 0 = top
bool load_envmap(const std::string& filename, int* width, int* height, std::vector<float>* rgb, std::string* error);

// ---- PNG (8-bit RGBA, stored deflate blocks: no compression library in the image)
bool write_png(const std::string& filename, const uint32_t* rgba, size_t w, size_t h);

// ---- operators
class BVHAccel {
 public:
  // builds the flattened BVH on the host and uploads scene + BVH to the device (reference: BVHAccel(primitives, 4))
  explicit BVHAccel(const FlatScene& scene, size_t max_leaf_size = 4, int device = 0);
  ~BVHAccel();
  BVHAccel(const BVHAccel&) = delete;
  BVHAccel& operator=(const BVHAccel&) = delete;
  bool ok() const { return ctx_ != nullptr && error_.empty(); }
  const std::string& error() const { return error_; }
  void get_bbox(Vector3D* lo, Vector3D* hi) const;
  void set_blackhole(const Vector3D& centre, double r, double delta_theta, bool flat = false);
  bool intersect(const Ray& ray);                       // any hit
  bool intersect(const Ray& ray, Intersection* isect);  // closest hit
  // batch form of the two calls above (what a caller with more than one ray should use)
  bool intersect(const std::vector<Ray>& rays, std::vector<Intersection>* out, bool any_hit = false);
  int32_t ref_nodes = 0, ref_leaves = 0, ref_depth = 0;

 private:
  rrt_ctx* ctx_ = nullptr;
  std::string error_;
  Vector3D lo_, hi_;
};

// RRT_HOST_TIMING=1: wall-clock milestones of a run on stderr (tools/cli_timing.sh)
void rrt_host_tick(const char* what);

class PathTracer {
 public:
  enum State { INIT, READY, RENDERING, DONE };
  PathTracer(size_t ns_aa = 1, size_t max_ray_depth = 4, size_t ns_area_light = 1, size_t ns_diff = 1, size_t ns_glsy = 1, size_t ns_refr = 1,
             size_t num_threads = 1, size_t samples_per_batch = 32, float max_tolerance = 0.05f, const void* envmap = nullptr,
             bool direct_hemisphere_sample = false, std::string filename = "", double lensRadius = 0.25, double focalDistance = 4.7);
  ~PathTracer();
  PathTracer(const PathTracer&) = delete;
  PathTracer& operator=(const PathTracer&) = delete;

  void set_scene(const FlatScene* scene);    // copies to the device; only in INIT (silently ignored otherwise, like the reference)
  void set_camera(Camera* camera);           // not owned; lens parameters are overwritten (reference src/pathtracer.cpp:127-128)
  void set_frame_size(size_t width, size_t height);
  void set_blackhole(const Vector3D& centre, double r, double delta_theta, bool flat = false);   // -B (reference src/main.cpp:139-145)
  // -e: the reference takes an HDRImageBuffer* in its constructor (src/pathtracer.cpp:61-65); here the pixels are
  // handed over before set_scene.  rgb: [height][width][3], row 0 = top.
  void set_envmap(int width, int height, const float* rgb);
  void set_seed(uint64_t seed) { seed_ = seed; }
  // rrt_set_option pass-through (RRT_OPT_*: stepper, BVH builder, A/B variants); applied before the next set_scene
  void set_option(int option, int64_t value) { options_.push_back({option, value}); }
  void set_device(int device) { device_ = device; devices_.clear(); }
  // hand over a context created elsewhere (host/main.cpp creates it on a thread while the scene file is parsed); owned from here on
  void adopt_context(rrt_ctx* ctx) { if (!ctx_) ctx_ = ctx; }
  // several GPUs of this box (rrt_create_multi): the reference's N worker threads become N GPUs, tiles interleaved, one NCCL reduce per frame
  void set_devices(const std::vector<int>& devices) { devices_ = devices; if (!devices.empty()) device_ = devices[0]; }
  void set_tile_shard(int tile, int rank, int world) { tile_ = tile; rank_ = rank; world_ = world; }
  void start_raytracing();                   // asynchronous: returns at once, state becomes DONE when the frame is ready
  void stop();                               // abandons a running render after its current adaptive batch (reference: continueRaytracing
                                             // = false, src/pathtracer.cpp:180-200) and returns to READY; the partial frame stays readable
  // Progressive output (stands in for the frame the viewer shows while RENDERING): after every `every`-th
  // adaptive batch the running estimate is tone-mapped and written to <prefix>_<samples>.png.  Empty prefix = off.
  void set_progress_snapshots(const std::string& prefix, int every = 1) { snap_prefix_ = prefix; snap_every_ = every > 0 ? every : 1; }
  int snapshots_written() const { return snaps_written_; }
  bool use_thin_lens = false;                // raytrace_pixel(x, y, useThinLens): the reference compiles it out (THIN_LENS 0, src/pathtracer.h:6)
  void clear();
  bool is_done();
  State get_state();
  void render_to_file(std::string filename, size_t x = (size_t)-1, size_t y = 0, size_t dx = 0, size_t dy = 0);   // blocking
  void save_image(std::string filename = "");
  void save_sampling_rate_image(std::string filename);
  const std::vector<float>& hdr() const { return hdr_; }                 // W*H*3, row 0 = image bottom
  const std::vector<int32_t>& sample_counts() const { return counts_; }
  const rrt_stats& stats() const { return stats_; }
  const std::string& error() const { return error_; }

  bool render_cell = false;
  double cell_tl[2] = {0, 0}, cell_br[2] = {0, 0};

 private:
  bool has_valid_configuration() const;
  void worker();
  static int progress_trampoline(void* self, int32_t done, int32_t total, int64_t active);
  int on_progress(int32_t done, int32_t total, int64_t active);
  std::atomic<bool> stop_requested_{false};
  std::string snap_prefix_; int snap_every_ = 1, snap_calls_ = 0, snaps_written_ = 0;
  rrt_ctx* ctx_ = nullptr;
  int device_ = 0, tile_ = 8, rank_ = 0, world_ = 1;
  std::vector<int> devices_;
  bool have_scene_ = false;
  Camera* camera_ = nullptr;
  size_t W_ = 0, H_ = 0;
  size_t ns_aa_, max_ray_depth_, ns_area_light_, samples_per_batch_;
  float max_tolerance_;
  bool hemisphere_;
  std::string filename_;
  double lensRadius_, focalDistance_;
  uint64_t seed_ = 0;
  std::vector<std::pair<int, int64_t>> options_;
  Vector3D bh_centre_ = Vector3D(0, 1, 0);
  double bh_r_ = 0.1, bh_dtheta_ = 0.1;
  bool bh_flat_ = false;
  State state_ = INIT;
  std::mutex m_done_;
  std::condition_variable cv_done_;
  std::thread thread_;
  std::vector<float> hdr_;
  std::vector<int32_t> counts_;
  std::vector<uint32_t> frame_;    // tone-mapped RGBA, row 0 = image bottom
  rrt_stats stats_;
  std::string error_;
  size_t n_prims_ = 0;
  int env_w_ = 0, env_h_ = 0;
  std::vector<float> env_;
};

}  // namespace rrt_host

#endif
