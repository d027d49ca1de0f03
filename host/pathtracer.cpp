// pathtracer.cpp — PathTracer and BVHAccel of the reference, re-hosted on the C-ABI (include/rrt_b200.h).
//
// PathTracer keeps the reference's contract (src/pathtracer.cpp): calls made in the wrong state return silently,
// start_raytracing() is asynchronous and completion is signalled through a condition variable, render_to_file() blocks,
// save_image() flips rows and also writes the *_rate.png heat map.  What used to be a pool of worker threads pulling
// 32x32 tiles (src/pathtracer.cpp:224-282, 611-644) is ONE thread that hands the whole frame to rrt_render.
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <time.h>

#include <algorithm>
#include <sstream>

#include "rrt_host.h"

namespace rrt_host {

// ------------------------------------------------------------------------------------------------------------ BVHAccel
BVHAccel::BVHAccel(const FlatScene& scene, size_t max_leaf_size, int device) {
  lo_ = Vector3D(INFINITY, INFINITY, INFINITY); hi_ = Vector3D(-INFINITY, -INFINITY, -INFINITY);
  if (max_leaf_size != 4) { error_ = "BVHAccel: the device BVH keeps the reference's max_leaf_size = 4 for tie-breaking ranks"; return; }
  if (rrt_create(&ctx_, device) != RRT_OK) { error_ = rrt_last_error(nullptr); ctx_ = nullptr; return; }
  rrt_scene s = scene.view();
  if (rrt_upload_scene(ctx_, &s) != RRT_OK) { error_ = rrt_last_error(ctx_); return; }
  rrt_bvh_info(ctx_, &ref_nodes, &ref_leaves, &ref_depth, nullptr);
  for (size_t t = 0; t < scene.tri_pos.size(); t += 3)
    for (int k = 0; k < 3; ++k) { lo_[k] = std::min(lo_[k], scene.tri_pos[t + k]); hi_[k] = std::max(hi_[k], scene.tri_pos[t + k]); }
  for (size_t q = 0; q + 3 < scene.sph.size(); q += 4)
    for (int k = 0; k < 3; ++k) { lo_[k] = std::min(lo_[k], scene.sph[q + k] - scene.sph[q + 3]); hi_[k] = std::max(hi_[k], scene.sph[q + k] + scene.sph[q + 3]); }
}

BVHAccel::~BVHAccel() { if (ctx_) rrt_destroy(ctx_); }

void BVHAccel::get_bbox(Vector3D* lo, Vector3D* hi) const { *lo = lo_; *hi = hi_; }

void BVHAccel::set_blackhole(const Vector3D& c, double r, double delta_theta, bool flat) {
  if (!ctx_) return;
  double ctr[3] = {c.x, c.y, c.z};
  if (rrt_set_blackhole(ctx_, ctr, r, delta_theta, flat ? RRT_MODE_FLAT : RRT_MODE_LITERAL) != RRT_OK) error_ = rrt_last_error(ctx_);
}

bool BVHAccel::intersect(const std::vector<Ray>& rays, std::vector<Intersection>* out, bool any_hit) {
  if (!ctx_) return false;
  const size_t n = rays.size();
  std::vector<double> o(3 * n), d(3 * n), tmin(n), tmax(n);
  for (size_t i = 0; i < n; ++i) {
    for (int k = 0; k < 3; ++k) { o[3 * i + k] = rays[i].o[k]; d[3 * i + k] = rays[i].d[k]; }
    tmin[i] = rays[i].min_t; tmax[i] = rays[i].max_t;
  }
  std::vector<rrt_hit> hits(n);
  if (rrt_trace_rays(ctx_, (int64_t)n, o.data(), d.data(), tmin.data(), tmax.data(), any_hit ? 1 : 0, hits.data(), nullptr) != RRT_OK) {
    error_ = rrt_last_error(ctx_);
    return false;
  }
  out->resize(n);
  for (size_t i = 0; i < n; ++i) {
    Intersection& q = (*out)[i];
    const rrt_hit& h = hits[i];
    q.status = h.status; q.primitive = h.prim; q.chord = h.chord; q.steps = h.steps; q.t = h.t;
    q.hit_p = Vector3D(h.hit_p[0], h.hit_p[1], h.hit_p[2]); q.w_out = Vector3D(h.w_out[0], h.w_out[1], h.w_out[2]);
    q.n = Vector3D(h.n[0], h.n[1], h.n[2]);
  }
  return true;
}

bool BVHAccel::intersect(const Ray& ray) {
  std::vector<Intersection> out;
  return intersect(std::vector<Ray>(1, ray), &out, true) && out[0].status == 1;
}

bool BVHAccel::intersect(const Ray& ray, Intersection* isect) {
  std::vector<Intersection> out;
  if (!intersect(std::vector<Ray>(1, ray), &out, false) || out[0].status != 1) return false;
  if (isect) *isect = out[0];
  return true;
}

// ---------------------------------------------------------------------------------------------------------- PathTracer
PathTracer::PathTracer(size_t ns_aa, size_t max_ray_depth, size_t ns_area_light, size_t, size_t, size_t, size_t, size_t samples_per_batch,
                       float max_tolerance, const void* envmap, bool direct_hemisphere_sample, std::string filename, double lensRadius,
                       double focalDistance)
    : ns_aa_(ns_aa), max_ray_depth_(max_ray_depth), ns_area_light_(ns_area_light), samples_per_batch_(samples_per_batch),
      max_tolerance_(max_tolerance), hemisphere_(direct_hemisphere_sample), filename_(filename), lensRadius_(lensRadius),
      focalDistance_(focalDistance) {
  memset(&stats_, 0, sizeof stats_);
  (void)envmap;   /* the pixels come through set_envmap: an opaque HDRImageBuffer* cannot cross this boundary */
}

void rrt_host_tick(const char* what) {
  static const bool on = getenv("RRT_HOST_TIMING") != nullptr;
  static struct timespec t0; static bool have = false;
  if (!on) return;
  struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t);
  if (!have) { t0 = t; have = true; }
  fprintf(stderr, "[timing] %8.1f ms  %s\n", 1e3 * (t.tv_sec - t0.tv_sec) + 1e-6 * (t.tv_nsec - t0.tv_nsec), what);
}

PathTracer::~PathTracer() {
  rrt_host_tick("~PathTracer: before rrt_destroy");
  if (thread_.joinable()) thread_.join();
  if (ctx_) rrt_destroy(ctx_);
  rrt_host_tick("~PathTracer: after rrt_destroy");
}

bool PathTracer::has_valid_configuration() const { return have_scene_ && camera_ && W_ && H_; }

void PathTracer::set_scene(const FlatScene* scene) {
  if (state_ != INIT) return;
  rrt_host_tick("set_scene: before rrt_create");
  if (!ctx_ && (devices_.size() > 1 ? rrt_create_multi(&ctx_, (int)devices_.size(), devices_.data()) : rrt_create(&ctx_, device_)) != RRT_OK) {
    error_ = rrt_last_error(nullptr); ctx_ = nullptr; return;
  }
  rrt_host_tick("set_scene: context created");
  for (const auto& o : options_) if (rrt_set_option(ctx_, o.first, o.second) != RRT_OK) { error_ = rrt_last_error(ctx_); return; }
  fprintf(stdout, "[PathTracer] Collecting primitives... Done! (%.4f sec)\n", 0.0);
  fprintf(stdout, "[PathTracer] Building BVH from %lu primitives... ", (unsigned long)scene->n_prims()); fflush(stdout);
  struct timespec t0, t1; clock_gettime(CLOCK_MONOTONIC, &t0);
  rrt_scene s = scene->view();
  if (rrt_upload_scene(ctx_, &s) != RRT_OK) { error_ = rrt_last_error(ctx_); fprintf(stdout, "failed: %s\n", error_.c_str()); return; }
  clock_gettime(CLOCK_MONOTONIC, &t1);
  fprintf(stdout, "Done! (%.4f sec)\n", (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec));
  if (!env_.empty()) {
    fprintf(stdout, "[PathTracer] Initializing environment light..."); fflush(stdout);
    if (rrt_set_envmap(ctx_, env_w_, env_h_, env_.data()) != RRT_OK) { error_ = rrt_last_error(ctx_); fprintf(stdout, "failed: %s\n", error_.c_str()); return; }
    fprintf(stdout, "done.\n");
  }
  rrt_host_tick("set_scene: scene uploaded, BVH built");
  have_scene_ = true; n_prims_ = scene->n_prims();
  if (has_valid_configuration()) state_ = READY;
}

void PathTracer::set_camera(Camera* camera) {
  if (state_ != INIT) return;
  camera_ = camera;
  camera_->lensRadius = lensRadius_;
  camera_->focalDistance = focalDistance_;
  if (has_valid_configuration()) state_ = READY;
}

void PathTracer::set_frame_size(size_t width, size_t height) {
  if (state_ != INIT && state_ != READY) stop();
  W_ = width; H_ = height;
  hdr_.assign(3 * W_ * H_, 0.f); counts_.assign(W_ * H_, 0); frame_.assign(W_ * H_, 0xFF000000u);
  if (has_valid_configuration()) state_ = READY;
}

void PathTracer::set_envmap(int width, int height, const float* rgb) {
  if (state_ != INIT) return;
  env_w_ = width; env_h_ = height;
  if (rgb && width > 0 && height > 0) env_.assign(rgb, rgb + 3 * (size_t)width * (size_t)height); else { env_.clear(); env_w_ = env_h_ = 0; }
}

void PathTracer::set_blackhole(const Vector3D& c, double r, double dtheta, bool flat) { bh_centre_ = c; bh_r_ = r; bh_dtheta_ = dtheta; bh_flat_ = flat; }

void PathTracer::worker() {
  auto finish = [&](const char* what) {
    if (what) error_ = std::string(what) + ": " + (ctx_ ? rrt_last_error(ctx_) : "no context");
    std::lock_guard<std::mutex> lk(m_done_);
    state_ = DONE;
    cv_done_.notify_all();
  };
  rrt_camera cam = camera_->view();
  double ctr[3] = {bh_centre_.x, bh_centre_.y, bh_centre_.z};
  if (rrt_set_camera(ctx_, &cam) != RRT_OK) return finish("rrt_set_camera");
  if (rrt_set_frame(ctx_, (int)W_, (int)H_) != RRT_OK) return finish("rrt_set_frame");
  if (rrt_set_blackhole(ctx_, ctr, bh_r_, bh_dtheta_, bh_flat_ ? RRT_MODE_FLAT : RRT_MODE_LITERAL) != RRT_OK) return finish("rrt_set_blackhole");
  const bool multi = rrt_gpu_count(ctx_) > 1;   // a multi-GPU context deals the tiles over its own GPUs
  if (rrt_set_tile_shard(ctx_, tile_, multi ? 0 : rank_, multi ? 1 : world_) != RRT_OK) return finish("rrt_set_tile_shard");
  int rc = render_cell ? rrt_set_cell(ctx_, (int)cell_tl[0], (int)cell_tl[1], (int)(cell_br[0] - cell_tl[0]), (int)(cell_br[1] - cell_tl[1]))
                       : rrt_set_cell(ctx_, 0, 0, 0, 0);
  if (rc != RRT_OK) return finish("rrt_set_cell");
  rrt_params p; memset(&p, 0, sizeof p);
  p.ns_aa = (int32_t)ns_aa_; p.max_ray_depth = (int32_t)max_ray_depth_; p.ns_area_light = (int32_t)ns_area_light_;
  p.samples_per_batch = (int32_t)samples_per_batch_; p.max_tolerance = max_tolerance_; p.direct_hemisphere_sample = hemisphere_ ? 1 : 0;
  p.thin_lens = use_thin_lens ? 1 : 0; p.seed = seed_;
  snap_calls_ = 0; snaps_written_ = 0;
  // progressive output and stop() work between adaptive batches of ONE GPU; a multi-GPU context renders its tiles independently
  if (!multi && rrt_set_progress(ctx_, &PathTracer::progress_trampoline, this) != RRT_OK) return finish("rrt_set_progress");
  rrt_host_tick("worker: before rrt_render");
  if (rrt_render(ctx_, &p, hdr_.data(), counts_.data(), &stats_) != RRT_OK) return finish("rrt_render");
  rrt_host_tick("worker: rrt_render returned");
  if (rrt_tonemap(ctx_, (int64_t)(W_ * H_), hdr_.data(), frame_.data()) != RRT_OK) return finish("rrt_tonemap");
  finish(nullptr);
}

int PathTracer::progress_trampoline(void* self, int32_t done, int32_t total, int64_t active) {
  return static_cast<PathTracer*>(self)->on_progress(done, total, active);
}

// runs on the render thread between adaptive batches
int PathTracer::on_progress(int32_t done, int32_t total, int64_t active) {
  if (stop_requested_.load()) return 1;
  if (!snap_prefix_.empty() && ++snap_calls_ % snap_every_ == 0) {
    std::vector<float> hdr(W_ * H_ * 3);
    std::vector<uint32_t> rgba(W_ * H_), out(W_ * H_);
    if (rrt_snapshot(ctx_, hdr.data(), nullptr) == RRT_OK && rrt_tonemap(ctx_, (int64_t)(W_ * H_), hdr.data(), rgba.data()) == RRT_OK) {
      for (size_t i = 0; i < H_; ++i)
        for (size_t j = 0; j < W_; ++j) out[i * W_ + j] = rgba[(H_ - i - 1) * W_ + j] | 0xFF000000u;
      char name[64]; snprintf(name, sizeof name, "_%06d.png", (int)done);
      if (write_png(snap_prefix_ + name, out.data(), W_, H_)) {
        ++snaps_written_;
        fprintf(stderr, "[PathTracer] %d / %d samples, %lld pixels still sampling: %s%s\n", (int)done, (int)total, (long long)active, snap_prefix_.c_str(), name);
      }
    }
  }
  return 0;
}

void PathTracer::start_raytracing() {
  if (state_ != READY) return;
  if (thread_.joinable()) thread_.join();
  error_.clear();
  stop_requested_.store(false);
  state_ = RENDERING;
  fprintf(stdout, "[PathTracer] Rendering... "); fflush(stdout);
  thread_ = std::thread(&PathTracer::worker, this);
}

void PathTracer::stop() {
  stop_requested_.store(true);
  if (thread_.joinable()) thread_.join();
  stop_requested_.store(false);
  std::lock_guard<std::mutex> lk(m_done_);
  if (state_ == RENDERING || state_ == DONE) state_ = READY;
}

void PathTracer::clear() {
  if (state_ != READY) return;
  have_scene_ = false; camera_ = nullptr;
  hdr_.clear(); counts_.clear(); frame_.clear();
  state_ = INIT;
}

bool PathTracer::is_done() { std::lock_guard<std::mutex> lk(m_done_); return state_ == DONE; }
PathTracer::State PathTracer::get_state() { std::lock_guard<std::mutex> lk(m_done_); return state_; }

void PathTracer::render_to_file(std::string filename, size_t x, size_t y, size_t dx, size_t dy) {
  if (x == (size_t)-1) {
    render_cell = false;
  } else {
    render_cell = true;
    cell_tl[0] = (double)x; cell_tl[1] = (double)y; cell_br[0] = (double)(x + dx); cell_br[1] = (double)(y + dy);
  }
  if (state_ != READY) { if (error_.empty()) error_ = "render_to_file: scene, camera and frame size must be set first"; return; }
  start_raytracing();
  {
    std::unique_lock<std::mutex> lk(m_done_);
    cv_done_.wait(lk, [this] { return state_ == DONE; });
  }
  thread_.join();
  if (!error_.empty()) { fprintf(stdout, "failed: %s\n", error_.c_str()); return; }
  // the reference prints its own counters here (src/pathtracer.cpp:636-638); same line shapes, device numbers
  fprintf(stdout, "100%%! (%.4fs)\n", stats_.seconds);
  fprintf(stdout, "[PathTracer] BVH traced %llu rays.\n", (unsigned long long)stats_.rays);
  fprintf(stdout, "[PathTracer] Averaged %f intersection tests per ray.\n",
          stats_.rays ? (double)(stats_.tri_tests + stats_.sphere_tests) / (double)stats_.rays : 0.0);
  fprintf(stdout, "[PathTracer] %.3f Mpaths/s, %.3f G geodesic steps/s on the device.\n", stats_.seconds > 0 ? stats_.paths / stats_.seconds / 1e6 : 0.0,
          stats_.seconds > 0 ? stats_.geodesic_steps / stats_.seconds / 1e9 : 0.0);
  save_image(filename);
  fprintf(stdout, render_cell ? "[PathTracer] Cell job completed.\n" : "[PathTracer] Job completed.\n");
}

void PathTracer::save_image(std::string filename) {
  if (state_ != DONE) return;
  if (filename == "") {
    time_t t = time(nullptr);
    tm* lt = localtime(&t);
    std::stringstream ss;
    ss << filename_ << "_screenshot_" << lt->tm_mon + 1 << "-" << lt->tm_mday << "_" << lt->tm_hour << "-" << lt->tm_min << "-" << lt->tm_sec << ".png";
    filename = ss.str();
  }
  size_t x0 = 0, y0 = 0, w = W_, h = H_;
  if (render_cell) {   // raytrace_cell renders into a buffer of the cell's size (src/pathtracer.cpp:583-609)
    x0 = (size_t)cell_tl[0]; y0 = (size_t)cell_tl[1];
    w = (size_t)(cell_br[0] - cell_tl[0]); h = (size_t)(cell_br[1] - cell_tl[1]);
    if (x0 + w > W_) w = W_ > x0 ? W_ - x0 : 0;
    if (y0 + h > H_) h = H_ > y0 ? H_ - y0 : 0;
  }
  std::vector<uint32_t> out(w * h);
  for (size_t i = 0; i < h; ++i)
    for (size_t j = 0; j < w; ++j) out[i * w + j] = frame_[(y0 + h - i - 1) * W_ + x0 + j] | 0xFF000000u;
  fprintf(stderr, "[PathTracer] Saving to file: %s... ", filename.c_str());
  bool ok = write_png(filename, out.data(), w, h);
  fprintf(stderr, ok ? "Done!\n" : "FAILED\n");
  if (!ok) error_ = "cannot write " + filename;
  save_sampling_rate_image(filename);
}

void PathTracer::save_sampling_rate_image(std::string filename) {
  // reference src/pathtracer.cpp:686-717: rate = samples taken / ns_aa; blue (few) -> green -> red (all)
  const size_t w = W_, h = H_;
  std::vector<uint32_t> out(w * h, 0xFF000000u);
  auto chan = [](float c) { return (uint32_t)(std::min(1.f, c) * 255); };
  for (size_t x = 0; x < w; ++x)
    for (size_t y = 0; y < h; ++y) {
      float rate = counts_[y * w + x] * 1.0f / ns_aa_;
      float cr, cg, cb;
      if (rate <= 0.5) { float r = (0.5 - rate) / 0.5; cr = 0.f; cg = 1.0f * (float)(1.0 - r); cb = 1.0f * r; }
      else { float r = (1.0 - rate) / 0.5; cr = 1.0f * (float)(1.0 - r); cg = 1.0f * r; cb = 0.f; }
      out[x + (h - 1 - y) * w] = (chan(cb) << 16) | (chan(cg) << 8) | chan(cr) | 0xFF000000u;
    }
  if (filename.size() >= 4) write_png(filename.substr(0, filename.size() - 4) + "_rate.png", out.data(), w, h);
}

}  // namespace rrt_host
