// main.cpp — the reference's command line (src/main.cpp:28-40, 88-188) over the B200 path.
//
//   pathtracer [options] <scene.dae | scene.rrtscene>
//     -s <INT>  -l <INT>  -m <INT>  -t <INT>  -a <INT> <FLOAT>  -H  -r <W> <H>  -p <x> <y> <dx> <dy>  -f <out.png>
//     -b <FLOAT> -d <FLOAT> -c <camera settings file>  -B <x> <y> <z> <r> <dtheta>        exactly as the reference
//   additions (not in the reference): -F (flat spacetime, the reference has no working switch), -S <seed>,
//     -G <device>, --dump-scene <out.rrtscene> (load + flatten only: needs no GPU), --dump-camera <file>
//   -e <map.exr | map.pfm>: scan-line OpenEXR (NONE / RLE / ZIPS / ZIP / PIZ) or PFM.
//   not available: the interactive viewer (no graphics stack on this box): without -f the program renders to
//     <scene>_screenshot_<time>.png instead of opening a window.
// -t is accepted and ignored: the device replaces the worker threads.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <thread>

#include "rrt_host.h"

using namespace rrt_host;

static void usage(const char* binary) {
  printf("Usage: %s [options] <scenefile>\n", binary);
  printf("Program Options:\n");
  printf("  -s  <INT>        Number of camera rays per pixel\n");
  printf("  -l  <INT>        Number of samples per area light\n");
  printf("  -t  <INT>        Number of render threads (ignored: the GPU renders)\n");
  printf("  -m  <INT>        Maximum ray depth\n");
  printf("  -a  <INT> <FLOAT> Samples per batch and tolerance for adaptive sampling\n");
  printf("  -H               Enable uniform hemisphere sampling for direct lighting\n");
  printf("  -f  <FILENAME>   Image (.png) file to save output to\n");
  printf("  -r  <INT> <INT>  Width and height of output image\n");
  printf("  -p  <x> <y> <dx> <dy> Render only this cell of the frame\n");
  printf("  -c  <FILENAME>   Camera settings file\n");
  printf("  -b  <FLOAT>      Lens radius      -d  <FLOAT>  Focal distance\n");
  printf("  -e  <FILENAME>   Environment map (uncompressed .exr or .pfm)\n");
  printf("  -B  <x> <y> <z> <r> <dtheta>  Black hole centre, Schwarzschild radius, angular step\n");
  printf("  -F               Flat spacetime (no hole)      -S <INT> RNG seed      -G <INT[,INT...]> CUDA device(s)\n");
  printf("  --progress <PREFIX> [--progress-every <INT>]  write <PREFIX>_<samples>.png after every n-th adaptive batch\n");
  printf("  --thin-lens      sample the lens (-b, -d) for every camera ray\n");
  printf("  --stepper <fp64|fp32>   geodesic stepper: fp64 = bit-identical to the reference (default), fp32 = within 1e-4, faster\n");
  printf("  --bvh-builder <auto|host|device>   where the BVH is built (default auto: device for >= 4096 primitives)\n");
  printf("  --dump-scene <FILE>  write the flattened scene and exit (no GPU needed)\n");
  printf("  -h               Print this help message\n\n");
}

int main(int argc, char** argv) {
  rrt_host_tick("main");
  size_t ns_aa = 1, max_ray_depth = 1, ns_area_light = 1, num_threads = 1, samples_per_batch = 32;
  float max_tolerance = 0.05f;
  bool hemisphere = false, write_to_file = false, flat = false;
  double lens_radius = 0.25, focal_distance = 4.7;
  size_t w = 0, h = 0, x = (size_t)-1, y = 0, dx = 0, dy = 0;
  std::string filename, cam_settings, dump_scene, dump_camera, scene_path, envmap_path;
  Vector3D bh_c(0, 1, 0); double bh_r = 0.1, bh_dtheta = 0.1;
  unsigned long long seed = 0; int device = 0; std::vector<int> devices;
  std::string progress_prefix; int progress_every = 1; bool thin_lens = false; bool stepper_fp32 = false; int bvh_builder = 0;

  auto need = [&](int i, int k) { if (i + k >= argc) { usage(argv[0]); exit(1); } };
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    if (a == "-f") { need(i, 1); write_to_file = true; filename = argv[++i]; }
    else if (a == "-r") { need(i, 2); w = (size_t)atoi(argv[i + 1]); h = (size_t)atoi(argv[i + 2]); i += 2; }
    else if (a == "-p") { need(i, 4); x = (size_t)atoi(argv[i + 1]); y = (size_t)atoi(argv[i + 2]); dx = (size_t)atoi(argv[i + 3]); dy = (size_t)atoi(argv[i + 4]); i += 4; }
    else if (a == "-s") { need(i, 1); ns_aa = (size_t)atoi(argv[++i]); }
    else if (a == "-l") { need(i, 1); ns_area_light = (size_t)atoi(argv[++i]); }
    else if (a == "-t") { need(i, 1); num_threads = (size_t)atoi(argv[++i]); }
    else if (a == "-m") { need(i, 1); max_ray_depth = (size_t)atoi(argv[++i]); }
    else if (a == "-b") { need(i, 1); lens_radius = atof(argv[++i]); }
    else if (a == "-d") { need(i, 1); focal_distance = atof(argv[++i]); }
    else if (a == "-c") { need(i, 1); cam_settings = argv[++i]; }
    else if (a == "-a") { need(i, 2); samples_per_batch = (size_t)atoi(argv[i + 1]); max_tolerance = (float)atof(argv[i + 2]); i += 2; }
    else if (a == "-H") hemisphere = true;
    else if (a == "-B") {
      need(i, 5);
      bh_c = Vector3D(atof(argv[i + 1]), atof(argv[i + 2]), atof(argv[i + 3])); bh_r = atof(argv[i + 4]); bh_dtheta = atof(argv[i + 5]);
      i += 5;
    }
    else if (a == "-F") flat = true;
    else if (a == "-S") { need(i, 1); seed = strtoull(argv[++i], nullptr, 10); }
    else if (a == "-G") {   // one CUDA device, or a comma-separated list: the frame's tiles are dealt over those GPUs
      need(i, 1);
      devices.clear();
      for (const char* q = argv[++i]; *q;) { devices.push_back(atoi(q)); while (*q && *q != ',') ++q; if (*q == ',') ++q; }
      if (devices.empty()) { usage(argv[0]); return 1; }
      device = devices[0];
    }
    else if (a == "--progress") { need(i, 1); progress_prefix = argv[++i]; }
    else if (a == "--progress-every") { need(i, 1); progress_every = atoi(argv[++i]); }
    else if (a == "--thin-lens") thin_lens = true;
    else if (a == "--stepper") { need(i, 1); std::string v = argv[++i]; if (v != "fp64" && v != "fp32") { usage(argv[0]); return 1; } stepper_fp32 = v == "fp32"; }
    else if (a == "--bvh-builder") { need(i, 1); std::string v = argv[++i]; bvh_builder = v == "device" ? 1 : v == "host" ? 2 : v == "auto" ? 0 : -1; if (bvh_builder < 0) { usage(argv[0]); return 1; } }
    else if (a == "--dump-scene") { need(i, 1); dump_scene = argv[++i]; }
    else if (a == "--dump-camera") { need(i, 1); dump_camera = argv[++i]; }
    else if (a == "--write-test-png") {   // host self-check used by tests/test_host_cpp.py: a 5x3 gradient through write_png
      need(i, 1);
      uint32_t px[15];
      for (uint32_t k = 0; k < 15; ++k) px[k] = 0xFF000000u | (k * 17u) | ((255u - k * 17u) << 8) | ((k * 5u) << 16);
      return write_png(argv[i + 1], px, 5, 3) ? 0 : 2;
    }
    else if (a == "-e") { need(i, 1); envmap_path = argv[++i]; }
    else if (a == "--dump-envmap") {      // host self-check: decode a map, write int32 w, h + floats (tests/test_host_cpp.py)
      need(i, 2);
      int ew = 0, eh = 0; std::vector<float> env; std::string e;
      if (!load_envmap(argv[i + 1], &ew, &eh, &env, &e)) { fprintf(stderr, "%s\n", e.c_str()); return 2; }
      FILE* f = fopen(argv[i + 2], "wb"); if (!f) return 2;
      int32_t wh[2] = {ew, eh};
      fwrite(wh, 4, 2, f); fwrite(env.data(), 4, env.size(), f); fclose(f);
      return 0;
    }
    else if (a == "-h") { usage(argv[0]); return 1; }
    else if (!a.empty() && a[0] == '-') { usage(argv[0]); return 1; }
    else scene_path = a;
  }
  if (scene_path.empty()) { usage(argv[0]); return 1; }
  (void)num_threads;

  printf("[PathTracer] Input scene file: %s\n", scene_path.c_str());
  std::string scene_file = scene_path.substr(scene_path.find_last_of('/') + 1);
  scene_file = scene_file.substr(0, scene_file.find(".dae"));

  // The first CUDA call of a process (driver initialisation + primary context) takes 1.2 - 2.5 s on a B200 box — 20x everything else a C2 frame
  // needs (profiles/r02_j_cli_timing.txt) — and needs nothing from the scene: it runs on a thread of its own while the .dae is parsed and flattened.
  rrt_ctx* warm_ctx = nullptr; std::string warm_err;
  std::thread warm;
  if (dump_scene.empty())
    warm = std::thread([&]() {
      const int rc = devices.size() > 1 ? rrt_create_multi(&warm_ctx, (int)devices.size(), devices.data()) : rrt_create(&warm_ctx, device);
      if (rc != RRT_OK) { warm_err = rrt_last_error(nullptr); warm_ctx = nullptr; }
      rrt_host_tick("context ready (warm-up thread)");
    });
  struct Joiner { std::thread& t; ~Joiner() { if (t.joinable()) t.join(); } } joiner{warm};
  FlatScene scene; Camera camera; std::string err;
  bool is_dae = scene_path.size() >= 4 && scene_path.compare(scene_path.size() - 4, 4, ".dae") == 0;
  bool ok = is_dae ? load_dae(scene_path, &scene, &camera, &err) : load_rrtscene(scene_path, &scene, &camera, &err);
  if (!ok) { fprintf(stderr, "[PathTracer] %s\n", err.c_str()); return is_dae ? 0 : 2; }   // the reference exits 0 on a missing file
  if (w && h) camera.set_screen_size(w, h);
  if (!cam_settings.empty() && !camera.load_settings(cam_settings)) { fprintf(stderr, "[Camera] cannot load settings from %s\n", cam_settings.c_str()); return 2; }
  if (!cam_settings.empty()) printf("[Camera] Loaded settings from %s\n", cam_settings.c_str());

  if (!dump_camera.empty()) { if (!camera.dump_settings(dump_camera)) return 2; printf("[Camera] Dumped settings to %s\n", dump_camera.c_str()); }
  if (!dump_scene.empty()) {
    // like PathTracer::set_camera, the dump carries the lens parameters of the command line
    camera.lensRadius = lens_radius; camera.focalDistance = focal_distance;
    if (!save_rrtscene(dump_scene, scene, camera, &err)) { fprintf(stderr, "%s\n", err.c_str()); return 2; }
    printf("[PathTracer] %s: %zu primitives (%zu triangles, %zu spheres), %zu materials, %zu lights\n", dump_scene.c_str(), scene.n_prims(),
           scene.tri_mat.size(), scene.sph_mat.size(), scene.mats.size(), scene.lights.size());
    return 0;
  }

  rrt_host_tick("scene file loaded and flattened");
  PathTracer pt(ns_aa, max_ray_depth, ns_area_light, 1, 1, 1, num_threads, samples_per_batch, max_tolerance, nullptr, hemisphere,
                scene_file, lens_radius, focal_distance);
  if (warm.joinable()) warm.join();
  if (!warm_ctx && !warm_err.empty()) { fprintf(stderr, "[PathTracer] %s\n", warm_err.c_str()); return 3; }
  pt.set_device(device); if (devices.size() > 1) pt.set_devices(devices);
  pt.adopt_context(warm_ctx);
  pt.set_seed(seed); pt.set_blackhole(bh_c, bh_r, bh_dtheta, flat);
  pt.set_option(RRT_OPT_STEPPER, stepper_fp32 ? 1 : 0); pt.set_option(RRT_OPT_BVH_BUILDER, bvh_builder);
  if (!envmap_path.empty()) {
    int ew = 0, eh = 0; std::vector<float> env;
    if (!load_envmap(envmap_path, &ew, &eh, &env, &err)) { fprintf(stderr, "[PathTracer] Error loading environment map: %s\n", err.c_str()); return 1; }
    pt.set_envmap(ew, eh, env.data());
  }
  pt.set_progress_snapshots(progress_prefix, progress_every); pt.use_thin_lens = thin_lens;
  pt.set_camera(&camera);
  pt.set_scene(&scene);
  pt.set_frame_size(camera.screenW, camera.screenH);
  if (!pt.error().empty()) { fprintf(stderr, "[PathTracer] %s\n", pt.error().c_str()); return 3; }
  pt.render_to_file(write_to_file ? filename : std::string(""), x, y, dx, dy);
  rrt_host_tick("render_to_file returned (PNG written)");
  if (!pt.error().empty()) { fprintf(stderr, "[PathTracer] %s\n", pt.error().c_str()); return 3; }
  return 0;
}
