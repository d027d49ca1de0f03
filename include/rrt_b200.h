/* rrt_b200.h — C-ABI of the B200-native hot path of the relativistic ray tracer.
 *
 * One data-parallel path is served: per-pixel Monte Carlo radiance estimation in which every ray
 * (camera, bounce, shadow) is integrated as a Schwarzschild null geodesic in short chords and each
 * chord is tested against the BVH.  Everything below is plain C: pointers and sizes, no C++/torch
 * types, no exceptions.  All functions return 0 on success or a negative RRT_E_* code; the message
 * is available from rrt_last_error().  There is no CPU fallback: without a CUDA device every
 * compute entry point fails with RRT_E_CUDA.
 *
 * A context (rrt_ctx) is driven by one host thread at a time.  rrt_create binds it to ONE GPU;
 * rrt_create_multi makes one context drive several GPUs of the box from one process (tiles dealt
 * over the GPUs, scene replicated and the frame combined with NCCL inside the library).  Multi-
 * process runs (one context per rank) use rrt_set_tile_shard and combine the frames with a single
 * sum-reduce themselves — see INTEGRATION.md.
 *
 * Reference interfaces replaced (paths relative to /root/reference/pathtracer):
 *   PathTracer::set_scene  + build_accel   src/pathtracer.cpp:95-117,304-328   -> rrt_upload_scene
 *   BVHAccel::BVHAccel / construct_bvh     src/bvh.cpp:14-19,49-96             -> rrt_upload_scene
 *   PathTracer::set_camera                 src/pathtracer.cpp:119-134          -> rrt_set_camera
 *   global_black_hole / -B flag            src/static_scene/blackhole.cpp:5, src/main.cpp:139-145 -> rrt_set_blackhole
 *   PathTracer::set_frame_size             src/pathtracer.cpp:136-149          -> rrt_set_frame
 *   render_cell / cell_tl / cell_br        src/pathtracer.cpp:256-274          -> rrt_set_cell
 *   start_raytracing + worker_thread + raytrace_tile + raytrace_pixel
 *                                          src/pathtracer.cpp:224-282,549-644; src/part1_code.cpp:125-163 -> rrt_render[_device]
 *   BVHAccel::intersect(ray[,isect])       src/bvh.h:86-109, src/bvh.cpp:99-138 -> rrt_trace_rays
 *   BlackHole::next_micro_ray              src/static_scene/blackhole.cpp:17-40 -> rrt_geodesic
 *   Camera::generate_ray                   src/part1_code.cpp:182-187           -> rrt_generate_rays
 *   Camera::generate_ray_for_thin_lens     src/camera.cpp:176-184               -> rrt_generate_rays_thin_lens
 *   BSDF::f / sample_f                     src/bsdf.cpp:33-171, src/part1_code.cpp:167-173 -> rrt_bsdf_sample, rrt_bsdf_eval
 *   SceneLight::sample_L                   src/static_scene/light.cpp:17-92     -> rrt_light_sample
 *   HDRImageBuffer::toColor                src/image.h:183-198,53-62            -> rrt_tonemap
 *   update_screen (RENDERING) / stop       src/pathtracer.cpp:156-200           -> rrt_set_progress, rrt_snapshot
 */
#ifndef RRT_B200_H
#define RRT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RRT_ABI_VERSION 5

/* error codes */
#define RRT_OK 0
#define RRT_E_ARG (-1)     /* bad argument / wrong state */
#define RRT_E_CUDA (-2)    /* CUDA runtime error or no device */
#define RRT_E_STATE (-3)   /* scene/camera/frame not set */
#define RRT_E_IO (-4)      /* scene file problem */

typedef struct rrt_ctx rrt_ctx;

/* BSDF kinds (reference classes in src/bsdf.h:118-258) */
enum {
  RRT_MAT_DIFFUSE = 0,    /* a = reflectance */
  RRT_MAT_MIRROR = 1,     /* a = reflectance */
  RRT_MAT_GLASS = 2,      /* a = reflectance, b = transmittance, ior */
  RRT_MAT_MICROFACET = 3, /* a = eta, b = k, alpha */
  RRT_MAT_EMISSION = 4,   /* a = radiance */
  RRT_MAT_REFRACTION = 5  /* b = transmittance, ior; reference stub: f = 0, sample_f = 0 (src/bsdf.cpp:100-106) */
};

typedef struct rrt_material {
  int32_t type;
  float a[3];
  float b[3];
  float alpha;
  float ior;
} rrt_material; /* 36 bytes */

/* light kinds (reference classes in src/static_scene/light.h) */
enum {
  RRT_LIGHT_DIRECTIONAL = 0, /* direction = dirToLight (unit, already negated) */
  RRT_LIGHT_HEMISPHERE = 1,  /* InfiniteHemisphereLight: fixed sampleToWorld (src/static_scene/light.cpp:27-32) */
  RRT_LIGHT_POINT = 2,       /* position */
  RRT_LIGHT_AREA = 3,        /* position, direction, dim_x, dim_y, area */
  RRT_LIGHT_NULL = 4         /* Spot/Sphere/Mesh lights: reference stubs returning zero radiance (light.cpp:61-115) */
};

typedef struct rrt_light {
  int32_t type;
  float radiance[3];
  double position[3];
  double direction[3];
  double dim_x[3];
  double dim_y[3];
  float area;
  int32_t is_delta; /* SceneLight::is_delta_light() */
} rrt_light; /* 120 bytes */

/* Scene in the reference's primitive order (the order PathTracer::build_accel collects them,
 * src/pathtracer.cpp:311-316).  Geometry is the reference's own doubles.  The library copies
 * everything; the caller keeps ownership. */
typedef struct rrt_scene {
  uint32_t n_prims;
  const uint8_t* prim_kind;   /* [n_prims] 0 = triangle, 1 = sphere */
  const uint32_t* prim_index; /* [n_prims] index into the triangle or sphere arrays */
  uint32_t n_tris;
  const double* tri_pos;  /* [n_tris*9] p0,p1,p2 in Triangle's (v1,v2,v3) order */
  const double* tri_nrm;  /* [n_tris*9] vertex normals n0,n1,n2 (not necessarily unit) */
  const int32_t* tri_mat; /* [n_tris] */
  uint32_t n_spheres;
  const double* sph;      /* [n_spheres*4] centre xyz, radius */
  const int32_t* sph_mat; /* [n_spheres] */
  uint32_t n_mats;
  const rrt_material* mats;
  uint32_t n_lights;
  const rrt_light* lights;
} rrt_scene;

/* Camera after Application::load/resize (src/camera.cpp:22-40,68-74,96-119).  c2w holds the three
 * COLUMNS of the reference Matrix3x3 (c2w[3*j+i] = entries[j][i]).  Fov in degrees. */
typedef struct rrt_camera {
  double pos[3];
  double c2w[9];
  double hFov, vFov;
  double nClip, fClip;
  double lensRadius, focalDistance; /* used only with rrt_params.thin_lens */
} rrt_camera;

#define RRT_MODE_FLAT 0    /* BVHAccel::intersect_micro semantics: straight rays, min_t/max_t honoured */
#define RRT_MODE_LITERAL 1 /* reference-literal: every ray bent, caller's min_t/max_t discarded (src/bvh.cpp:103-113) */

typedef struct rrt_params {
  int32_t ns_aa;            /* -s */
  int32_t max_ray_depth;    /* -m */
  int32_t ns_area_light;    /* -l */
  int32_t samples_per_batch;/* -a batch (default 32) */
  float max_tolerance;      /* -a tol (default 0.05) */
  int32_t direct_hemisphere_sample; /* -H */
  int32_t thin_lens;        /* THIN_LENS compile switch of the reference, run-time here */
  uint64_t seed;            /* Philox key */
  int32_t max_paths_in_flight; /* 0 = library default; wavefront chunk size */
  int32_t reserved;
} rrt_params;

typedef struct rrt_stats {
  double seconds;          /* device time of the whole render (CUDA events) */
  double trace_seconds;    /* device time inside the geodesic+BVH trace kernels */
  uint64_t paths;          /* est_radiance_global_illumination calls (= sum of sample counts) */
  uint64_t rays;           /* BVHAccel::intersect calls (camera + bounce + shadow) */
  uint64_t geodesic_steps; /* BlackHole::next_micro_ray calls (literal mode) or straight segments (flat) */
  uint64_t node_visits;    /* BVH node records fetched: 64 bytes each (binary nodes) or 32 bytes per child slot (wide nodes: 256) */
  uint64_t tri_tests;      /* 80-byte triangle records tested */
  uint64_t sphere_tests;   /* sphere records tested */
  uint64_t kernel_launches;/* kernels of this library launched during the render */
  uint64_t trace_launches; /* of which trace kernels */
  uint64_t algorithmic_bytes; /* node record bytes*node_visits + 80*tri_tests + 32*sphere_tests + 160*rays */
} rrt_stats;

/* Result of one BVHAccel::intersect(ray, &isect) call. */
typedef struct rrt_hit {
  int32_t status; /* 0 = miss, 1 = hit, 2 = swallowed by the hole (reference returns false) */
  int32_t prim;   /* index in the reference primitive order, -1 if none */
  int32_t chord;  /* index of the geodesic chord that produced the result (0 in flat mode) */
  int32_t steps;  /* next_micro_ray calls made for this ray */
  double t;       /* ray parameter within that chord (Ray::max_t after the hit) */
  double hit_p[3];
  double w_out[3];
  double n[3];    /* as stored by the reference: un-normalised for triangles */
} rrt_hit; /* 96 bytes */

/* tuning / instrumentation switches for rrt_set_option */
enum {
  RRT_OPT_BVH_TOPOLOGY = 0, /* 0 = the reference's midpoint-split tree shape, 1 = binned SAH (default); takes effect at the next rrt_upload_scene */
  RRT_OPT_COUNTERS = 1,     /* 1 (default) = trace kernels count rays / chords / node visits / primitive tests for rrt_stats */
  RRT_OPT_WORK_BYTES = 2,   /* upper bound for the wavefront work buffers in bytes (default 12 GiB) */
  RRT_OPT_TRACE_KERNEL = 3, /* trace kernel variant: 0 = chord-lockstep loop (default).  A/B builds of the library only (make EXTRA=-DRRT_AB=1; otherwise
                               RRT_E_ARG): 1 = vote-scheduled state machine, 2 = decoupled stepping, both measured slower */
  RRT_OPT_LEAF_SIZE = 4,    /* primitives per leaf for topology 1 (1..7, default 2); takes effect at the next rrt_upload_scene */
  RRT_OPT_CLEARANCE_GRID = 5, /* cells along the longest axis of the empty-space grid (0 = off, the default); next rrt_upload_scene */
  RRT_OPT_BVH_WIDTH = 6,      /* children per node on the device: 2, 8 (4 in A/B builds only), or 0 (default) = 8 for scenes of <= 128 primitives, else 2; topology 0 and trace kernel 1 always use 2; next rrt_upload_scene */
  RRT_OPT_SHADE_SPLIT = 7,    /* direct lighting: 1 = its own kernel, one thread per light sample; 0 = inside k_shade's per-vertex loop; 2 (default) = 1 when the scene has a microfacet material, else 0 */
  RRT_OPT_STEPPER = 8,        /* geodesic stepper with the hole on: 0 (default) = the reference's recurrence in IEEE double, bit-identical chords;
                                 1 = the same recurrence carried as (orbital frame, u, u') in FP32 registers (BASELINE.json north_star item 3):
                                 chords within ~1e-6 of the reference's (bound asserted: 1e-4), NOT bit-identical; trace kernel 0 only */
  RRT_OPT_STREAMS = 10,       /* CUDA streams the wavefront chunks of one adaptive round are dealt over (1..4, default 2): a trace launch ends ragged
                                 (rays need 1..63 chords) and another stream's kernels fill the idle SMs; bit-identical results for any value */
  RRT_OPT_ESCAPE = 11,        /* A/B builds only (otherwise RRT_E_ARG for a non-zero value): 1 = rays that have left the scene box are parked by the trace kernel
                                 and carried on by a lean kernel (stepper + horizon + root-box test only); measured slower, default 0.  Bit-identical results */
  RRT_OPT_TOP_NODE = 12,      /* binary trees: 1 (default) = the top of the tree is also kept as ONE 8-wide record (groupings of large primitives opened into
                                 their flat boxes + the root of the rest) that every traversal tests first; 0 = start at the binary root.  Same results; next rrt_upload_scene */
  RRT_OPT_OVERLAP = 13,       /* 1 (default) = the shadow rays of depth i are traced on their own stream beside the extension rays of depths i + 1 and i + 2 (they only feed
                                 k_accumulate(i), which makes every addition into a path's radiance, in vertex order; the shadow side of the queues is double-buffered and
                                 k_shade(i + 2) waits for k_accumulate(i)); 0 = one kernel at a time per chunk.  Bit-identical */
  RRT_OPT_SORT_SHADOW = 14,   /* hole on, meshes (binary nodes): the shadow rays of a launch are handed to the trace kernel cell by cell of their origin (16^3 grid
                                 over the scene box, Morton order; a counting sort on the device builds the launch order, the rays stay where they are), so
                                 that the rays a warp holds need the same part of the tree and end together.  1 (default) = in chunks of fewer than 2 M paths,
                                 where the shorter ragged end shows (an 8-GPU rank's share of C4: -2.3 %; C4 on one GPU: +-0); 2 = in every chunk
                                 (C4 on one GPU: +0.5 %); 0 = never, queue order.  Bit-identical */
  RRT_OPT_BVH_BUILDER = 9     /* where the BVH is built.  1 = on the device: the reference's depth-first ranks by level-synchronous stable partition
                                 + an LBVH for the traversal (scenes of >= 64 primitives, binary nodes; falls back to the host if the tree comes
                                 out deeper than the traversal stack); 2 = on the host (reference tree shape or binned SAH, threaded);
                                 0 (default) = device for scenes of >= 4096 primitives unless a host-tree option (topology, leaf size, width) is set.
                                 Same images either way; next rrt_upload_scene */
};

/* ---- lifecycle ---- */
int rrt_abi_version(void);
int rrt_device_count(void);
int rrt_create(rrt_ctx** out, int device);
/* One context driving n_gpus GPUs of this box from one process (devices == NULL: 0 .. n_gpus-1).  Replaces the reference's pool of N worker
 * threads pulling tiles from one queue (src/pathtracer.cpp:224-282, 611-644): GPU k renders the interleaved 8x8 tiles it owns
 * ((tx + ty * s) % n_gpus == k, see rrt_set_tile_shard), the scene is built once on the first GPU and replicated with ncclBroadcast,
 * and ONE ncclReduce per frame sums the per-GPU frames onto the first GPU, where rrt_render reads them back (rrt_render_device: d_rgbn /
 * d_counts must live on the first GPU).  Every other call behaves as on a single-GPU context (configuration goes to every GPU, parity
 * hooks run on the first); rrt_set_progress / rrt_snapshot are not available.  Results are bit-identical to a single-GPU context.
 * NCCL (libnccl.so.2) is loaded at this call; n_gpus == 1 is rrt_create. */
int rrt_create_multi(rrt_ctx** out, int n_gpus, const int* devices);
int rrt_gpu_count(const rrt_ctx* ctx);
void rrt_destroy(rrt_ctx* ctx);
const char* rrt_last_error(const rrt_ctx* ctx); /* ctx may be NULL: last creation error */
int rrt_set_option(rrt_ctx* ctx, int option, int64_t value);

/* ---- configuration (host buffers) ---- */
int rrt_upload_scene(rrt_ctx* ctx, const rrt_scene* scene);
int rrt_set_camera(rrt_ctx* ctx, const rrt_camera* cam);
/* centre, r finite, r >= 0; delta_theta finite and >= 2 pi / 65536 (RRT_E_ARG otherwise: with a NaN or non-positive step the reference's
 * loop `j * delta_theta < 2 pi`, src/bvh.cpp:105, never ends) */
int rrt_set_blackhole(rrt_ctx* ctx, const double centre[3], double r, double delta_theta, int mode);
int rrt_set_frame(rrt_ctx* ctx, int width, int height);
/* cell render (-p x y dx dy); w <= 0 restores the full frame; w > 0 needs h > 0 */
int rrt_set_cell(rrt_ctx* ctx, int x, int y, int w, int h);
/* Environment light, the reference's -e <map.exr> (src/main.cpp:42-79, src/static_scene/environment_light.cpp):
 * rgb = [height*width*3] floats, row 0 = top of the lat-long map (theta = 0 is +y).  The light is appended after the
 * scene's lights, is importance sampled for direct lighting, and camera rays that hit nothing (or fall into the
 * hole) return the map's radiance along their own, unbent direction.  rgb == NULL removes it.  Survives
 * rrt_upload_scene.  Maps with an all-black row are not usable (the reference divides by that row's marginal). */
int rrt_set_envmap(rrt_ctx* ctx, int width, int height, const float* rgb);
/* interleaved tile ownership for multi-GPU: tile (tx, ty) belongs to rank (tx + ty * s) % world, s = the smallest integer >= round(sqrt(world))
 * coprime with world (8 -> 3, 4 -> 3, 2 -> 1): every run of `world` tiles along a row or a column holds every rank once */
int rrt_set_tile_shard(rrt_ctx* ctx, int tile_size, int rank, int world);

/* ---- the hot path ---- */
/* Blocking.  hdr_rgb: [W*H*3] float, row 0 = image bottom (sampleBuffer layout, src/pathtracer.cpp:575);
 * sample_counts: [W*H] int (sampleCountBuffer).  Pixels not owned by this shard are written as 0.
 * Either output may be NULL.  Host buffers; the copies are inside the call. */
int rrt_render(rrt_ctx* ctx, const rrt_params* p, float* hdr_rgb, int32_t* sample_counts, rrt_stats* stats);
/* Same, results left on the device: d_rgbn is a CUDA device pointer to [W*H] float4 (mean r,g,b, sample
 * count as float), d_counts a device pointer to [W*H] int32 or NULL.  BLOCKING: the call first synchronises
 * `stream` (a cudaStream_t passed as void*, may be NULL: work the caller queued there, e.g. clearing d_rgbn, is
 * finished before the render starts), renders on the context's own non-blocking streams (the adaptive rounds need host
 * read-backs), and returns when d_rgbn / d_counts are complete — any stream may read them afterwards. */
int rrt_render_device(rrt_ctx* ctx, const rrt_params* p, void* d_rgbn, void* d_counts, void* stream, rrt_stats* stats);

/* ---- parity hooks: the same device code the render uses, on caller-supplied inputs ---- */
/* n rays, o/d: [n*3] doubles; min_t/max_t may be NULL (0, +inf); used only in flat mode. any_hit != 0
 * gives BVHAccel::intersect(ray) (shadow-ray form): only status is meaningful. */
int rrt_trace_rays(rrt_ctx* ctx, int64_t n, const double* o, const double* d, const double* min_t,
                   const double* max_t, int any_hit, rrt_hit* out, rrt_stats* stats);
/* rrt_trace_rays without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same kernel, same arguments and results as rrt_trace_rays (no rrt_stats); o, d, min_t, max_t and out may be
 * host or device pointers.  All device scratch — staged rays, hit records, the work cursor, the counters — lives in one
 * file-scope `static __device__` array of 4 KiB + RRT_STATIC_MAX_RAYS * 160 B (10 MiB) that every GPU loading the
 * library holds until exit; the call makes no cudaMalloc / cudaFree / allocating library call (the scene must have
 * been uploaded before, which does allocate).  n > RRT_STATIC_MAX_RAYS fails with RRT_E_ARG and a message naming the
 * limit.  Every call (from any context on that GPU) shares the array and finds the previous call's contents in it;
 * the call clears what must start at zero, but calls that overlap on one GPU RACE: serialise them. */
#define RRT_STATIC_MAX_RAYS 65536
int rrt_trace_rays_static(rrt_ctx* ctx, int64_t n, const double* o, const double* d, const double* min_t,
                          const double* max_t, int any_hit, rrt_hit* out);
/* First nsteps chords of each ray: seg_out [n*nsteps*8] = origin xyz, unit direction xyz, length, and 1.0 when the
 * chord crosses the horizon (the reference's capture test, src/bvh.cpp:107), else 0.0. */
int rrt_geodesic(rrt_ctx* ctx, int64_t n, const double* o, const double* d, int nsteps, double* seg_out);
/* Camera::generate_ray for n normalised sensor positions xy [n*2] -> o,d [n*3]. */
int rrt_generate_rays(rrt_ctx* ctx, int64_t n, const double* xy, double* o, double* d);
/* Camera::generate_ray_for_thin_lens (src/camera.cpp:176-184; what raytrace_pixel calls with useThinLens, -b / -d):
 * lens [n*2] = (rndR, rndTheta) as the reference passes them (a [0,1) sample and a [0,2 pi) angle). */
int rrt_generate_rays_thin_lens(rrt_ctx* ctx, int64_t n, const double* xy, const double* lens, double* o, double* d);
/* BSDF::sample_f of material `mat` for n outgoing directions wo [n*3] with uniforms u [n*4] in [0,1)
 * (u0,u1 = 2-D sample, u2 = glass coin).  Outputs f [n*3] float, wi [n*3] double, pdf [n] float. */
int rrt_bsdf_sample(rrt_ctx* ctx, int mat, int64_t n, const double* wo, const double* u, float* f,
                    double* wi, float* pdf);
/* BSDF::f */
int rrt_bsdf_eval(rrt_ctx* ctx, int mat, int64_t n, const double* wo, const double* wi, float* f);
/* SceneLight::sample_L of light `light` from n points p [n*3] with uniforms u [n*2]:
 * radiance [n*3] float, wi [n*3] double, dist [n] float, pdf [n] float. */
int rrt_light_sample(rrt_ctx* ctx, int light, int64_t n, const double* p, const double* u,
                     float* radiance, double* wi, float* dist, float* pdf);
/* Philox4x32-10 block for (key = seed, counter = c0..c3) as the render uses it: out [n*4] uint32. */
/* EnvironmentLight::sample_dir for n directions dir [n*3] -> radiance [n*3]; the light itself is sampled through
 * rrt_light_sample with light == the scene's light count. */
int rrt_env_lookup(rrt_ctx* ctx, int64_t n, const double* dir, float* radiance);
int rrt_philox(rrt_ctx* ctx, uint64_t seed, int64_t n, const uint32_t* counters, uint32_t* out);
/* toColor: [n*3] float HDR -> [n] RGBA8 (alpha 0xFF), on the device. */
int rrt_tonemap(rrt_ctx* ctx, int64_t n, const float* hdr_rgb, uint32_t* rgba);

/* ---- progressive output ----
 * Stands in for what the viewer shows while the reference is in its RENDERING state (the frame buffer fills tile by
 * tile, src/pathtracer.cpp:156-178) and for PathTracer::stop() (:180-200, `continueRaytracing = false`).
 * fn runs on the thread that called rrt_render[_device], after every adaptive batch (samples_per_batch samples of every
 * still-active pixel) except the last; samples_done counts samples per still-active pixel.  Inside fn, rrt_snapshot
 * copies the running estimate (sum / n per pixel, 0 where n == 0) and the per-pixel sample counts to the host.  A
 * non-zero return abandons the remaining batches: rrt_render then returns RRT_OK with the estimate reached so far.
 * fn == NULL removes the callback.  Without a callback the render makes no extra synchronisation. */
typedef int (*rrt_progress_fn)(void* user, int32_t samples_done, int32_t samples_total, int64_t active_pixels);
int rrt_set_progress(rrt_ctx* ctx, rrt_progress_fn fn, void* user);
/* only valid inside an rrt_progress_fn call; either pointer may be NULL.  hdr_rgb [W*H*3], sample_counts [W*H]. */
int rrt_snapshot(rrt_ctx* ctx, float* hdr_rgb, int32_t* sample_counts);

/* ---- introspection ---- */
/* BVH facts after rrt_upload_scene: nodes, leaves, depth; dfs_rank [n_prims] (may be NULL) receives, for
 * every primitive in reference order, its position in the reference BVH's depth-first leaf order. */
int rrt_bvh_info(rrt_ctx* ctx, int32_t* n_nodes, int32_t* n_leaves, int32_t* depth, uint32_t* dfs_rank);
/* Host-only (no device needed): runs the BVH construction rrt_upload_scene uses.  info[0..7] = reference-shape nodes,
 * leaves, depth, flattened interior nodes, leaf records, root child code, interior levels of the flattened tree, the
 * device traversal stack size those levels must stay below; dfs_rank as in rrt_bvh_info (may be NULL). */
int rrt_bvh_build_host(const rrt_scene* scene, int topology, int32_t info[8], uint32_t* dfs_rank);
/* number of geodesic chords per ray: count of j >= 0 with j*delta_theta < 2*pi (src/bvh.cpp:105) */
int rrt_chords_per_ray(rrt_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* RRT_B200_H */
