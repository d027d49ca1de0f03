// rrt_api.cu — the C-ABI of include/rrt_b200.h: context, uploads, the wavefront render loop, parity hooks.
// Host code here only orchestrates; every number is produced by the kernels in rrt_kernels.cuh.
// There is no CPU fallback: without a CUDA device rrt_create fails with RRT_E_CUDA.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <vector>
#include <chrono>

#include "rrt_bvh.h"
#include "rrt_kernels.cuh"
#include "rrt_bvh_device.cuh"

static thread_local std::string g_create_error;

struct DevBuf {
  void* p = nullptr; size_t bytes = 0;
  cudaError_t ensure(size_t need) {
    if (need <= bytes) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; bytes = 0;
    cudaError_t e = cudaMalloc(&p, need ? need : 16);
    if (e == cudaSuccess) bytes = need;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
  template <class T> T* as() const { return (T*)p; }
};

struct rrt_ctx {
  int device = 0, sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaStream_t aux[3] = {nullptr, nullptr, nullptr};   /* wavefront chunks of one round are dealt over stream + aux[0 .. n-2] (RRT_OPT_STREAMS) */
  cudaEvent_t ev_fork = nullptr, ev_join[3] = {nullptr, nullptr, nullptr};
  cudaStream_t shadow_stream[4] = {nullptr, nullptr, nullptr, nullptr};   /* RRT_OPT_OVERLAP: shadow rays of depth i beside extension rays of depth i + 1 */
  std::vector<cudaEvent_t> sync_events; size_t sync_used = 0;
  std::string err;
  // configuration
  bool have_scene = false, have_cam = false;
  DevScene ds; DevHole bh; DevCamera cam;
  rrt::FlatBvh bvh;
  int n_lights = 0, n_delta_lights = 0;
  std::vector<rrt_light> host_lights;
  bool has_microfacet = false;
  int W = 0, H = 0;
  int cell_x = 0, cell_y = 0, cell_w = 0, cell_h = 0;
  // progressive output: callback between adaptive batches; snap_* are valid while it runs
  rrt_progress_fn progress_fn = nullptr; void* progress_user = nullptr;
  bool in_progress_cb = false; PixelAcc snap_px; size_t snap_npix = 0;
  int tile = 32, rank = 0, world = 1;
  int64_t opt[16];
  double sync_wait_ms = 0;   /* host time spent blocked on the round-end read-backs of the current render (RRT_ENQUEUE_TIMING) */
  // device memory
  DevBuf d_env, d_env_pdf, d_env_cond, d_env_marg;
  bool have_env = false;
  DevBuf d_top, d_tripos, d_build, d_grid, d_nodes, d_wide, d_leaf, d_nrm, d_kind, d_index, d_pmat, d_sph, d_mats, d_lights;
  DevBuf d_work, d_px, d_act, d_blk, d_cnt, d_out_rgb, d_out_cnt, d_out_rgbn, d_hook_in, d_hook_out;
  std::vector<cudaEvent_t> events;
  size_t events_used = 0;
  uint32_t* h_pinned = nullptr;  // small pinned scratch for counts
  // multi-GPU context (rrt_create_multi): this object only owns the per-GPU contexts and their NCCL communicators
  std::vector<rrt_ctx*> subs;
  std::vector<void*> comms;      // ncclComm_t per sub-context
};

#define CK(call)                                                                                           \
  do {                                                                                                     \
    cudaError_t e_ = (call);                                                                               \
    if (e_ != cudaSuccess) {                                                                               \
      char b_[512]; snprintf(b_, sizeof b_, "%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      ctx->err = b_; return RRT_E_CUDA;                                                                    \
    }                                                                                                      \
  } while (0)
#define FAIL(code, msg) do { ctx->err = (msg); return (code); } while (0)
/* multi-GPU contexts: calls that concern one GPU (parity hooks, introspection) go to the first sub-context */
#define RRT_MULTI_FWD(ctx, CALL)                                         \
  if ((ctx) && !(ctx)->subs.empty()) {                                   \
    rrt_ctx* m_ = (ctx); rrt_ctx* ctx0 = m_->subs[0];                    \
    int rc_ = (CALL);                                                    \
    if (rc_ < 0) m_->err = ctx0->err;                                    \
    return rc_;                                                          \
  }
/* ... and configuration calls go to every sub-context */
#define RRT_MULTI_ALL(ctx, CALL)                                         \
  if ((ctx) && !(ctx)->subs.empty()) {                                   \
    rrt_ctx* m_ = (ctx);                                                 \
    for (rrt_ctx* sub : m_->subs) { int rc_ = (CALL); if (rc_ < 0) { m_->err = sub->err; return rc_; } } \
    return RRT_OK;                                                       \
  }
static int multi_upload_scene(rrt_ctx* m, const rrt_scene* s);
static int multi_render(rrt_ctx* m, const rrt_params* p, float* hdr_rgb, int32_t* sample_counts, void* d_rgbn_user, void* d_counts_user, rrt_stats* stats);
static void multi_destroy(rrt_ctx* m);

static int chords_for(double dtheta) {
  int j = 0;
  while (j * dtheta < 2 * M_PI && j < (1 << 20)) ++j;  /* reference src/bvh.cpp:105 */
  return j;
}

/* FP32 copies of the hole's constants for fast_step (rrt_trace.cuh), rounded once from the doubles */
static void hole_fp32_constants(DevHole& b) {
  const double h = b.h;
  b.f_cx = (float)b.cx; b.f_cy = (float)b.cy; b.f_cz = (float)b.cz; b.f_r2 = (float)b.r2; b.f_r15 = (float)(1.5 * b.r);
  b.f_h = (float)h; b.f_h2 = (float)(h / 2); b.f_h24 = (float)(h * h / 4); b.f_h26 = (float)(h * h / 6);
  b.f_ch = (float)b.cos_h; b.f_sh = (float)b.sin_h; b.f_k = (float)(2.0 * sin(h / 2) * sin(h / 2));
  b.f_cot = (float)(b.cos_h / b.sin_h); b.f_tan2 = (float)tan(h / 2);
}
extern "C" {

int rrt_abi_version(void) { return RRT_ABI_VERSION; }

int rrt_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

const char* rrt_last_error(const rrt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int rrt_create(rrt_ctx** out, int device) {
  if (!out) { g_create_error = "rrt_create: out is NULL"; return RRT_E_ARG; }
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    g_create_error = std::string("rrt_create: no CUDA device (") + (e != cudaSuccess ? cudaGetErrorString(e) : "count = 0") +
                     "); this library has no CPU fallback";
    return RRT_E_CUDA;
  }
  if (device < 0 || device >= n) { g_create_error = "rrt_create: device index out of range"; return RRT_E_ARG; }
  if ((e = cudaSetDevice(device)) != cudaSuccess) { g_create_error = cudaGetErrorString(e); return RRT_E_CUDA; }
  rrt_ctx* ctx = new rrt_ctx();
  ctx->device = device;
  memset(ctx->opt, 0, sizeof ctx->opt);
  ctx->opt[RRT_OPT_COUNTERS] = 1;
  ctx->opt[RRT_OPT_LEAF_SIZE] = 2;
  ctx->opt[RRT_OPT_CLEARANCE_GRID] = 0;   /* measured: no gain while chords of a warp are traversed in lockstep */
  ctx->opt[RRT_OPT_BVH_TOPOLOGY] = 1;
  ctx->opt[RRT_OPT_BVH_WIDTH] = 0;
  ctx->opt[RRT_OPT_SHADE_SPLIT] = 2;
  ctx->opt[RRT_OPT_STREAMS] = 2;
  ctx->opt[RRT_OPT_ESCAPE] = 0;
  ctx->opt[RRT_OPT_TOP_NODE] = 1;
  ctx->opt[RRT_OPT_OVERLAP] = 1;
  ctx->opt[RRT_OPT_SORT_SHADOW] = 1;
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess || (e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess ||
      (e = cudaMallocHost((void**)&ctx->h_pinned, 4096)) != cudaSuccess) {
    g_create_error = cudaGetErrorString(e); delete ctx; return RRT_E_CUDA;
  }
  ctx->sm_count = prop.multiProcessorCount;
  memset(&ctx->ds, 0, sizeof ctx->ds); memset(&ctx->cam, 0, sizeof ctx->cam);
  /* global_black_hole(nullptr, Vector3D(0, 1, 0), 0.1, 0.1)  (reference src/static_scene/blackhole.cpp:5) */
  ctx->bh.cx = 0; ctx->bh.cy = 1; ctx->bh.cz = 0; ctx->bh.r = 0.1; ctx->bh.r2 = 0.1 * 0.1; ctx->bh.h = 0.1;
  ctx->bh.cos_h = cos(0.1); ctx->bh.sin_h = sin(0.1); ctx->bh.mode = RRT_MODE_LITERAL; ctx->bh.nchords = chords_for(0.1);
  hole_fp32_constants(ctx->bh);
  *out = ctx;
  return RRT_OK;
}

void rrt_destroy(rrt_ctx* ctx) {
  if (!ctx) return;
  if (!ctx->subs.empty()) { multi_destroy(ctx); return; }
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  DevBuf* bufs[] = {&ctx->d_top, &ctx->d_env, &ctx->d_env_pdf, &ctx->d_env_cond, &ctx->d_env_marg, &ctx->d_grid, &ctx->d_nodes, &ctx->d_tripos, &ctx->d_build, &ctx->d_wide, &ctx->d_leaf, &ctx->d_nrm, &ctx->d_kind, &ctx->d_index, &ctx->d_pmat, &ctx->d_sph, &ctx->d_mats, &ctx->d_lights,
                    &ctx->d_work, &ctx->d_px, &ctx->d_act, &ctx->d_blk, &ctx->d_cnt, &ctx->d_out_rgb, &ctx->d_out_cnt, &ctx->d_out_rgbn,
                    &ctx->d_hook_in, &ctx->d_hook_out};
  for (DevBuf* b : bufs) b->release();
  for (cudaEvent_t ev : ctx->events) cudaEventDestroy(ev);
  if (ctx->h_pinned) cudaFreeHost(ctx->h_pinned);
  for (int k = 0; k < 3; ++k) { if (ctx->aux[k]) cudaStreamDestroy(ctx->aux[k]); if (ctx->ev_join[k]) cudaEventDestroy(ctx->ev_join[k]); }
  if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
  for (int k = 0; k < 4; ++k) if (ctx->shadow_stream[k]) cudaStreamDestroy(ctx->shadow_stream[k]);
  for (cudaEvent_t ev : ctx->sync_events) cudaEventDestroy(ev);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

int rrt_set_option(rrt_ctx* ctx, int option, int64_t value) {
  if (!ctx) return RRT_E_ARG;
  RRT_MULTI_ALL(ctx, rrt_set_option(sub, option, value));
  if (option < 0 || option > RRT_OPT_SORT_SHADOW) FAIL(RRT_E_ARG, "rrt_set_option: unknown option");
#if !RRT_AB
  if ((option == RRT_OPT_TRACE_KERNEL && value != 0) || (option == RRT_OPT_BVH_WIDTH && value == 4) || (option == RRT_OPT_ESCAPE && value != 0))
    FAIL(RRT_E_ARG, "rrt_set_option: this variant is compiled only into A/B builds (make EXTRA=-DRRT_AB=1)");
#endif
  ctx->opt[option] = value;
  return RRT_OK;
}

static int upload(rrt_ctx* ctx, DevBuf& b, const void* src, size_t bytes) {
  CK(b.ensure(bytes));
  if (bytes) CK(cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
  return RRT_OK;
}

}  /* extern "C" */
/* interior levels of the flattened tree = the most entries a traversal can have on its stack, plus one */
static int flat_depth(const rrt::FlatBvh& b) {
  int depth = 0;
  std::vector<std::pair<int, int>> st;
  if (b.root >= 0) st.push_back({b.root, 1});
  while (!st.empty()) {
    auto [nd, dp] = st.back(); st.pop_back();
    depth = std::max(depth, dp);
    int32_t c[2]; memcpy(c, &b.nodes[16 * (size_t)nd + 12], 8);
    for (int k = 0; k < 2; ++k) if (c[k] >= 0) st.push_back({c[k], dp + 1});
  }
  return depth;
}


/* ---------------------------------------------------------------- device BVH build (RRT_OPT_BVH_BUILDER = 1, rrt_bvh_device.cuh)
 * Needs d_kind / d_index / d_sph uploaded.  Fills d_nodes and d_leaf on the device and the host-side summary in ctx->bvh
 * (root box, counts, reference-tree statistics, depth-first ranks for rrt_bvh_info).  *built = false: the tree came out
 * deeper than the traversal stack (or the scene is tiny) and the caller falls back to the host builder. */
static int build_bvh_on_device(rrt_ctx* ctx, const rrt_scene* s, bool* built) {
  using namespace rrtdev;
  *built = false;
  const uint32_t n = s->n_prims;
  if (n < 64) return RRT_OK;
  cudaStream_t st = ctx->stream;
  int rc;
  if ((rc = upload(ctx, ctx->d_tripos, s->tri_pos, 72 * (size_t)s->n_tris))) return rc;
  /* scratch layout */
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
  const size_t o_plo = take(24 * (size_t)n), o_phi = take(24 * (size_t)n), o_cen = take(24 * (size_t)n);
  const size_t o_ord[2] = {take(4 * (size_t)n), take(4 * (size_t)n)}, o_ss[2] = {take(4 * (size_t)n), take(4 * (size_t)n)}, o_se[2] = {take(4 * (size_t)n), take(4 * (size_t)n)};
  const size_t o_isl = take(4 * ((size_t)n + 1)), o_pre = take(4 * ((size_t)n + 1)), o_seg = take(48 * (size_t)n), o_rank = take(4 * (size_t)n);
  const size_t o_key[2] = {take(8 * (size_t)n), take(8 * (size_t)n)}, o_val[2] = {take(4 * (size_t)n), take(4 * (size_t)n)};
  const size_t o_child = take(8 * (size_t)n), o_pint = take(4 * (size_t)n), o_pleaf = take(4 * (size_t)n), o_vis = take(4 * (size_t)n), o_dep = take(4 * (size_t)n);
  const size_t o_lbox = take(24 * (size_t)n), o_nbox = take(24 * (size_t)n), o_flags = take(256);
  size_t scan_bytes = 0, sort_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n + 1, st);
  cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n, 0, 63, st);
  const size_t o_tmp = take(std::max(scan_bytes, sort_bytes) + 256);
  CK(ctx->d_build.ensure(off));
  char* base = (char*)ctx->d_build.p;
  double* plo = (double*)(base + o_plo); double* phi = (double*)(base + o_phi); double* cen = (double*)(base + o_cen);
  uint32_t* ord[2] = {(uint32_t*)(base + o_ord[0]), (uint32_t*)(base + o_ord[1])};
  uint32_t* ss[2] = {(uint32_t*)(base + o_ss[0]), (uint32_t*)(base + o_ss[1])};
  uint32_t* se[2] = {(uint32_t*)(base + o_se[0]), (uint32_t*)(base + o_se[1])};
  uint32_t* isl = (uint32_t*)(base + o_isl); uint32_t* pre = (uint32_t*)(base + o_pre);
  unsigned long long* segbox = (unsigned long long*)(base + o_seg);
  uint32_t* rank = (uint32_t*)(base + o_rank);
  unsigned long long* key[2] = {(unsigned long long*)(base + o_key[0]), (unsigned long long*)(base + o_key[1])};
  uint32_t* val[2] = {(uint32_t*)(base + o_val[0]), (uint32_t*)(base + o_val[1])};
  int* child = (int*)(base + o_child); int* pint = (int*)(base + o_pint); int* pleaf = (int*)(base + o_pleaf); int* vis = (int*)(base + o_vis); int* dep = (int*)(base + o_dep);
  float* lbox = (float*)(base + o_lbox); float* nbox = (float*)(base + o_nbox);
  int* flags = (int*)(base + o_flags);   /* [0] any segment left to split, [1] reference nodes, [2] reference leaves */
  void* tmp = base + o_tmp; size_t tmp_bytes = std::max(scan_bytes, sort_bytes) + 256;
  const int B = 256; const int G = (int)((n + B - 1) / B), G1 = (int)((n + 1 + B - 1) / B);

  k_prim_bounds<<<G, B, 0, st>>>(n, ctx->d_kind.as<uint8_t>(), ctx->d_index.as<uint32_t>(), ctx->d_tripos.as<double>(), ctx->d_sph.as<double>(), plo, phi, cen);
  /* 1. the reference's depth-first order */
  k_ref_init<<<G, B, 0, st>>>(n, ord[0], ss[0], se[0]);
  CK(cudaMemsetAsync(flags, 0, 256, st));
  int cur = 0, levels = 0;
  Box6 sb;
  int* h_flags = (int*)ctx->h_pinned;
  /* the root box = the scene box (doubles, exact): first half of level 0, read back once */
  k_ref_reset<<<G, B, 0, st>>>(n, ss[cur], se[cur], segbox);
  k_ref_accum<<<G, B, 0, st>>>(n, ord[cur], ss[cur], se[cur], plo, phi, segbox);
  {
    unsigned long long h[6];
    for (int k = 0; k < 6; ++k) CK(cudaMemcpyAsync(&h[k], segbox + (size_t)k * n, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    for (int k = 0; k < 3; ++k) {
      unsigned long long e0 = h[k], e1 = h[3 + k];
      unsigned long long b0 = (e0 >> 63) ? (e0 & 0x7fffffffffffffffull) : ~e0, b1 = (e1 >> 63) ? (e1 & 0x7fffffffffffffffull) : ~e1;
      memcpy(&sb.lo[k], &b0, 8); memcpy(&sb.hi[k], &b1, 8);
    }
  }
  /* levels run in groups of 8 without a host round trip (C5's reference tree is 194 levels deep); a level with nothing left to
   * split is a no-op that copies the order through, and each level leaves its own "new segments to split" flag behind */
  const int GROUP = 8;
  bool first = true;
  for (bool done = false; !done;) {
    if (levels > 1 << 16) FAIL(RRT_E_ARG, "rrt_upload_scene: device BVH build did not terminate");
    CK(cudaMemsetAsync(flags + 8, 0, 4 * GROUP, st));
    for (int g = 0; g < GROUP; ++g) {
      if (!first) {
        k_ref_reset<<<G, B, 0, st>>>(n, ss[cur], se[cur], segbox);
        k_ref_accum<<<G, B, 0, st>>>(n, ord[cur], ss[cur], se[cur], plo, phi, segbox);
      }
      first = false;
      k_ref_flag<<<G1, B, 0, st>>>(n, ord[cur], ss[cur], se[cur], segbox, cen, isl);
      size_t tb = tmp_bytes;
      cub::DeviceScan::ExclusiveSum(tmp, tb, isl, pre, (int)n + 1, st);
      k_ref_scatter<<<G, B, 0, st>>>(n, ord[cur], ss[cur], se[cur], isl, pre, ord[cur ^ 1], ss[cur ^ 1], se[cur ^ 1], flags + 8 + g, flags + 1);
      cur ^= 1;
    }
    CK(cudaMemcpyAsync(h_flags, flags, 4 * (8 + GROUP), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    for (int g = 0; g < GROUP; ++g) {
      ++levels;
      if (!h_flags[8 + g]) { done = true; break; }
    }
  }
  k_ref_rank<<<G, B, 0, st>>>(n, ord[cur], rank);
  /* 2. LBVH */
  k_morton<<<G, B, 0, st>>>(n, sb, plo, phi, cen, key[0], val[0]);
  {
    size_t tb = tmp_bytes;
    cub::DeviceRadixSort::SortPairs(tmp, tb, key[0], key[1], val[0], val[1], (int)n, 0, 63, st);
  }
  CK(cudaMemsetAsync(vis, 0, 4 * (size_t)n, st));
  k_lbvh_topology<<<G, B, 0, st>>>((int)n, key[1], child, pint, pleaf);
  k_lbvh_refit<<<G, B, 0, st>>>((int)n, val[1], plo, phi, child, pint, pleaf, lbox, nbox, dep, vis);
  CK(cudaMemcpyAsync(h_flags + 4, dep, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  CK(cudaGetLastError());
  const int depth = h_flags[4];
  if (depth >= RRT_STACK) return RRT_OK;   /* too deep for the shared-memory stack: host builder */
  CK(ctx->d_nodes.ensure(64 * (size_t)(n - 1)));
  CK(ctx->d_leaf.ensure(80 * (size_t)n));
  k_lbvh_emit<<<G, B, 0, st>>>((int)n, child, lbox, nbox, ctx->d_nodes.as<float4>());
  k_leaf_records<<<G, B, 0, st>>>(n, val[1], ctx->d_kind.as<uint8_t>(), ctx->d_index.as<uint32_t>(), ctx->d_tripos.as<double>(), ctx->d_sph.as<double>(), rank,
                                  ctx->d_leaf.as<double>());
  rrt::FlatBvh& b = ctx->bvh;
  b = rrt::FlatBvh();
  b.dfs_rank.resize(n);
  CK(cudaMemcpyAsync(b.dfs_rank.data(), rank, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  CK(cudaGetLastError());
  b.prim_mat.resize(n);
  for (uint32_t p = 0; p < n; ++p) b.prim_mat[p] = s->prim_kind[p] == 0 ? s->tri_mat[s->prim_index[p]] : s->sph_mat[s->prim_index[p]];
  b.root = 0; b.n_nodes = (int)n - 1; b.width = 2; b.device_depth = depth; b.device_built = true;
  for (int k = 0; k < 3; ++k) { b.root_lo[k] = sb.lo[k]; b.root_hi[k] = sb.hi[k]; }
  b.ref_nodes = 1 + h_flags[1]; b.ref_leaves = n <= 4 ? 1 : h_flags[2]; b.ref_depth = n <= 4 ? 0 : levels;
  *built = true;
  return RRT_OK;
}

/* The top of a binary tree as one 8-wide record (rrt_trace.cuh traverse()).  Starting from the root's two children, the interior child with the
 * largest surface area that is not flat (a grouping of walls, not the two coplanar triangles of one wall) is replaced by its two children until 8
 * slots are used or nothing is left to open.  `fetch(i, rec)` delivers the
 * 64-byte record of interior node i.  out: 6 arrays of 8 floats (lo.x lo.y lo.z hi.x hi.y hi.z) + 8 child codes + 8 pad; returns the child count. */
template <class Fetch>
static int build_top_record(Fetch fetch, int32_t root_code, float out[64], int max_children = 8) {
  struct Child { int32_t code; float lo[3], hi[3]; };
  auto area = [](const Child& c) {
    const double dx = (double)c.hi[0] - c.lo[0], dy = (double)c.hi[1] - c.lo[1], dz = (double)c.hi[2] - c.lo[2];
    return (dx < 0 || dy < 0 || dz < 0) ? 0.0 : 2.0 * (dx * dy + dy * dz + dz * dx);
  };
  auto kids_of = [&](int32_t node, Child* L, Child* R) {
    float q[16]; fetch(node, q);
    L->lo[0] = q[0]; L->lo[1] = q[1]; L->lo[2] = q[2]; L->hi[0] = q[3]; L->hi[1] = q[4]; L->hi[2] = q[5];
    R->lo[0] = q[6]; R->lo[1] = q[7]; R->lo[2] = q[8]; R->hi[0] = q[9]; R->hi[1] = q[10]; R->hi[2] = q[11];
    memcpy(&L->code, &q[12], 4); memcpy(&R->code, &q[13], 4);
  };
  std::vector<Child> ch(2);
  kids_of(root_code, &ch[0], &ch[1]);
  auto flat = [](const Child& c) {   /* a wall: opening it yields two triangles with the same box */
    const double e[3] = {(double)c.hi[0] - c.lo[0], (double)c.hi[1] - c.lo[1], (double)c.hi[2] - c.lo[2]};
    const double mx = std::max(e[0], std::max(e[1], e[2])), mn = std::min(e[0], std::min(e[1], e[2]));
    return mn <= 1e-3 * mx;
  };
  while ((int)ch.size() < max_children) {
    int pick = -1; double best = 0.0;
    for (size_t k = 0; k < ch.size(); ++k) {
      if (ch[k].code < 0 || flat(ch[k])) continue;
      const double a = area(ch[k]);
      if (a > best) { best = a; pick = (int)k; }
    }
    if (pick < 0) break;
    Child l, r; kids_of(ch[pick].code, &l, &r);
    ch[pick] = l; ch.push_back(r);
  }
  /* a child whose two children are leaves with adjacent records (the two triangles of one wall) becomes that leaf itself: one binary visit less
   * for every chord that comes near the wall */
  for (Child& c : ch) {
    if (c.code < 0) continue;
    Child l, r; kids_of(c.code, &l, &r);
    if (l.code >= 0 || r.code >= 0 || l.code == rrt::kEmptyChild || r.code == rrt::kEmptyChild) continue;
    const uint32_t bl = (uint32_t)~l.code, br = (uint32_t)~r.code;
    uint32_t f0 = bl >> 3, n0 = bl & 7u, f1 = br >> 3, n1 = br & 7u;
    if (f1 + n1 == f0) { std::swap(f0, f1); std::swap(n0, n1); }
    if (f0 + n0 == f1 && n0 + n1 <= 7u) c.code = ~(int32_t)((f0 << 3) | (n0 + n1));
  }
  for (int k = 0; k < 8; ++k) {
    int32_t code = rrt::kEmptyChild;
    if (k < (int)ch.size()) {
      for (int a = 0; a < 3; ++a) { out[a * 8 + k] = ch[k].lo[a]; out[(3 + a) * 8 + k] = ch[k].hi[a]; }
      code = ch[k].code;
    } else {
      for (int a = 0; a < 3; ++a) { out[a * 8 + k] = INFINITY; out[(3 + a) * 8 + k] = -INFINITY; }
    }
    memcpy(&out[48 + k], &code, 4); out[56 + k] = 0.f;
  }
  return (int)ch.size();
}

extern "C" {
int rrt_upload_scene(rrt_ctx* ctx, const rrt_scene* s) {
  if (!ctx || !s) return RRT_E_ARG;
  if (!ctx->subs.empty()) return multi_upload_scene(ctx, s);
  CK(cudaSetDevice(ctx->device));
  if (s->n_prims && (!s->prim_kind || !s->prim_index)) FAIL(RRT_E_ARG, "rrt_upload_scene: primitive arrays missing");
  if ((s->n_tris && (!s->tri_pos || !s->tri_nrm || !s->tri_mat)) || (s->n_spheres && (!s->sph || !s->sph_mat)))
    FAIL(RRT_E_ARG, "rrt_upload_scene: geometry arrays missing");
  if (s->n_prims >= (1u << 28)) FAIL(RRT_E_ARG, "rrt_upload_scene: more than 2^28 primitives");
  for (uint32_t p = 0; p < s->n_prims; ++p) {
    uint32_t k = s->prim_kind[p], idx = s->prim_index[p];
    if (k > 1 || idx >= (k == 0 ? s->n_tris : s->n_spheres)) FAIL(RRT_E_ARG, "rrt_upload_scene: primitive table out of range");
    int32_t m = k == 0 ? s->tri_mat[idx] : s->sph_mat[idx];
    if (m < 0 || (uint32_t)m >= s->n_mats) FAIL(RRT_E_ARG, "rrt_upload_scene: material index out of range");
  }
  int width = (int)ctx->opt[RRT_OPT_BVH_WIDTH];
  /* 0 = choose: a scene small enough that a couple of 8-wide nodes cover every leaf is walked ~7 % faster that way
   * (C2: 37.3 vs 40.1 ms per frame); larger trees are faster over binary nodes with ordered descent (bunny, 28 k
   * triangles: 110.8 vs 116.2 ms; C3 trace time 16.0 vs 18.1 ms).  Measured on B200, DESIGN.md section 4.2. */
  if (width == 0) width = s->n_prims <= 128 ? 8 : 2;
  if (ctx->opt[RRT_OPT_TRACE_KERNEL] == 1 || ctx->opt[RRT_OPT_BVH_TOPOLOGY] == 0) width = 2;   /* A/B variants keep the binary nodes */
  int rc;
  CK(cudaStreamSynchronize(ctx->stream));
  if ((rc = upload(ctx, ctx->d_kind, s->prim_kind, s->n_prims))) return rc;
  if ((rc = upload(ctx, ctx->d_index, s->prim_index, 4 * (size_t)s->n_prims))) return rc;
  if ((rc = upload(ctx, ctx->d_sph, s->sph, 32 * (size_t)s->n_spheres))) return rc;
  bool on_device = false;
  /* 0 = auto: on the device for scenes of >= 4096 primitives when no host-tree option (shape, leaf size, node width, A/B kernel)
   * is set: measured on B200, the LBVH is walked 0.5-3 % slower than the host's binned-SAH tree and is ready ~100x sooner
   * (C5, 1 M triangles: e2e 32.9 -> 339.7 Mpaths/s at 4 spp); 1 = device whenever possible; 2 = host */
  const int64_t builder = ctx->opt[RRT_OPT_BVH_BUILDER];
  const bool host_opts = ctx->opt[RRT_OPT_BVH_TOPOLOGY] != 1 || ctx->opt[RRT_OPT_LEAF_SIZE] != 2 || ctx->opt[RRT_OPT_BVH_WIDTH] != 0;
  const bool want_device = ctx->opt[RRT_OPT_TRACE_KERNEL] != 1 && (builder == 1 || (builder == 0 && !host_opts && s->n_prims >= 4096));
  if (want_device && (rc = build_bvh_on_device(ctx, s, &on_device))) return rc;
  if (!on_device) {
    rrt::build_bvh(s, (int)ctx->opt[RRT_OPT_BVH_TOPOLOGY], (int)ctx->opt[RRT_OPT_LEAF_SIZE], &ctx->bvh, width);
    if (ctx->bvh.width > 2 && ctx->bvh.width * ctx->bvh.wide_depth + 1 > RRT_STACK + RRT_SPILL) {
      /* cannot happen below 2^28 primitives with the depth budget of the builder; keep the binary nodes if it ever does */
      rrt::build_bvh(s, (int)ctx->opt[RRT_OPT_BVH_TOPOLOGY], (int)ctx->opt[RRT_OPT_LEAF_SIZE], &ctx->bvh, 2);
    }
  }
  rrt::build_clearance_grid(s, (int)ctx->opt[RRT_OPT_CLEARANCE_GRID], &ctx->bvh);
  const rrt::FlatBvh& b = ctx->bvh;
  if (!on_device) {
    const int tree_depth = flat_depth(b);   /* bounds the traversal stack */
    if (tree_depth >= RRT_STACK) FAIL(RRT_E_ARG, "rrt_upload_scene: BVH deeper than the traversal stack; set RRT_OPT_BVH_TOPOLOGY = 1");
    if ((rc = upload(ctx, ctx->d_nodes, b.nodes.data(), b.nodes.size() * 4))) return rc;
    if ((rc = upload(ctx, ctx->d_wide, b.wide.data(), b.wide.size() * 4))) return rc;
    if ((rc = upload(ctx, ctx->d_leaf, b.leaf.data(), b.leaf.size() * 8))) return rc;
  }
  if ((rc = upload(ctx, ctx->d_grid, b.clearance.data(), b.clearance.size() * 4))) return rc;
  if ((rc = upload(ctx, ctx->d_nrm, s->tri_nrm, 72 * (size_t)s->n_tris))) return rc;
  if ((rc = upload(ctx, ctx->d_pmat, b.prim_mat.data(), 4 * (size_t)s->n_prims))) return rc;
  if ((rc = upload(ctx, ctx->d_mats, s->mats, sizeof(rrt_material) * (size_t)s->n_mats))) return rc;
  if ((rc = upload(ctx, ctx->d_lights, s->lights, sizeof(rrt_light) * (size_t)s->n_lights))) return rc;
  CK(cudaStreamSynchronize(ctx->stream));
  DevScene& d = ctx->ds;
  d.nodes = ctx->d_nodes.as<float4>(); d.leaf = ctx->d_leaf.as<double2>(); d.tri_nrm = ctx->d_nrm.as<double>();
  d.prim_kind = ctx->d_kind.as<uint8_t>(); d.prim_index = ctx->d_index.as<uint32_t>(); d.prim_mat = ctx->d_pmat.as<int32_t>();
  d.sph = ctx->d_sph.as<double>(); d.mats = ctx->d_mats.as<rrt_material>(); d.lights = ctx->d_lights.as<rrt_light>();
  d.clearance = b.clearance.empty() ? nullptr : ctx->d_grid.as<float>();
  d.gnx = b.grid_n[0]; d.gny = b.grid_n[1]; d.gnz = b.grid_n[2];
  for (int k = 0; k < 3; ++k) { d.glo[k] = b.grid_lo[k]; d.ginv[k] = b.grid_inv_h[k]; }
  d.n_prims = (int)s->n_prims; d.n_mats = (int)s->n_mats; d.n_lights = (int)s->n_lights;
  d.width = b.width; d.wide = b.width > 2 ? ctx->d_wide.as<float>() : nullptr;
  d.root = b.root; d.wroot = b.wide_root;
  d.top = nullptr;
  {
    const int tree_depth = on_device ? b.device_depth : flat_depth(b);
    /* stack: the top visit leaves at most (children - 1) entries below whatever the binary descent pushes (<= tree_depth - 1) */
    const int max_children = std::min(8, RRT_STACK - tree_depth + 1);
    if (ctx->opt[RRT_OPT_TOP_NODE] != 0 && b.width <= 2 && b.root >= 0 && b.n_nodes >= 8 && max_children >= 3 && ctx->opt[RRT_OPT_TRACE_KERNEL] != 1) {
      float rec[64];
      cudaError_t fe = cudaSuccess;
      auto fetch = [&](int32_t node, float* q) {
        if (on_device) { cudaError_t e = cudaMemcpy(q, ctx->d_nodes.as<float>() + 16 * (size_t)node, 64, cudaMemcpyDeviceToHost); if (e != cudaSuccess) fe = e; }
        else memcpy(q, &b.nodes[16 * (size_t)node], 64);
      };
      const int nch = build_top_record(fetch, b.root, rec, max_children);
      CK(fe);
      if (nch > 2) {
        if ((rc = upload(ctx, ctx->d_top, rec, sizeof rec))) return rc;
        CK(cudaStreamSynchronize(ctx->stream));
        d.top = ctx->d_top.as<float>();
      }
    }
  }
  for (int k = 0; k < 3; ++k) { d.root_lo[k] = rrt::round_down_f(b.root_lo[k]); d.root_hi[k] = rrt::round_up_f(b.root_hi[k]); }
  {
    double r2 = 0.0;
    for (int k = 0; k < 3; ++k) {
      d.ctr[k] = 0.5 * ((double)d.root_lo[k] + (double)d.root_hi[k]);
      double e = 0.5 * ((double)d.root_hi[k] - (double)d.root_lo[k]);
      r2 += e * e;
    }
    d.rad = std::sqrt(r2) * 1.0001 + 1e-9;   /* only precision depends on it (rrt_trace.cuh boxctx_make) */
    if (!(d.rad < 1e300)) { d.rad = 0.0; d.ctr[0] = d.ctr[1] = d.ctr[2] = 0.0; }
  }
  ctx->n_lights = (int)s->n_lights; ctx->n_delta_lights = 0;
  ctx->host_lights.assign(s->lights, s->lights + s->n_lights);
  ctx->has_microfacet = false;
  for (uint32_t i = 0; i < s->n_mats; ++i) if (s->mats[i].type == RRT_MAT_MICROFACET) ctx->has_microfacet = true;
  for (uint32_t i = 0; i < s->n_lights; ++i) if (s->lights[i].is_delta) ctx->n_delta_lights++;
  ctx->have_scene = true;
  return RRT_OK;
}

int rrt_set_camera(rrt_ctx* ctx, const rrt_camera* cam) {
  if (!ctx || !cam) return RRT_E_ARG;
  RRT_MULTI_ALL(ctx, rrt_set_camera(sub, cam));
  memcpy(ctx->cam.pos, cam->pos, sizeof cam->pos); memcpy(ctx->cam.c2w, cam->c2w, sizeof cam->c2w);
  ctx->cam.hFov = cam->hFov; ctx->cam.vFov = cam->vFov; ctx->cam.nClip = cam->nClip; ctx->cam.fClip = cam->fClip;
  ctx->cam.lensRadius = cam->lensRadius; ctx->cam.focalDistance = cam->focalDistance;
  ctx->have_cam = true;
  return RRT_OK;
}

int rrt_set_blackhole(rrt_ctx* ctx, const double c[3], double r, double dtheta, int mode) {
  if (!ctx || !c) return RRT_E_ARG;
  RRT_MULTI_ALL(ctx, rrt_set_blackhole(sub, c, r, dtheta, mode));
  if (mode != RRT_MODE_FLAT && mode != RRT_MODE_LITERAL) FAIL(RRT_E_ARG, "rrt_set_blackhole: unknown mode");
  /* a NaN or non-positive step makes `j * dtheta < 2 pi` true for ever (reference src/bvh.cpp:105: the reference would hang); a tiny one
   * makes every ray walk millions of chords.  2 pi / 65536 is 650 times finer than the reference's default of 0.1 */
  if (!(dtheta >= 9.5873799e-5) || !std::isfinite(dtheta)) FAIL(RRT_E_ARG, "rrt_set_blackhole: delta_theta must be a finite number >= 2 pi / 65536");
  if (!(r >= 0.0) || !std::isfinite(r) || !std::isfinite(c[0]) || !std::isfinite(c[1]) || !std::isfinite(c[2]))
    FAIL(RRT_E_ARG, "rrt_set_blackhole: centre and radius must be finite, radius >= 0");
  ctx->bh.cx = c[0]; ctx->bh.cy = c[1]; ctx->bh.cz = c[2]; ctx->bh.r = r; ctx->bh.r2 = r * r; ctx->bh.h = dtheta;
  ctx->bh.cos_h = cos(dtheta); ctx->bh.sin_h = sin(dtheta);  /* host libm, as the reference evaluates them (src/static_scene/blackhole.cpp:34-35) */
  ctx->bh.mode = mode; ctx->bh.nchords = chords_for(dtheta);
  hole_fp32_constants(ctx->bh);
  return RRT_OK;
}

int rrt_set_frame(rrt_ctx* ctx, int w, int h) {
  if (!ctx) return RRT_E_ARG;
  if (!ctx->subs.empty()) { ctx->W = w; ctx->H = h; }
  RRT_MULTI_ALL(ctx, rrt_set_frame(sub, w, h));
  if (w <= 0 || h <= 0 || (int64_t)w * h >= (1ll << 31)) FAIL(RRT_E_ARG, "rrt_set_frame: bad size");
  ctx->W = w; ctx->H = h; ctx->cell_w = 0;
  return RRT_OK;
}

int rrt_set_cell(rrt_ctx* ctx, int x, int y, int w, int h) {
  if (!ctx) return RRT_E_ARG;
  RRT_MULTI_ALL(ctx, rrt_set_cell(sub, x, y, w, h));
  if (w > 0 && h <= 0) FAIL(RRT_E_ARG, "rrt_set_cell: a cell needs a positive height (w <= 0 restores the full frame)");
  ctx->cell_x = x; ctx->cell_y = y; ctx->cell_w = w; ctx->cell_h = h;
  return RRT_OK;
}

int rrt_set_progress(rrt_ctx* ctx, rrt_progress_fn fn, void* user) {
  if (!ctx) return RRT_E_ARG;
  if (!ctx->subs.empty()) { if (!fn) return RRT_OK; FAIL(RRT_E_ARG, "rrt_set_progress: not available on a multi-GPU context (the GPUs render their tiles independently)"); }
  if (ctx->in_progress_cb) FAIL(RRT_E_STATE, "rrt_set_progress: called from inside the callback");
  ctx->progress_fn = fn; ctx->progress_user = user;
  return RRT_OK;
}

int rrt_snapshot(rrt_ctx* ctx, float* hdr_rgb, int32_t* sample_counts) {
  if (!ctx) return RRT_E_ARG;
  if (!ctx->subs.empty()) FAIL(RRT_E_STATE, "rrt_snapshot: not available on a multi-GPU context");
  if (!ctx->in_progress_cb) FAIL(RRT_E_STATE, "rrt_snapshot: only valid inside an rrt_progress_fn call");
  const size_t npix = ctx->snap_npix;
  CK(cudaSetDevice(ctx->device));
  CK(ctx->d_out_rgb.ensure(12 * npix)); CK(ctx->d_out_cnt.ensure(4 * npix));
  k_finalize<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(ctx->snap_px, npix, ctx->d_out_rgb.as<float>(), ctx->d_out_cnt.as<int>(), nullptr, nullptr);
  if (hdr_rgb) CK(cudaMemcpyAsync(hdr_rgb, ctx->d_out_rgb.p, 12 * npix, cudaMemcpyDeviceToHost, ctx->stream));
  if (sample_counts) CK(cudaMemcpyAsync(sample_counts, ctx->d_out_cnt.p, 4 * npix, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  return RRT_OK;
}

int rrt_set_tile_shard(rrt_ctx* ctx, int tile, int rank, int world) {
  if (!ctx) return RRT_E_ARG;
  if (!ctx->subs.empty()) {   /* the context deals its own GPUs; only the tile size can be chosen */
    if (tile <= 0 || rank != 0 || world != 1) FAIL(RRT_E_ARG, "rrt_set_tile_shard: a multi-GPU context shards over its own GPUs; pass (tile, 0, 1)");
    for (size_t k = 0; k < ctx->subs.size(); ++k) { ctx->subs[k]->tile = tile; }
    return RRT_OK;
  }
  if (tile <= 0 || world <= 0 || rank < 0 || rank >= world) FAIL(RRT_E_ARG, "rrt_set_tile_shard: bad arguments");
  ctx->tile = tile; ctx->rank = rank; ctx->world = world;
  return RRT_OK;
}

int rrt_bvh_build_host(const rrt_scene* scene, int topology, int32_t info[8], uint32_t* dfs_rank) {
  if (!scene || !info) return RRT_E_ARG;
  rrt::FlatBvh b;
  rrt::build_bvh(scene, topology, 2, &b);
  info[0] = b.ref_nodes; info[1] = b.ref_leaves; info[2] = b.ref_depth; info[3] = b.n_nodes; info[4] = (int32_t)(b.leaf.size() / 10); info[5] = b.root;
  info[6] = flat_depth(b); info[7] = RRT_STACK;
  if (dfs_rank && !b.dfs_rank.empty()) memcpy(dfs_rank, b.dfs_rank.data(), 4 * b.dfs_rank.size());
  return RRT_OK;
}

int rrt_chords_per_ray(rrt_ctx* ctx) { if (ctx && !ctx->subs.empty()) return rrt_chords_per_ray(ctx->subs[0]); return ctx ? ctx->bh.nchords : RRT_E_ARG; }

int rrt_bvh_info(rrt_ctx* ctx, int32_t* n_nodes, int32_t* n_leaves, int32_t* depth, uint32_t* dfs_rank) {
  if (!ctx) return RRT_E_ARG;
  RRT_MULTI_FWD(ctx, rrt_bvh_info(ctx0, n_nodes, n_leaves, depth, dfs_rank));
  if (!ctx->have_scene) FAIL(RRT_E_STATE, "rrt_bvh_info: no scene");
  if (n_nodes) *n_nodes = ctx->bvh.ref_nodes;
  if (n_leaves) *n_leaves = ctx->bvh.ref_leaves;
  if (depth) *depth = ctx->bvh.ref_depth;
  if (dfs_rank) memcpy(dfs_rank, ctx->bvh.dfs_rank.data(), 4 * ctx->bvh.dfs_rank.size());
  return RRT_OK;
}

}  // extern "C"

/* ---------------------------------------------------------------- render */
static cudaEvent_t next_sync_event(rrt_ctx* ctx) {   /* ordering only, no timing */
  if (ctx->sync_used == ctx->sync_events.size()) { cudaEvent_t ev; cudaEventCreateWithFlags(&ev, cudaEventDisableTiming); ctx->sync_events.push_back(ev); }
  return ctx->sync_events[ctx->sync_used++];
}
static cudaEvent_t next_event(rrt_ctx* ctx) {
  if (ctx->events_used == ctx->events.size()) {
    cudaEvent_t ev; cudaEventCreate(&ev); ctx->events.push_back(ev);
  }
  return ctx->events[ctx->events_used++];
}

template <class Keep>
static int compact(rrt_ctx* ctx, Keep keep, const uint32_t* in, uint32_t n, uint32_t* out, uint32_t* n_out) {
  uint32_t nblocks = (n + RRT_CBLOCK - 1) / RRT_CBLOCK;
  if (n == 0) { *n_out = 0; return RRT_OK; }
  CK(ctx->d_blk.ensure(4 * (size_t)(nblocks + 1)));
  uint32_t* blk = ctx->d_blk.as<uint32_t>();
  k_compact_count<<<nblocks, RRT_CBLOCK, 0, ctx->stream>>>(keep, in, n, blk);
  k_compact_scan<<<1, 1024, 0, ctx->stream>>>(blk, nblocks, blk + nblocks);
  k_compact_scatter<<<nblocks, RRT_CBLOCK, 0, ctx->stream>>>(keep, in, n, blk, out);
  CK(cudaMemcpyAsync(ctx->h_pinned, blk + nblocks, 4, cudaMemcpyDeviceToHost, ctx->stream));
  const auto t_wait = std::chrono::steady_clock::now();
  CK(cudaStreamSynchronize(ctx->stream));   /* the one host round trip of an adaptive round: the next round's launch sizes depend on it */
  ctx->sync_wait_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_wait).count();
  *n_out = ctx->h_pinned[0];
  return RRT_OK;
}

static size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

/* ---- trace-kernel selection.  Template arguments of the kernels in rrt_wavetrace.cuh:
 *   V    = node form / scheduler: 0 lockstep-or-decoupled over binary nodes, 8 over 8-wide nodes; A/B builds (-DRRT_AB=1) add 1 = vote-scheduled
 *          state machine (binary nodes) and 4 = 4-wide nodes, both measured slower
 *   STEP = 0 hole on, FP64 stepper, chord-lockstep loop (the default); 1 hole on, FP32 stepper (RRT_OPT_STEPPER = 1); 2 flat spacetime;
 *          A/B builds add 3 = hole on, FP64 stepper, decoupled stepping (RRT_OPT_TRACE_KERNEL = 2), measured slower (rrt_wavetrace.cuh)
 * bytes of one node record as the running kernel fetches it: 64 (two child boxes) or 32 per child of a wide node */
#ifndef RRT_AB
#define RRT_AB 0
#endif
static uint64_t node_record_bytes(const rrt_ctx* ctx) {
  if (ctx->opt[RRT_OPT_TRACE_KERNEL] == 1 || ctx->ds.width <= 2) return 64;
  return 32ull * (uint64_t)ctx->ds.width;
}
static int trace_variant(const rrt_ctx* ctx) {
  if (ctx->opt[RRT_OPT_TRACE_KERNEL] == 1) return 1;
  return ctx->ds.width == 8 ? 8 : ctx->ds.width == 4 ? 4 : 0;
}
/* RRT_OPT_STEPPER = 1: the FP32 stepper, for the lockstep kernels over binary and 8-wide nodes, hole on */
static bool fast_stepper(const rrt_ctx* ctx) {
  return ctx->opt[RRT_OPT_STEPPER] == 1 && ctx->bh.mode != RRT_MODE_FLAT && ctx->opt[RRT_OPT_TRACE_KERNEL] != 1 && ctx->ds.width != 4;
}
static int trace_step_mode(const rrt_ctx* ctx) {
  if (ctx->bh.mode == RRT_MODE_FLAT) return 2;
  if (fast_stepper(ctx)) return 1;
  return (RRT_AB && ctx->opt[RRT_OPT_TRACE_KERNEL] == 2) ? 3 : 0;
}
typedef void (*closest_fn)(const DevScene, const DevHole, const double*, size_t, const int*, int*, int*, double*, Counters*, const ParkQ);
typedef void (*shadow_fn)(const DevScene, const DevHole, const double*, size_t, const int*, int*, const uint32_t*, int*, const uint32_t*, Counters*, const ParkQ);
typedef void (*hook_fn)(const DevScene, const DevHole, int, const double*, const double*, const double*, const double*, rrt_hit*, int*, Counters*, const ParkQ);

/* park: 0 = rays run to their end in the trace kernel, 1 = escaping rays are parked for k_escape, 2 = the re-entry pass (rrt_wavetrace.cuh ParkQ);
 * 1 and 2 exist for STEP 0 (hole on, FP64 stepper, lockstep loop) over binary and 8-wide nodes */
template <bool COUNT, int V> static closest_fn closest_of_step(int step, int park) {
  switch (step) {
    case 1: if constexpr (V != 4) return k_trace_closest<COUNT, V, 1>; else return k_trace_closest<COUNT, V, 0>;
    case 2: return k_trace_closest<COUNT, V, 2>;
#if RRT_AB
    case 3: return k_trace_closest<COUNT, V, 3>;
#endif
    default:
#if RRT_AB
      if constexpr (V != 4) { if (park == 1) return k_trace_closest<COUNT, V, 0, 1>; if (park == 2) return k_trace_closest<COUNT, V, 0, 2>; }
#endif
      return k_trace_closest<COUNT, V, 0>;
  }
}
template <bool COUNT> static closest_fn closest_kernel(int v, int step, int park) {
  switch (v) {
#if RRT_AB
    case 1: return k_trace_closest<COUNT, 1, 0>;
    case 4: return closest_of_step<COUNT, 4>(step, 0);
#endif
    case 8: return closest_of_step<COUNT, 8>(step, park);
    default: return closest_of_step<COUNT, 0>(step, park);
  }
}
template <bool ANY, bool COUNT, int V> static shadow_fn shadow_of_step(int step, int park) {
  switch (step) {
    case 1: if constexpr (V != 4) return k_trace_shadow<ANY, COUNT, V, 1>; else return k_trace_shadow<ANY, COUNT, V, 0>;
    case 2: return k_trace_shadow<ANY, COUNT, V, 2>;
#if RRT_AB
    case 3: return k_trace_shadow<ANY, COUNT, V, 3>;
#endif
    default:
#if RRT_AB
      if constexpr (V != 4) { if (park == 1) return k_trace_shadow<ANY, COUNT, V, 0, 1>; if (park == 2) return k_trace_shadow<ANY, COUNT, V, 0, 2>; }
#endif
      return k_trace_shadow<ANY, COUNT, V, 0>;
  }
}
template <bool ANY, bool COUNT> static shadow_fn shadow_kernel(int v, int step, int park) {
  switch (v) {
#if RRT_AB
    case 1: return k_trace_shadow<ANY, COUNT, 1, 0>;
    case 4: return shadow_of_step<ANY, COUNT, 4>(step, 0);
#endif
    case 8: return shadow_of_step<ANY, COUNT, 8>(step, park);
    default: return shadow_of_step<ANY, COUNT, 0>(step, park);
  }
}
template <bool ANY, int V> static hook_fn hook_of_step(int step, int park) {
  switch (step) {
    case 1: if constexpr (V != 4) return k_hook_trace<ANY, V, 1>; else return k_hook_trace<ANY, V, 0>;
    case 2: return k_hook_trace<ANY, V, 2>;
#if RRT_AB
    case 3: return k_hook_trace<ANY, V, 3>;
#endif
    default:
#if RRT_AB
      if constexpr (V != 4) { if (park == 1) return k_hook_trace<ANY, V, 0, 1>; if (park == 2) return k_hook_trace<ANY, V, 0, 2>; }
#endif
      return k_hook_trace<ANY, V, 0>;
  }
}
template <bool ANY> static hook_fn hook_kernel(int v, int step, int park) {
  switch (v) {
#if RRT_AB
    case 1: return k_hook_trace<ANY, 1, 0>;
    case 4: return hook_of_step<ANY, 4>(step, 0);
#endif
    case 8: return hook_of_step<ANY, 8>(step, park);
    default: return hook_of_step<ANY, 0>(step, park);
  }
}
/* the escape hand-off applies to the default trace loop with the hole on and the FP64 stepper (RRT_OPT_ESCAPE, default 1) */
static bool escape_on(const rrt_ctx* ctx) {
#if RRT_AB
  const int v = trace_variant(ctx);
  return ctx->opt[RRT_OPT_ESCAPE] != 0 && trace_step_mode(ctx) == 0 && (v == 0 || v == 8);
#else
  (void)ctx; return false;   /* measured slower (rrt_wavetrace.cuh, ParkQ): compiled into A/B builds only */
#endif
}
/* park + re-entry queues in one allocation: 2 x (7 doubles + 2 ints) per ray + 4 counters; returns the bytes used */
static size_t park_bytes(size_t cap) { return 256 + 2 * (align_up(56 * cap) + align_up(8 * cap)); }
static void park_carve(char* base, size_t cap, ParkQ* park, ParkQ* reentry) {
  int* cnt = (int*)base; char* p = base + 256;
  park->cnt = cnt; reentry->cnt = cnt + 2; park->cap = reentry->cap = cap;
  park->st = (double*)p; p += align_up(56 * cap); park->meta = (int*)p; p += align_up(8 * cap);
  reentry->st = (double*)p; p += align_up(56 * cap); reentry->meta = (int*)p;
}

static void launch_closest(rrt_ctx* ctx, bool count, int grid, cudaStream_t st, const double* ray, size_t cap, const int* cnt_ptr, int* cursor, int* h_prim,
                           double* h_geo, Counters* gc, const ParkQ* park = nullptr, const ParkQ* reentry = nullptr) {
  const int v = trace_variant(ctx), step = trace_step_mode(ctx);
  const bool esc = park && escape_on(ctx);
  closest_fn fn = count ? closest_kernel<true>(v, step, esc ? 1 : 0) : closest_kernel<false>(v, step, esc ? 1 : 0);
  if (!esc) { fn<<<grid, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, ray, cap, cnt_ptr, cursor, h_prim, h_geo, gc, ParkQ()); return; }
#if RRT_AB
  cudaMemsetAsync(park->cnt, 0, 16, st);
  fn<<<grid, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, ray, cap, cnt_ptr, cursor, h_prim, h_geo, gc, *park);
  const int ge = ctx->sm_count * RRT_MINB_ESCAPE;
  if (count) k_escape_closest<true><<<ge, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, cap, h_prim, h_geo, gc, *park, *reentry);
  else k_escape_closest<false><<<ge, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, cap, h_prim, h_geo, gc, *park, *reentry);
  closest_fn fr = count ? closest_kernel<true>(v, step, 2) : closest_kernel<false>(v, step, 2);
  fr<<<grid, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, ray, cap, cnt_ptr, cursor, h_prim, h_geo, gc, *reentry);
#else
  (void)reentry;
#endif
}
template <bool ANY>
static void launch_shadow(rrt_ctx* ctx, bool count, int grid, cudaStream_t st, const double* ray, size_t cap, const int* cnt_ptr, int* cursor,
                          const uint32_t* cidx, int* res, const uint32_t* perm, Counters* gc, const ParkQ* park = nullptr, const ParkQ* reentry = nullptr) {
  const int v = trace_variant(ctx), step = trace_step_mode(ctx);
  const bool esc = park && escape_on(ctx);
  shadow_fn fn = count ? shadow_kernel<ANY, true>(v, step, esc ? 1 : 0) : shadow_kernel<ANY, false>(v, step, esc ? 1 : 0);
  if (!esc) { fn<<<grid, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, ray, cap, cnt_ptr, cursor, cidx, res, perm, gc, ParkQ()); return; }
#if RRT_AB
  cudaMemsetAsync(park->cnt, 0, 16, st);
  fn<<<grid, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, ray, cap, cnt_ptr, cursor, cidx, res, nullptr, gc, *park);
  const int ge = ctx->sm_count * RRT_MINB_ESCAPE;
  if (count) k_escape_shadow<true><<<ge, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, cap, cidx, res, gc, *park, *reentry);
  else k_escape_shadow<false><<<ge, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, cap, cidx, res, gc, *park, *reentry);
  shadow_fn fr = count ? shadow_kernel<ANY, true>(v, step, 2) : shadow_kernel<ANY, false>(v, step, 2);
  fr<<<grid, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, ray, cap, cnt_ptr, cursor, cidx, res, nullptr, gc, *reentry);
#else
  (void)reentry;
#endif
}
static void launch_hook(rrt_ctx* ctx, bool any_hit, int grid, cudaStream_t st, int n, const double* d_o, const double* d_d, const double* d_mn, const double* d_mx,
                        rrt_hit* d_out, int* d_cursor, Counters* gc, const ParkQ* park = nullptr, const ParkQ* reentry = nullptr) {
  const int v = trace_variant(ctx), step = trace_step_mode(ctx);
  const bool esc = park && escape_on(ctx);
  hook_fn fn = any_hit ? hook_kernel<true>(v, step, esc ? 1 : 0) : hook_kernel<false>(v, step, esc ? 1 : 0);
  if (!esc) { fn<<<grid, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, n, d_o, d_d, d_mn, d_mx, d_out, d_cursor, gc, ParkQ()); return; }
#if RRT_AB
  cudaMemsetAsync(park->cnt, 0, 16, st);
  fn<<<grid, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, n, d_o, d_d, d_mn, d_mx, d_out, d_cursor, gc, *park);
  const int ge = ctx->sm_count * RRT_MINB_ESCAPE;
  if (any_hit) k_escape_hook<true><<<ge, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, d_out, gc, *park, *reentry);
  else k_escape_hook<false><<<ge, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, d_out, gc, *park, *reentry);
  hook_fn fr = any_hit ? hook_kernel<true>(v, step, 2) : hook_kernel<false>(v, step, 2);
  fr<<<grid, RRT_BLOCK, 0, st>>>(ctx->ds, ctx->bh, n, d_o, d_d, d_mn, d_mx, d_out, d_cursor, gc, *reentry);
#else
  (void)reentry;
#endif
}

static int render_impl(rrt_ctx* ctx, const rrt_params* p, float* d_rgb, int* d_counts, float4* d_rgbn, cudaStream_t user_stream, rrt_stats* stats) {
  if (!ctx->have_scene || !ctx->have_cam || ctx->W <= 0) FAIL(RRT_E_STATE, "rrt_render: scene, camera and frame must be set first");
  if (p->ns_aa <= 0 || p->max_ray_depth < 0 || p->ns_area_light <= 0 || p->samples_per_batch <= 0) FAIL(RRT_E_ARG, "rrt_render: bad parameters");
  CK(cudaSetDevice(ctx->device));
  (void)user_stream;
  cudaStream_t st = ctx->stream;
  const bool count = ctx->opt[RRT_OPT_COUNTERS] != 0;
  const size_t npix = (size_t)ctx->W * ctx->H;

  RenderK rk;
  rk.W = ctx->W; rk.H = ctx->H; rk.ns_aa = p->ns_aa; rk.D = p->max_ray_depth; rk.ns_area_light = p->ns_area_light; rk.batch = p->samples_per_batch;
  rk.tol = p->max_tolerance; rk.hemi = p->direct_hemisphere_sample != 0; rk.thin_lens = p->thin_lens != 0; rk.seed = p->seed;
  const int n_lights_all = ctx->n_lights + (ctx->have_env ? 1 : 0);   /* scene->lights.size(): the environment light counts */
  if (rk.hemi) rk.S = rk.total = n_lights_all * p->ns_area_light;
  else rk.S = rk.total = ctx->n_delta_lights + (n_lights_all - ctx->n_delta_lights) * p->ns_area_light;
  if (rk.D == 0) rk.S = 0;
  /* 2 = choose: the per-sample kernel pays when evaluating the BSDF is expensive (microfacet: C3 34.2 -> 26.6 ms per
   * 32 spp) and costs a little when it is not (diffuse C1: 4.43 -> 4.74 ms), DESIGN.md section 4.4 */
  rk.defer_emit = (rk.D >= 1 && rk.S > 0) ? 1 : 0;   /* k_accumulate runs for every vertex index */
  rk.split = ctx->opt[RRT_OPT_SHADE_SPLIT] == 2 ? (ctx->has_microfacet ? 1 : 0) : (ctx->opt[RRT_OPT_SHADE_SPLIT] != 0);
  rk.direct_draws = 0;
  if (rk.hemi) rk.direct_draws = 2 * rk.S;
  else {
    for (const rrt_light& l : ctx->host_lights)
      if (l.type == RRT_LIGHT_AREA || l.type == RRT_LIGHT_HEMISPHERE) rk.direct_draws += 2 * (l.is_delta ? 1 : p->ns_area_light);
    if (ctx->have_env) rk.direct_draws += 2 * p->ns_area_light;
  }
  const int S = std::max(rk.S, 1);
  const int nb_max = std::min(rk.batch, rk.ns_aa);
  const int iters = std::max(rk.D, 1);

  /* ---- work buffers: capacity from the in-flight budget */
  const size_t per_path = 2 * (8 * 8 + 4 + 4 + 4 + 12) + 4 + 72 + 12 + 2 * (12 + 1 + 12 + 4) + 2 * (size_t)S * (56 + 4 + 4 + 16 + 2 + 4) + (escape_on(ctx) ? (size_t)S * 128 : 0);
  size_t want = p->max_paths_in_flight > 0 ? (size_t)p->max_paths_in_flight : (size_t)4 << 20;
  size_t budget = (size_t)(ctx->opt[RRT_OPT_WORK_BYTES] > 0 ? ctx->opt[RRT_OPT_WORK_BYTES] : (int64_t)12 << 30);
  want = std::min(want, std::max<size_t>(budget / per_path / (size_t)std::min<int64_t>(std::max<int64_t>(ctx->opt[RRT_OPT_STREAMS], 1), 4), 1024));
  size_t pix_per_chunk = std::max<size_t>(want / nb_max, 1);
  pix_per_chunk = std::min(pix_per_chunk, npix);
  /* buffers follow the frame, not a fixed budget: when rounds are split over NS streams no chunk is larger than a stream's share (or the
   * 1 M-path threshold below which a round stays whole), and a frame too small to ever be split gets ONE set of queues */
  const int NS_opt = (int)std::min<int64_t>(std::max<int64_t>(ctx->opt[RRT_OPT_STREAMS], 1), 4);
  const bool ever_split = NS_opt > 1 && npix * (size_t)nb_max >= (size_t)NS_opt * ((size_t)1 << 20);
  if (ever_split) pix_per_chunk = std::min(pix_per_chunk, std::max((npix + NS_opt - 1) / NS_opt, (((size_t)NS_opt << 20) + nb_max - 1) / nb_max));
  const size_t cap = pix_per_chunk * nb_max, capS = cap * S;
  if (capS >= ((size_t)1 << 31)) FAIL(RRT_E_ARG, "rrt_render: max_paths_in_flight x light samples exceeds 2^31");
  rk.cap = (int)cap;

  /* Wavefront chunks of one adaptive round are independent (disjoint pixels, slot-addressed results), so they are dealt round-robin over
   * NS streams, each with its own queues: a trace launch ends ragged — rays need 1 to 63 chords, and the last ones keep a few lanes alive —
   * and the other stream's kernels fill the SMs that tail leaves idle.  Matters most where launches are small: strong scaling (C4 on 8 GPUs:
   * 20 k pixels per rank and round).  RRT_OPT_STREAMS (1..4, default 2); results are bit-identical for any value. */
  const int NS = ever_split ? NS_opt : 1;
  cudaStream_t streams[4] = {st, nullptr, nullptr, nullptr};
  for (int k = 1; k < NS; ++k) {
    if (!ctx->aux[k - 1]) { CK(cudaStreamCreateWithFlags(&ctx->aux[k - 1], cudaStreamNonBlocking)); CK(cudaEventCreateWithFlags(&ctx->ev_join[k - 1], cudaEventDisableTiming)); }
    streams[k] = ctx->aux[k - 1];
  }
  if (!ctx->ev_fork) CK(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
  /* Shadow rays of depth i only feed k_accumulate(i), so they run on a stream of their own beside the extension rays of the NEXT TWO depths: the
   * shadow side of the queues (rays, contribution table, verdicts, the vertex's throughput / slot / emission term) exists once per parity of the
   * vertex index, so k_shade(i + 1) and k_shade_direct(i + 1) fill the other set while shadow(i) and k_accumulate(i) still read theirs, and
   * k_shade(i + 2) waits for k_accumulate(i).  Two or three big persistent launches are in flight and fill the SMs each other's ragged ends leave
   * idle.  Every addition into a path's radiance is made by k_accumulate, in vertex order, on one stream (k_shade hands the emission term of a
   * vertex reached through a delta BSDF over instead of adding it): the order of every float addition is what it was, results are bit-identical. */
  const bool overlap = ctx->opt[RRT_OPT_OVERLAP] != 0 && !escape_on(ctx) && rk.D >= 2 && rk.S > 0;
  cudaStream_t sstreams[4] = {nullptr, nullptr, nullptr, nullptr};
  for (int k = 0; k < NS; ++k) {
    if (overlap && !ctx->shadow_stream[k]) CK(cudaStreamCreateWithFlags(&ctx->shadow_stream[k], cudaStreamNonBlocking));
    sstreams[k] = overlap ? ctx->shadow_stream[k] : streams[k];
  }
  ctx->sync_used = 0;
  Queues qs[4]; memset(qs, 0, sizeof qs);
  ParkQ parks[4], reents[4];
  size_t off = 0;
  auto carve = [&](size_t bytes) { size_t o = off; off += align_up(bytes); return o; };
  size_t o_ray[2], o_slot[2], o_rngn[2], o_flags[2], o_beta[2];
  for (int b = 0; b < 2; ++b) { o_ray[b] = carve(64 * cap); o_slot[b] = carve(4 * cap); o_rngn[b] = carve(4 * cap); o_flags[b] = carve(4 * cap); o_beta[b] = carve(12 * cap); }
  size_t o_hprim = carve(4 * cap), o_hgeo = carve(72 * cap), o_L = carve(12 * cap);
  size_t o_dbeta[2], o_vflag[2], o_emit[2], o_aslot[2], o_sray[2], o_scidx[2], o_sres[2], o_scon[2];   /* shadow side: one set per parity of the vertex index */
  for (int b = 0; b < 2; ++b) {
    o_dbeta[b] = carve(12 * cap); o_vflag[b] = carve(cap); o_emit[b] = carve(12 * cap); o_aslot[b] = carve(4 * cap);
    o_sray[b] = carve(56 * capS); o_scidx[b] = carve(4 * capS); o_sres[b] = carve(4 * capS); o_scon[b] = carve(16 * capS);
  }
  const int ctl_stride = iters + 2;
  size_t o_ctl = carve(4 * 4 * (size_t)ctl_stride);
  const bool esc = escape_on(ctx);
  /* shadow rays handed out cell by cell (rrt_kernels.cuh, k_bin_*): hole on — a flat shadow ray is one traversal, too short to repay a sort */
  const bool sort_sh = ctx->opt[RRT_OPT_SORT_SHADOW] != 0 && ctx->bh.mode != RRT_MODE_FLAT && !esc && rk.D >= 1 && rk.S > 0 && trace_variant(ctx) == 0;   /* meshes */
  /* 1: only chunks under 2 M paths.  What the sort buys is mostly a shorter ragged end — rays that start together end together — and that is a
   * visible share of a small launch only: C4 whole, 4 M-path chunks, +-0; an 8-GPU rank's share of it, rounds of 8 M paths shrinking to 0.2 M, -2.3 % */
  const size_t sort_below = ctx->opt[RRT_OPT_SORT_SHADOW] == 1 ? ((size_t)2 << 20) : ~(size_t)0;
  size_t o_sbin[2] = {0, 0}, o_sperm[2] = {0, 0}, o_shist[2] = {0, 0};
  if (sort_sh) for (int b = 0; b < 2; ++b) { o_sbin[b] = carve(2 * capS); o_sperm[b] = carve(4 * capS); o_shist[b] = carve(8 * RRT_BINS); }
  BinGrid bing;
  for (int a3 = 0; a3 < 3; ++a3) {
    const float ext = ctx->ds.root_hi[a3] - ctx->ds.root_lo[a3];
    bing.lo[a3] = ctx->ds.root_lo[a3]; bing.inv[a3] = ext > 0.f ? (float)(1 << RRT_BIN_BITS) / ext : 0.f;
  }
  const size_t cap_park = std::max(cap, capS);
  size_t o_park = esc ? carve(park_bytes(cap_park)) : 0;
  const size_t per_stream = off;
  CK(ctx->d_work.ensure(per_stream * NS));
  for (int k = 0; k < NS; ++k) {
    Queues& q = qs[k];
    char* wb = ctx->d_work.as<char>() + per_stream * k;
    q.ctl_stride = ctl_stride;
    for (int b = 0; b < 2; ++b) {
      q.ray[b] = (double*)(wb + o_ray[b]); q.slot[b] = (uint32_t*)(wb + o_slot[b]); q.rngn[b] = (uint32_t*)(wb + o_rngn[b]);
      q.flags[b] = (uint32_t*)(wb + o_flags[b]); q.beta[b] = (float*)(wb + o_beta[b]);
    }
    q.h_prim = (int*)(wb + o_hprim); q.h_geo = (double*)(wb + o_hgeo);
    for (int b = 0; b < 2; ++b) {
      q.dbeta[b] = (float*)(wb + o_dbeta[b]); q.vflag[b] = (uint8_t*)(wb + o_vflag[b]); q.emit[b] = (float*)(wb + o_emit[b]); q.aslot[b] = (uint32_t*)(wb + o_aslot[b]);
      q.sray[b] = (double*)(wb + o_sray[b]); q.s_cidx[b] = (uint32_t*)(wb + o_scidx[b]); q.s_res[b] = (int*)(wb + o_sres[b]); q.s_contrib[b] = (float4*)(wb + o_scon[b]);
    }
    q.Lsample = (float*)(wb + o_L); q.ctl = (int*)(wb + o_ctl);
    if (sort_sh) for (int b = 0; b < 2; ++b) {
      q.s_bin[b] = (uint16_t*)(wb + o_sbin[b]); q.s_perm[b] = (uint32_t*)(wb + o_sperm[b]); q.s_hist[b] = (uint32_t*)(wb + o_shist[b]);
      CK(cudaMemsetAsync(q.s_hist[b], 0, 8 * RRT_BINS, st));   /* k_bin_scan leaves the counts at zero after every use */
    }
    if (esc) park_carve(wb + o_park, cap_park, &parks[k], &reents[k]);
  }
  const size_t ctl_bytes = 4 * 4 * (size_t)ctl_stride;

  /* ---- per-pixel accumulators and the active list */
  PixelAcc px;
  {
    size_t o_sum = 0, o_s1 = align_up(12 * npix), o_s2 = o_s1 + align_up(8 * npix), o_n = o_s2 + align_up(8 * npix), o_done = o_n + align_up(4 * npix),
           total = o_done + align_up(npix);
    CK(ctx->d_px.ensure(total));
    char* pb = ctx->d_px.as<char>();
    px.sum = (float*)(pb + o_sum); px.s1 = (double*)(pb + o_s1); px.s2 = (double*)(pb + o_s2); px.n = (int*)(pb + o_n); px.done = (uint8_t*)(pb + o_done);
    CK(cudaMemsetAsync(pb, 0, total, st));
  }
  CK(ctx->d_act.ensure(2 * 4 * npix));
  uint32_t* act[2] = {ctx->d_act.as<uint32_t>(), ctx->d_act.as<uint32_t>() + npix};
  CK(ctx->d_cnt.ensure(sizeof(Counters)));
  Counters* gcnt = ctx->d_cnt.as<Counters>();
  CK(cudaMemsetAsync(gcnt, 0, sizeof(Counters), st));

  ctx->events_used = 0;
  cudaEvent_t ev_begin = next_event(ctx), ev_end = next_event(ctx);
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> trace_ev;
  CK(cudaEventRecord(ev_begin, st));
  const auto t_render0 = std::chrono::steady_clock::now();
  ctx->sync_wait_ms = 0;

  KeepOwned own;
  own.W = ctx->W; own.tile = ctx->tile; own.ntx = rrt_shard_stride(ctx->world); own.rank = ctx->rank; own.world = ctx->world;
  own.x0 = 0; own.y0 = 0; own.x1 = ctx->W; own.y1 = ctx->H;
  if (ctx->cell_w > 0) {
    own.x0 = std::max(ctx->cell_x, 0); own.y0 = std::max(ctx->cell_y, 0);
    own.x1 = std::min(ctx->cell_x + ctx->cell_w, ctx->W); own.y1 = std::min(ctx->cell_y + ctx->cell_h, ctx->H);
  }
  uint32_t n_act = 0;
  int rc = compact(ctx, own, nullptr, (uint32_t)npix, act[0], &n_act);
  if (rc) return rc;

  const int grid_trace = ctx->sm_count * (trace_variant(ctx) >= 4 ? RRT_MINB_WIDE : RRT_MINB), grid_small = ctx->sm_count * 8;
  uint64_t launches = 3, trace_launches = 0;
  int a = 0;
  for (int base_sample = 0; base_sample < rk.ns_aa && n_act > 0;) {
    const int nb = std::min(rk.batch, rk.ns_aa - base_sample);
    /* chunks of this round: as few as the buffers allow, but a multiple of NS of equal size once every chunk still fills the machine */
    size_t nchunks = (n_act + pix_per_chunk - 1) / pix_per_chunk;
    /* (a chunk below ~1 M paths leaves under 10 rays per resident lane: launches that small are mostly ragged end, and splitting them over streams makes
     * them smaller still: C4 on 8 GPUs, 691 k paths per rank and round: 416 ms per frame in two chunks, 402 ms in one) */
    const int ns_round = (NS > 1 && (size_t)n_act * nb >= (size_t)NS * ((size_t)1 << 20)) ? NS : 1;
    nchunks = (nchunks + ns_round - 1) / ns_round * ns_round;
    const size_t chunk_px = ((size_t)n_act + nchunks - 1) / nchunks;
    if (ns_round > 1) {
      CK(cudaEventRecord(ctx->ev_fork, st));
      for (int k = 1; k < ns_round; ++k) CK(cudaStreamWaitEvent(streams[k], ctx->ev_fork, 0));
    }
    size_t ci = 0;
    for (size_t c0 = 0; c0 < n_act; c0 += chunk_px, ++ci) {
      const int npx = (int)std::min(chunk_px, (size_t)n_act - c0);
      const uint32_t* chunk = act[a] + c0;
      const int si = (int)(ci % ns_round);
      cudaStream_t cs = streams[si], ss = sstreams[si];
      const Queues& q = qs[si];
      cudaEvent_t ev_acc[2] = {nullptr, nullptr};   /* k_accumulate(it - 1), k_accumulate(it - 2) on ss: the latter frees the shadow-side buffers k_shade(it) fills */
      CK(cudaMemsetAsync(q.ctl, 0, ctl_bytes, cs));
      k_generate<<<grid_small, 256, 0, cs>>>(ctx->cam, rk, q, chunk, npx, base_sample, nb);
      launches++;
      for (int it = 0; it < iters; ++it) {
        const int cur = it & 1;
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (stats) { e0 = next_event(ctx); e1 = next_event(ctx); CK(cudaEventRecord(e0, cs)); }
        launch_closest(ctx, count, grid_trace, cs, q.ray[cur], cap, q.ctl + it, q.ctl + 2 * q.ctl_stride + it, q.h_prim, q.h_geo, gcnt, esc ? &parks[si] : nullptr,
                       esc ? &reents[si] : nullptr);
        if (stats) { CK(cudaEventRecord(e1, cs)); trace_ev.push_back({e0, e1}); }
        if (ev_acc[1]) { CK(cudaStreamWaitEvent(cs, ev_acc[1], 0)); ev_acc[1] = nullptr; }
        k_shade<<<grid_small, RRT_BLOCK, 0, cs>>>(ctx->ds, ctx->bh, rk, q, chunk, base_sample, nb, it);
        if (rk.split && rk.D >= 1 && rk.S > 0) {
          k_shade_direct<<<ctx->sm_count * 2 * RRT_DIRECT_MINB, 256, 0, cs>>>(ctx->ds, rk, q, chunk, base_sample, nb, it);
          launches++;
        }
        launches += 2 + (esc ? 2 : 0); trace_launches++;
        if (rk.D >= 1 && rk.S > 0) {
          if (ss != cs) { cudaEvent_t ev = next_sync_event(ctx); CK(cudaEventRecord(ev, cs)); CK(cudaStreamWaitEvent(ss, ev, 0)); }
          if (stats) { e0 = next_event(ctx); e1 = next_event(ctx); CK(cudaEventRecord(e0, ss)); }
          int* scount = q.ctl + q.ctl_stride + it; int* scursor = q.ctl + 3 * q.ctl_stride + it;
          const ParkQ* pkq = esc ? &parks[si] : nullptr; const ParkQ* req = esc ? &reents[si] : nullptr;
          const int sp = it & 1;
          const uint32_t* perm = nullptr;
          if (sort_sh && (size_t)npx * nb < sort_below) {
            k_bin_count<<<grid_small, 256, 0, ss>>>(q.sray[sp], capS, scount, bing, q.s_hist[sp], q.s_bin[sp]);
            k_bin_scan<<<1, 1024, 0, ss>>>(q.s_hist[sp], q.s_hist[sp] + RRT_BINS);
            k_bin_scatter<<<grid_small, 256, 0, ss>>>(q.s_bin[sp], scount, q.s_hist[sp] + RRT_BINS, q.s_perm[sp]);
            perm = q.s_perm[sp]; launches += 3;
          }
          if (rk.hemi) launch_shadow<false>(ctx, count, grid_trace, ss, q.sray[sp], capS, scount, scursor, q.s_cidx[sp], q.s_res[sp], perm, gcnt, pkq, req);
          else launch_shadow<true>(ctx, count, grid_trace, ss, q.sray[sp], capS, scount, scursor, q.s_cidx[sp], q.s_res[sp], perm, gcnt, pkq, req);
          if (stats) { CK(cudaEventRecord(e1, ss)); trace_ev.push_back({e0, e1}); }
          k_accumulate<<<grid_small, 256, 0, ss>>>(ctx->ds, rk, q, it);
          ev_acc[1] = ev_acc[0];
          if (ss != cs) { ev_acc[0] = next_sync_event(ctx); CK(cudaEventRecord(ev_acc[0], ss)); }
          launches += 2 + (esc ? 2 : 0); trace_launches++;
        }
        if (iters > 8 && it + 1 < iters && (it & 3) == 3) {  /* long depth budgets (-m 100): stop once every path has died */
          CK(cudaMemcpyAsync(ctx->h_pinned, q.ctl + it + 1, 4, cudaMemcpyDeviceToHost, cs));
          CK(cudaStreamSynchronize(cs));
          if (ctx->h_pinned[0] == 0) break;
        }
      }
      if (ev_acc[0]) CK(cudaStreamWaitEvent(cs, ev_acc[0], 0));   /* stream order on ss: every earlier k_accumulate is done too */
      k_resolve<<<grid_small, 256, 0, cs>>>(rk, q, px, chunk, npx, nb);
      launches++;
    }
    for (int k = 1; k < ns_round; ++k) {   /* join: the round's compaction and the next round read what every stream wrote */
      CK(cudaEventRecord(ctx->ev_join[k - 1], streams[k]));
      CK(cudaStreamWaitEvent(st, ctx->ev_join[k - 1], 0));
    }
    base_sample += nb;
    if (base_sample < rk.ns_aa) {
      KeepActive ka; ka.done = px.done;
      uint32_t n_next = 0;
      rc = compact(ctx, ka, act[a], n_act, act[a ^ 1], &n_next);
      if (rc) return rc;
      launches += 3;
      a ^= 1; n_act = n_next;
      if (ctx->progress_fn && n_act > 0) {   /* progressive output / stop(): include/rrt_b200.h */
        CK(cudaStreamSynchronize(st));
        ctx->snap_px = px; ctx->snap_npix = npix; ctx->in_progress_cb = true;
        const int stop = ctx->progress_fn(ctx->progress_user, base_sample, rk.ns_aa, (int64_t)n_act);
        ctx->in_progress_cb = false;
        if (stop) break;
      }
    }
  }
  k_finalize<<<grid_small, 256, 0, st>>>(px, npix, d_rgb, d_counts, d_rgbn, gcnt);
  launches++;
  CK(cudaEventRecord(ev_end, st));
  /* RRT_ENQUEUE_TIMING=1: how far ahead of the device the host's launch loop runs (is a frame of many small launches bound by the enqueue rate?) */
  static const bool enqueue_timing = getenv("RRT_ENQUEUE_TIMING") != nullptr;
  const auto t_enq = std::chrono::steady_clock::now();
  CK(cudaStreamSynchronize(st));
  if (enqueue_timing) {
    const auto t_done = std::chrono::steady_clock::now();
    const double host_ms = std::chrono::duration<double, std::milli>(t_enq - t_render0).count();
    fprintf(stderr, "[rrt] render: %llu launches; host %.2f ms in the launch loop = %.2f ms issuing + %.2f ms blocked on round-end read-backs; device done %.2f ms after the last launch\n",
            (unsigned long long)launches, host_ms, host_ms - ctx->sync_wait_ms, ctx->sync_wait_ms, std::chrono::duration<double, std::milli>(t_done - t_enq).count());
  }
  CK(cudaGetLastError());
  if (stats) {
    memset(stats, 0, sizeof *stats);
    float ms = 0; CK(cudaEventElapsedTime(&ms, ev_begin, ev_end));
    stats->seconds = ms * 1e-3;
    /* time during which at least one trace kernel was running: the launches of different streams overlap, so their durations are
     * merged as intervals on the device clock (relative to the render's first event) instead of summed */
    std::vector<std::pair<float, float>> iv;
    iv.reserve(trace_ev.size());
    for (auto& pr : trace_ev) {
      float t0 = 0, t1 = 0;
      CK(cudaEventElapsedTime(&t0, ev_begin, pr.first)); CK(cudaEventElapsedTime(&t1, ev_begin, pr.second));
      iv.push_back({t0, t1});
    }
    std::sort(iv.begin(), iv.end());
    double tms = 0; float lo = 0, hi = -1;
    for (auto& x : iv) {
      if (hi < lo || x.first > hi) { if (hi >= lo) tms += hi - lo; lo = x.first; hi = x.second; }
      else hi = std::max(hi, x.second);
    }
    if (hi >= lo) tms += hi - lo;
    stats->trace_seconds = tms * 1e-3;
    Counters hc;
    CK(cudaMemcpy(&hc, gcnt, sizeof hc, cudaMemcpyDeviceToHost));
    stats->rays = hc.rays; stats->geodesic_steps = hc.steps; stats->node_visits = hc.nodes; stats->tri_tests = hc.tris; stats->sphere_tests = hc.sph;
    stats->kernel_launches = launches; stats->trace_launches = trace_launches;
    stats->algorithmic_bytes = node_record_bytes(ctx) * hc.nodes + 80 * hc.tris + 32 * hc.sph + 160 * hc.rays;
    stats->paths = hc.paths; /* sum of sample counts */
  }
  return RRT_OK;
}

extern "C" {

int rrt_render(rrt_ctx* ctx, const rrt_params* p, float* hdr_rgb, int32_t* sample_counts, rrt_stats* stats) {
  if (!ctx || !p) return RRT_E_ARG;
  if (!ctx->subs.empty()) return multi_render(ctx, p, hdr_rgb, sample_counts, nullptr, nullptr, stats);
  const size_t npix = (size_t)ctx->W * ctx->H;
  CK(cudaSetDevice(ctx->device));
  CK(ctx->d_out_rgb.ensure(12 * npix)); CK(ctx->d_out_cnt.ensure(4 * npix));
  int rc = render_impl(ctx, p, ctx->d_out_rgb.as<float>(), ctx->d_out_cnt.as<int>(), nullptr, nullptr, stats);
  if (rc) return rc;
  if (hdr_rgb) CK(cudaMemcpy(hdr_rgb, ctx->d_out_rgb.p, 12 * npix, cudaMemcpyDeviceToHost));
  if (sample_counts) CK(cudaMemcpy(sample_counts, ctx->d_out_cnt.p, 4 * npix, cudaMemcpyDeviceToHost));
  return RRT_OK;
}

int rrt_render_device(rrt_ctx* ctx, const rrt_params* p, void* d_rgbn, void* d_counts, void* stream, rrt_stats* stats) {
  if (!ctx || !p) return RRT_E_ARG;
  if (!ctx->subs.empty()) {
    if (!d_rgbn) FAIL(RRT_E_ARG, "rrt_render_device: d_rgbn is NULL");
    if (stream) CK(cudaStreamSynchronize((cudaStream_t)stream));
    return multi_render(ctx, p, nullptr, nullptr, d_rgbn, d_counts, stats);
  }
  if (stream) CK(cudaStreamSynchronize((cudaStream_t)stream));
  return render_impl(ctx, p, nullptr, (int*)d_counts, (float4*)d_rgbn, (cudaStream_t)stream, stats);
}

/* ---------------------------------------------------------------- parity hooks */
struct HookStage {  // stage host arrays on the device, fetch results back
  rrt_ctx* ctx; size_t in_off = 0, out_off = 0;
  std::vector<std::pair<size_t, std::pair<void*, size_t>>> outs;
};

static int hook_begin(rrt_ctx* ctx, size_t in_bytes, size_t out_bytes) {
  CK(cudaSetDevice(ctx->device));
  CK(ctx->d_hook_in.ensure(in_bytes + 4096)); CK(ctx->d_hook_out.ensure(out_bytes + 4096));
  return RRT_OK;
}
static void* hook_in(rrt_ctx* ctx, size_t& off, const void* src, size_t bytes) {
  if (!src) return nullptr;
  char* dst = ctx->d_hook_in.as<char>() + off;
  cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream);
  off += align_up(bytes);
  return dst;
}
static void* hook_out(rrt_ctx* ctx, size_t& off, size_t bytes) {
  char* dst = ctx->d_hook_out.as<char>() + off;
  off += align_up(bytes);
  return dst;
}
static int grid_for(long long n, int block) { long long g = (n + block - 1) / block; return (int)std::max<long long>(1, std::min<long long>(g, 65535)); }

int rrt_trace_rays(rrt_ctx* ctx, int64_t n, const double* o, const double* d, const double* min_t, const double* max_t, int any_hit, rrt_hit* out,
                   rrt_stats* stats) {
  RRT_MULTI_FWD(ctx, rrt_trace_rays(ctx0, n, o, d, min_t, max_t, any_hit, out, stats));
  if (!ctx || n < 0 || (n && (!o || !d || !out))) return RRT_E_ARG;
  if (!ctx->have_scene) FAIL(RRT_E_STATE, "rrt_trace_rays: no scene");
  if (stats) memset(stats, 0, sizeof *stats);
  if (n == 0) return RRT_OK;
  if (n >= ((int64_t)1 << 30)) FAIL(RRT_E_ARG, "rrt_trace_rays: at most 2^30 rays per call");
  const bool esc = escape_on(ctx);
  int rc = hook_begin(ctx, 64 * (size_t)n + 4096, sizeof(rrt_hit) * (size_t)n + 1024 + (esc ? park_bytes((size_t)n) + 512 : 0));
  if (rc) return rc;
  size_t io = 0, oo = 0;
  double* d_o = (double*)hook_in(ctx, io, o, 24 * (size_t)n); double* d_d = (double*)hook_in(ctx, io, d, 24 * (size_t)n);
  double* d_mn = (double*)hook_in(ctx, io, min_t, 8 * (size_t)n); double* d_mx = (double*)hook_in(ctx, io, max_t, 8 * (size_t)n);
  rrt_hit* d_out = (rrt_hit*)hook_out(ctx, oo, sizeof(rrt_hit) * (size_t)n);
  Counters* gc = (Counters*)hook_out(ctx, oo, sizeof(Counters));
  int* d_cursor = (int*)hook_out(ctx, oo, 256);
  ParkQ park, reentry;
  if (esc) park_carve((char*)hook_out(ctx, oo, park_bytes((size_t)n)), (size_t)n, &park, &reentry);
  CK(cudaMemsetAsync(gc, 0, sizeof(Counters), ctx->stream));
  CK(cudaMemsetAsync(d_cursor, 0, 4, ctx->stream));
  cudaEvent_t e0, e1; ctx->events_used = 0; e0 = next_event(ctx); e1 = next_event(ctx);
  CK(cudaEventRecord(e0, ctx->stream));
  /* as many resident blocks as a render's trace launches get, so that a timing taken through this hook reads like theirs */
  int grid = (int)std::min<int64_t>((n + RRT_BLOCK - 1) / RRT_BLOCK, (int64_t)ctx->sm_count * (trace_variant(ctx) >= 4 ? RRT_MINB_WIDE : RRT_MINB));
  launch_hook(ctx, any_hit != 0, grid, ctx->stream, (int)n, d_o, d_d, d_mn, d_mx, d_out, d_cursor, gc, esc ? &park : nullptr, esc ? &reentry : nullptr);
  CK(cudaEventRecord(e1, ctx->stream));
  CK(cudaMemcpyAsync(out, d_out, sizeof(rrt_hit) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  if (stats) {
    Counters hc; CK(cudaMemcpy(&hc, gc, sizeof hc, cudaMemcpyDeviceToHost));
    float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
    stats->seconds = stats->trace_seconds = ms * 1e-3;
    stats->rays = hc.rays; stats->geodesic_steps = hc.steps; stats->node_visits = hc.nodes; stats->tri_tests = hc.tris; stats->sphere_tests = hc.sph;
    stats->kernel_launches = esc ? 3 : 1; stats->trace_launches = 1;
    stats->algorithmic_bytes = node_record_bytes(ctx) * hc.nodes + 80 * hc.tris + 32 * hc.sph + 160 * hc.rays;
  }
  return RRT_OK;
}

int rrt_geodesic(rrt_ctx* ctx, int64_t n, const double* o, const double* d, int nsteps, double* seg_out) {
  RRT_MULTI_FWD(ctx, rrt_geodesic(ctx0, n, o, d, nsteps, seg_out));
  if (!ctx || n < 0 || nsteps < 0 || (n && (!o || !d || !seg_out))) return RRT_E_ARG;
  if (n == 0 || nsteps == 0) return RRT_OK;
  size_t ob = 64 * (size_t)n * nsteps;
  int rc = hook_begin(ctx, 48 * (size_t)n + 1024, ob);
  if (rc) return rc;
  size_t io = 0, oo = 0;
  double* d_o = (double*)hook_in(ctx, io, o, 24 * (size_t)n); double* d_d = (double*)hook_in(ctx, io, d, 24 * (size_t)n);
  double* d_s = (double*)hook_out(ctx, oo, ob);
  if (ctx->opt[RRT_OPT_STEPPER] == 1) k_hook_geodesic_fp32<<<grid_for(n, 128), 128, 0, ctx->stream>>>(ctx->bh, n, d_o, d_d, nsteps, d_s);
  else k_hook_geodesic<<<grid_for(n, 128), 128, 0, ctx->stream>>>(ctx->bh, n, d_o, d_d, nsteps, d_s);
  CK(cudaMemcpyAsync(seg_out, d_s, ob, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  return RRT_OK;
}

int rrt_generate_rays(rrt_ctx* ctx, int64_t n, const double* xy, double* o, double* d) {
  RRT_MULTI_FWD(ctx, rrt_generate_rays(ctx0, n, xy, o, d));
  if (!ctx || n < 0 || (n && (!xy || !o || !d))) return RRT_E_ARG;
  if (!ctx->have_cam) FAIL(RRT_E_STATE, "rrt_generate_rays: no camera");
  if (n == 0) return RRT_OK;
  int rc = hook_begin(ctx, 16 * (size_t)n, 48 * (size_t)n + 1024);
  if (rc) return rc;
  size_t io = 0, oo = 0;
  double* d_xy = (double*)hook_in(ctx, io, xy, 16 * (size_t)n);
  double* d_o = (double*)hook_out(ctx, oo, 24 * (size_t)n); double* d_d = (double*)hook_out(ctx, oo, 24 * (size_t)n);
  k_hook_camera<<<grid_for(n, 128), 128, 0, ctx->stream>>>(ctx->cam, n, d_xy, d_o, d_d);
  CK(cudaMemcpyAsync(o, d_o, 24 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(d, d_d, 24 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  return RRT_OK;
}

int rrt_generate_rays_thin_lens(rrt_ctx* ctx, int64_t n, const double* xy, const double* lens, double* o, double* d) {
  RRT_MULTI_FWD(ctx, rrt_generate_rays_thin_lens(ctx0, n, xy, lens, o, d));
  if (!ctx || n < 0 || (n && (!xy || !lens || !o || !d))) return RRT_E_ARG;
  if (!ctx->have_cam) FAIL(RRT_E_STATE, "rrt_generate_rays_thin_lens: no camera");
  if (n == 0) return RRT_OK;
  int rc = hook_begin(ctx, 32 * (size_t)n + 1024, 48 * (size_t)n + 1024);
  if (rc) return rc;
  size_t io = 0, oo = 0;
  double* d_xy = (double*)hook_in(ctx, io, xy, 16 * (size_t)n); double* d_lens = (double*)hook_in(ctx, io, lens, 16 * (size_t)n);
  double* d_o = (double*)hook_out(ctx, oo, 24 * (size_t)n); double* d_d = (double*)hook_out(ctx, oo, 24 * (size_t)n);
  k_hook_camera_lens<<<grid_for(n, 128), 128, 0, ctx->stream>>>(ctx->cam, n, d_xy, d_lens, d_o, d_d);
  CK(cudaMemcpyAsync(o, d_o, 24 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(d, d_d, 24 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  return RRT_OK;
}

static int bsdf_hook(rrt_ctx* ctx, int mat, int64_t n, const double* wo, const double* wi_in, const double* u, float* f, double* wi, float* pdf) {
  if (!ctx->have_scene) FAIL(RRT_E_STATE, "rrt_bsdf: no scene");
  if (mat < 0 || mat >= ctx->ds.n_mats) FAIL(RRT_E_ARG, "rrt_bsdf: material index out of range");
  if (n == 0) return RRT_OK;
  int rc = hook_begin(ctx, 80 * (size_t)n + 2048, 40 * (size_t)n + 2048);
  if (rc) return rc;
  size_t io = 0, oo = 0;
  double* d_wo = (double*)hook_in(ctx, io, wo, 24 * (size_t)n); double* d_wi_in = (double*)hook_in(ctx, io, wi_in, 24 * (size_t)n);
  double* d_u = (double*)hook_in(ctx, io, u, 32 * (size_t)n);
  float* d_f = (float*)hook_out(ctx, oo, 12 * (size_t)n); double* d_wi = (double*)hook_out(ctx, oo, 24 * (size_t)n); float* d_pdf = (float*)hook_out(ctx, oo, 4 * (size_t)n);
  k_hook_bsdf<<<grid_for(n, 128), 128, 0, ctx->stream>>>(ctx->ds.mats, mat, n, d_wo, d_wi_in, d_u, d_f, d_wi, d_pdf);
  CK(cudaMemcpyAsync(f, d_f, 12 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  if (wi) CK(cudaMemcpyAsync(wi, d_wi, 24 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  if (pdf) CK(cudaMemcpyAsync(pdf, d_pdf, 4 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  return RRT_OK;
}

int rrt_bsdf_sample(rrt_ctx* ctx, int mat, int64_t n, const double* wo, const double* u, float* f, double* wi, float* pdf) {
  RRT_MULTI_FWD(ctx, rrt_bsdf_sample(ctx0, mat, n, wo, u, f, wi, pdf));
  if (!ctx || n < 0 || (n && (!wo || !u || !f || !wi || !pdf))) return RRT_E_ARG;
  return bsdf_hook(ctx, mat, n, wo, nullptr, u, f, wi, pdf);
}

int rrt_bsdf_eval(rrt_ctx* ctx, int mat, int64_t n, const double* wo, const double* wi, float* f) {
  RRT_MULTI_FWD(ctx, rrt_bsdf_eval(ctx0, mat, n, wo, wi, f));
  if (!ctx || n < 0 || (n && (!wo || !wi || !f))) return RRT_E_ARG;
  return bsdf_hook(ctx, mat, n, wo, wi, nullptr, f, nullptr, nullptr);
}

int rrt_light_sample(rrt_ctx* ctx, int light, int64_t n, const double* p, const double* u, float* radiance, double* wi, float* dist, float* pdf) {
  RRT_MULTI_FWD(ctx, rrt_light_sample(ctx0, light, n, p, u, radiance, wi, dist, pdf));
  if (!ctx || n < 0 || (n && (!p || !u || !radiance || !wi || !dist || !pdf))) return RRT_E_ARG;
  if (!ctx->have_scene) FAIL(RRT_E_STATE, "rrt_light_sample: no scene");
  const bool is_env = ctx->have_env && light == ctx->ds.n_lights;   /* the environment light is the last light */
  if (light < 0 || (light >= ctx->ds.n_lights && !is_env)) FAIL(RRT_E_ARG, "rrt_light_sample: light index out of range");
  if (n == 0) return RRT_OK;
  int rc = hook_begin(ctx, 40 * (size_t)n + 1024, 44 * (size_t)n + 2048);
  if (rc) return rc;
  size_t io = 0, oo = 0;
  double* d_p = (double*)hook_in(ctx, io, p, 24 * (size_t)n); double* d_u = (double*)hook_in(ctx, io, u, 16 * (size_t)n);
  float* d_rad = (float*)hook_out(ctx, oo, 12 * (size_t)n); double* d_wi = (double*)hook_out(ctx, oo, 24 * (size_t)n);
  float* d_dist = (float*)hook_out(ctx, oo, 4 * (size_t)n); float* d_pdf = (float*)hook_out(ctx, oo, 4 * (size_t)n);
  if (is_env) k_hook_env_sample<<<grid_for(n, 128), 128, 0, ctx->stream>>>(ctx->ds, n, d_u, d_rad, d_wi, d_dist, d_pdf);
  else k_hook_light<<<grid_for(n, 128), 128, 0, ctx->stream>>>(ctx->ds.lights, light, n, d_p, d_u, d_rad, d_wi, d_dist, d_pdf);
  CK(cudaMemcpyAsync(radiance, d_rad, 12 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(wi, d_wi, 24 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(dist, d_dist, 4 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(pdf, d_pdf, 4 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  return RRT_OK;
}

/* EnvironmentLight::EnvironmentLight + init (reference src/static_scene/environment_light.cpp:10-49): the map and its
 * sampling tables (per-texel pdf, per-row conditional CDF, marginal CDF of the rows), built on the host in the reference's
 * summation order and kept on the device. */
int rrt_set_envmap(rrt_ctx* ctx, int width, int height, const float* rgb) {
  RRT_MULTI_ALL(ctx, rrt_set_envmap(sub, width, height, rgb));
  if (!ctx) return RRT_E_ARG;
  DevScene& d = ctx->ds;
  if (!rgb || width <= 0 || height <= 0) {
    ctx->have_env = false; d.env = nullptr; d.env_pdf = d.env_cond = d.env_marg = nullptr; d.env_w = d.env_h = 0;
    return RRT_OK;
  }
  const size_t w = (size_t)width, h = (size_t)height;
  if (w * h >= ((size_t)1 << 31)) FAIL(RRT_E_ARG, "rrt_set_envmap: map too large");
  std::vector<double> pdf(w * h), cond(w * h), marg(h, 0.0);
  double sum = 0;
  for (size_t j = 0; j < h; ++j)
    for (size_t i = 0; i < w; ++i) {
      const float* q = rgb + 3 * (w * j + i);
      const float illum = 0.2126f * q[0] + 0.7152f * q[1] + 0.0722f * q[2];
      pdf[w * j + i] = illum * sin(3.14159265358979323 * ((int)j + .5) / (uint32_t)h);
      sum += pdf[w * j + i];
    }
  for (size_t j = 0; j < h; ++j) {
    for (size_t i = 0; i < w; ++i) marg[j] += (pdf[w * j + i] /= sum);
    for (size_t i = 0; i < w; ++i) {
      cond[w * j + i] = pdf[w * j + i] / marg[j];
      if (i > 0) cond[w * j + i] += cond[w * j + i - 1];
    }
    if (j > 0) marg[j] += marg[j - 1];
  }
  CK(cudaStreamSynchronize(ctx->stream));
  int rc;
  if ((rc = upload(ctx, ctx->d_env, rgb, 12 * w * h))) return rc;
  if ((rc = upload(ctx, ctx->d_env_pdf, pdf.data(), 8 * w * h))) return rc;
  if ((rc = upload(ctx, ctx->d_env_cond, cond.data(), 8 * w * h))) return rc;
  if ((rc = upload(ctx, ctx->d_env_marg, marg.data(), 8 * h))) return rc;
  CK(cudaStreamSynchronize(ctx->stream));
  d.env = ctx->d_env.as<float>(); d.env_pdf = ctx->d_env_pdf.as<double>(); d.env_cond = ctx->d_env_cond.as<double>();
  d.env_marg = ctx->d_env_marg.as<double>(); d.env_w = width; d.env_h = height;
  ctx->have_env = true;
  return RRT_OK;
}

/* EnvironmentLight::sample_dir for n directions d [n*3] -> radiance [n*3] (parity hook) */
int rrt_env_lookup(rrt_ctx* ctx, int64_t n, const double* dir, float* radiance) {
  RRT_MULTI_FWD(ctx, rrt_env_lookup(ctx0, n, dir, radiance));
  if (!ctx || n < 0 || (n && (!dir || !radiance))) return RRT_E_ARG;
  if (!ctx->have_env) FAIL(RRT_E_STATE, "rrt_env_lookup: no environment map");
  if (n == 0) return RRT_OK;
  int rc = hook_begin(ctx, 24 * (size_t)n + 1024, 12 * (size_t)n + 1024);
  if (rc) return rc;
  size_t io = 0, oo = 0;
  double* d_d = (double*)hook_in(ctx, io, dir, 24 * (size_t)n);
  float* d_rad = (float*)hook_out(ctx, oo, 12 * (size_t)n);
  k_hook_env_lookup<<<grid_for(n, 128), 128, 0, ctx->stream>>>(ctx->ds, n, d_d, d_rad);
  CK(cudaMemcpyAsync(radiance, d_rad, 12 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  return RRT_OK;
}

int rrt_philox(rrt_ctx* ctx, uint64_t seed, int64_t n, const uint32_t* counters, uint32_t* out) {
  RRT_MULTI_FWD(ctx, rrt_philox(ctx0, seed, n, counters, out));
  if (!ctx || n < 0 || (n && (!counters || !out))) return RRT_E_ARG;
  if (n == 0) return RRT_OK;
  int rc = hook_begin(ctx, 16 * (size_t)n, 16 * (size_t)n);
  if (rc) return rc;
  size_t io = 0, oo = 0;
  uint32_t* d_c = (uint32_t*)hook_in(ctx, io, counters, 16 * (size_t)n);
  uint32_t* d_o = (uint32_t*)hook_out(ctx, oo, 16 * (size_t)n);
  k_hook_philox<<<grid_for(n, 128), 128, 0, ctx->stream>>>(seed, n, d_c, d_o);
  CK(cudaMemcpyAsync(out, d_o, 16 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  return RRT_OK;
}

int rrt_tonemap(rrt_ctx* ctx, int64_t n, const float* hdr_rgb, uint32_t* rgba) {
  RRT_MULTI_FWD(ctx, rrt_tonemap(ctx0, n, hdr_rgb, rgba));
  if (!ctx || n < 0 || (n && (!hdr_rgb || !rgba))) return RRT_E_ARG;
  if (n == 0) return RRT_OK;
  int rc = hook_begin(ctx, 12 * (size_t)n, 4 * (size_t)n);
  if (rc) return rc;
  size_t io = 0, oo = 0;
  float* d_h = (float*)hook_in(ctx, io, hdr_rgb, 12 * (size_t)n);
  uint32_t* d_o = (uint32_t*)hook_out(ctx, oo, 4 * (size_t)n);
  k_tonemap<<<grid_for(n, 256), 256, 0, ctx->stream>>>(n, d_h, d_o);
  CK(cudaMemcpyAsync(rgba, d_o, 4 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  return RRT_OK;
}

}  // extern "C"

/* ---------------------------------------------------------------- rrt_trace_rays_static: the parity hook's trace kernel without any allocation.
 * No counterpart in the reference; added for callers that cannot call cudaMalloc.
 * All device scratch of the call (staged rays, hit records, the work cursor, the counters) lives in ONE file-scope array
 * that every GPU loading the library holds until exit.  Layout (offsets in bytes, n_max = RRT_STATIC_MAX_RAYS):
 *   [0, 256)  work cursor (int, must start at 0)     [256, 512)  Counters (must start at 0)     [4096 ...)  o, d, min_t, max_t, hits
 * In HOST code the name g_static_scratch is only a stand-in for the symbol, not a device address: the address the kernel
 * gets comes from cudaGetSymbolAddress, asked once per device with that device current. */
#define RRT_STATIC_SCRATCH_BYTES (4096 + (size_t)RRT_STATIC_MAX_RAYS * (24 + 24 + 8 + 8 + sizeof(rrt_hit)))
static __device__ __align__(256) unsigned char g_static_scratch[RRT_STATIC_SCRATCH_BYTES];

#include <mutex>
static std::mutex g_static_mu;
static unsigned char* g_static_ptr[64];   /* per device, filled on first use */

static int static_scratch_for(rrt_ctx* ctx, unsigned char** out) {
  if (ctx->device < 0 || ctx->device >= 64) FAIL(RRT_E_ARG, "rrt_trace_rays_static: device index beyond 63");
  std::lock_guard<std::mutex> lk(g_static_mu);
  CK(cudaSetDevice(ctx->device));
  if (!g_static_ptr[ctx->device]) {
    void* p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p) FAIL(RRT_E_CUDA, "rrt_trace_rays_static: cudaGetSymbolAddress returned a null address");
    g_static_ptr[ctx->device] = (unsigned char*)p;
  }
  *out = g_static_ptr[ctx->device];
  return RRT_OK;
}

extern "C" int rrt_trace_rays_static(rrt_ctx* ctx, int64_t n, const double* o, const double* d, const double* min_t, const double* max_t, int any_hit,
                                     rrt_hit* out) {
  RRT_MULTI_FWD(ctx, rrt_trace_rays_static(ctx0, n, o, d, min_t, max_t, any_hit, out));
  if (!ctx || n < 0 || (n && (!o || !d || !out))) return RRT_E_ARG;
  if (!ctx->have_scene) FAIL(RRT_E_STATE, "rrt_trace_rays_static: no scene");
  if (n > RRT_STATIC_MAX_RAYS) {
    char b[160]; snprintf(b, sizeof b, "rrt_trace_rays_static: %lld rays exceed the static scratch limit of %d rays per call", (long long)n, RRT_STATIC_MAX_RAYS);
    FAIL(RRT_E_ARG, b);
  }
  if (n == 0) return RRT_OK;
  unsigned char* base = nullptr;
  int rc = static_scratch_for(ctx, &base);
  if (rc) return rc;
  const size_t N = (size_t)RRT_STATIC_MAX_RAYS;
  int* d_cursor = (int*)base; Counters* gc = (Counters*)(base + 256);
  double* d_o = (double*)(base + 4096); double* d_d = d_o + 3 * N; double* d_mn = d_d + 3 * N; double* d_mx = d_mn + N;
  rrt_hit* d_out = (rrt_hit*)(d_mx + N);
  cudaStream_t st = ctx->stream;
  /* every call shares the array and finds the last call's contents in it: the cursor and the counters must start at zero */
  CK(cudaMemsetAsync(base, 0, 512, st));
  CK(cudaMemcpyAsync(d_o, o, 24 * (size_t)n, cudaMemcpyDefault, st)); CK(cudaMemcpyAsync(d_d, d, 24 * (size_t)n, cudaMemcpyDefault, st));
  if (min_t) CK(cudaMemcpyAsync(d_mn, min_t, 8 * (size_t)n, cudaMemcpyDefault, st));
  if (max_t) CK(cudaMemcpyAsync(d_mx, max_t, 8 * (size_t)n, cudaMemcpyDefault, st));
  const double* k_mn = min_t ? d_mn : nullptr; const double* k_mx = max_t ? d_mx : nullptr;
  int grid = (int)std::min<int64_t>((n + RRT_BLOCK - 1) / RRT_BLOCK, (int64_t)ctx->sm_count * 4);
  launch_hook(ctx, any_hit != 0, grid, st, (int)n, d_o, d_d, k_mn, k_mx, d_out, d_cursor, gc);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out, d_out, sizeof(rrt_hit) * (size_t)n, cudaMemcpyDefault, st));
  CK(cudaStreamSynchronize(st));
  return RRT_OK;
}

/* ---------------------------------------------------------------- multi-GPU context (rrt_create_multi)
 * One process drives G GPUs (the reference's N worker threads pulling tiles from one queue, src/pathtracer.cpp:224-282,
 * 611-644, become G GPUs each owning the interleaved tiles (tx + ty * stride) % G == k, KeepOwned in rrt_kernels.cuh).
 *   scene   built once on GPU 0 (BVH build included), then replicated to the other GPUs by ncclBroadcast of the device arrays
 *   render  one host thread per GPU runs the same wavefront loop on its tiles: no exchange while rendering, a pixel's estimate
 *           depends only on (seed, pixel, sample)
 *   frame   ONE collective per frame: ncclReduce(sum) of the per-GPU float4 frames (mean r, g, b, sample count; pixels a GPU
 *           does not own are zero) onto GPU 0, then one kernel there splits it into the reference's sampleBuffer /
 *           sampleCountBuffer layout.
 * NCCL is loaded with dlopen at rrt_create_multi time, so single-GPU users of the library do not need it installed. */
#include <dlfcn.h>
#include <nccl.h>
#include <thread>

namespace {
struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

const char* load_nccl() {   /* nullptr = ok, else what went wrong */
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl.lib) return nullptr;
  /* An NCCL the process already holds (a Python caller that imported torch, or api.py's pre-load of the NCCL torch ships) is used as it is: the
   * loader identifies libraries by SONAME, and a second, older libnccl.so.2 loaded first would later be handed to torch in place of its own
   * (undefined symbol ncclDevCommCreate).  RRT_NCCL_LIB names a specific file.  RTLD_LOCAL: nothing of NCCL leaks into the global scope. */
  void* h = nullptr;
  if (const char* path = getenv("RRT_NCCL_LIB")) h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL | RTLD_NOLOAD);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
  if (!h) return "libnccl.so.2 could not be loaded (multi-GPU contexts need NCCL)";
  NcclApi a; a.lib = h;
#define RRT_SYM(field, name) *(void**)(&a.field) = dlsym(h, name); if (!a.field) { dlclose(h); return "libnccl.so.2 lacks " name; }
  RRT_SYM(CommInitAll, "ncclCommInitAll") RRT_SYM(CommDestroy, "ncclCommDestroy") RRT_SYM(Reduce, "ncclReduce") RRT_SYM(Broadcast, "ncclBroadcast")
  RRT_SYM(GroupStart, "ncclGroupStart") RRT_SYM(GroupEnd, "ncclGroupEnd") RRT_SYM(GetErrorString, "ncclGetErrorString")
#undef RRT_SYM
  g_nccl = a;
  return nullptr;
}
}  // namespace

#define NCK(call)                                                                                         \
  do {                                                                                                    \
    ncclResult_t r_ = (call);                                                                             \
    if (r_ != ncclSuccess) {                                                                              \
      char b_[512]; snprintf(b_, sizeof b_, "%s:%d: %s: %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r_)); \
      ctx->err = b_; return RRT_E_CUDA;                                                                   \
    }                                                                                                     \
  } while (0)

/* float4 (mean r, g, b, n) -> sampleBuffer rgb + sampleCountBuffer (either may be NULL); n is an exact small integer in float */
__global__ void __launch_bounds__(256) k_split_rgbn(size_t npix, const float4* __restrict__ rgbn, float* __restrict__ rgb, int* __restrict__ counts) {
  for (size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x; p < npix; p += (size_t)gridDim.x * blockDim.x) {
    const float4 v = rgbn[p];
    if (rgb) { rgb[3 * p] = v.x; rgb[3 * p + 1] = v.y; rgb[3 * p + 2] = v.z; }
    if (counts) counts[p] = (int)v.w;
  }
}

extern "C" int rrt_create_multi(rrt_ctx** out, int n_gpus, const int* devices) {
  if (!out) { g_create_error = "rrt_create_multi: out is NULL"; return RRT_E_ARG; }
  *out = nullptr;
  if (n_gpus <= 0 || n_gpus > 64) { g_create_error = "rrt_create_multi: n_gpus must be 1..64"; return RRT_E_ARG; }
  std::vector<int> devs(n_gpus);
  for (int k = 0; k < n_gpus; ++k) devs[k] = devices ? devices[k] : k;
  for (int k = 0; k < n_gpus; ++k)
    for (int j = 0; j < k; ++j)
      if (devs[j] == devs[k]) { g_create_error = "rrt_create_multi: a device is listed twice"; return RRT_E_ARG; }
  if (n_gpus == 1) return rrt_create(out, devs[0]);   /* nothing to shard, no NCCL needed */
  if (const char* why = load_nccl()) { g_create_error = std::string("rrt_create_multi: ") + why; return RRT_E_CUDA; }
  rrt_ctx* m = new rrt_ctx();
  memset(m->opt, 0, sizeof m->opt);
  for (int k = 0; k < n_gpus; ++k) {
    rrt_ctx* sub = nullptr;
    int rc = rrt_create(&sub, devs[k]);
    if (rc) { for (rrt_ctx* s : m->subs) rrt_destroy(s); delete m; return rc; }
    sub->tile = 8; sub->rank = k; sub->world = n_gpus;
    m->subs.push_back(sub);
  }
  m->device = devs[0];
  std::vector<ncclComm_t> comms(n_gpus);
  ncclResult_t r = g_nccl.CommInitAll(comms.data(), n_gpus, devs.data());
  if (r != ncclSuccess) {
    g_create_error = std::string("rrt_create_multi: ncclCommInitAll: ") + g_nccl.GetErrorString(r);
    for (rrt_ctx* s : m->subs) rrt_destroy(s);
    delete m; return RRT_E_CUDA;
  }
  for (ncclComm_t c : comms) m->comms.push_back((void*)c);
  *out = m;
  return RRT_OK;
}

extern "C" int rrt_gpu_count(const rrt_ctx* ctx) { return !ctx ? RRT_E_ARG : ctx->subs.empty() ? 1 : (int)ctx->subs.size(); }

static void multi_destroy(rrt_ctx* m) {
  for (size_t k = 0; k < m->comms.size(); ++k) { cudaSetDevice(m->subs[k]->device); cudaStreamSynchronize(m->subs[k]->stream); g_nccl.CommDestroy((ncclComm_t)m->comms[k]); }
  for (rrt_ctx* s : m->subs) rrt_destroy(s);
  for (cudaEvent_t ev : m->events) cudaEventDestroy(ev);
  delete m;
}

/* device pointers of the scene arrays of ctx (everything rrt_upload_scene fills) */
static void bind_scene_pointers(rrt_ctx* c) {
  DevScene& d = c->ds;
  d.nodes = c->d_nodes.as<float4>(); d.leaf = c->d_leaf.as<double2>(); d.tri_nrm = c->d_nrm.as<double>();
  d.prim_kind = c->d_kind.as<uint8_t>(); d.prim_index = c->d_index.as<uint32_t>(); d.prim_mat = c->d_pmat.as<int32_t>();
  d.sph = c->d_sph.as<double>(); d.mats = c->d_mats.as<rrt_material>(); d.lights = c->d_lights.as<rrt_light>();
}

static int multi_upload_scene(rrt_ctx* ctx, const rrt_scene* s) {
  rrt_ctx* root = ctx->subs[0];
  int rc = rrt_upload_scene(root, s);
  if (rc) { ctx->err = root->err; return rc; }
  const size_t G = ctx->subs.size();
  /* the arrays the kernels read, as rrt_upload_scene left them on GPU 0 (capacities can exceed the bytes in use: harmless) */
  DevBuf rrt_ctx::* const bufs[] = {&rrt_ctx::d_top, &rrt_ctx::d_nodes, &rrt_ctx::d_wide, &rrt_ctx::d_leaf, &rrt_ctx::d_nrm, &rrt_ctx::d_kind, &rrt_ctx::d_index, &rrt_ctx::d_pmat,
                                    &rrt_ctx::d_sph, &rrt_ctx::d_mats, &rrt_ctx::d_lights, &rrt_ctx::d_grid};
  for (size_t k = 1; k < G; ++k) {
    rrt_ctx* c = ctx->subs[k];
    CK(cudaSetDevice(c->device));
    for (auto mp : bufs) CK((c->*mp).ensure((root->*mp).bytes));
  }
  for (auto mp : bufs) {
    const size_t bytes = (root->*mp).bytes;
    if (!bytes) continue;
    NCK(g_nccl.GroupStart());
    for (size_t k = 0; k < G; ++k) {
      rrt_ctx* c = ctx->subs[k];
      NCK(g_nccl.Broadcast((root->*mp).p, (c->*mp).p, bytes, ncclUint8, 0, (ncclComm_t)ctx->comms[k], c->stream));
    }
    NCK(g_nccl.GroupEnd());
  }
  for (size_t k = 0; k < G; ++k) { CK(cudaSetDevice(ctx->subs[k]->device)); CK(cudaStreamSynchronize(ctx->subs[k]->stream)); }
  for (size_t k = 1; k < G; ++k) {
    rrt_ctx* c = ctx->subs[k];
    const DevScene keep = c->ds;           /* the environment light lives in the same struct and survives rrt_upload_scene */
    c->ds = root->ds;
    c->ds.env = keep.env; c->ds.env_pdf = keep.env_pdf; c->ds.env_cond = keep.env_cond; c->ds.env_marg = keep.env_marg; c->ds.env_w = keep.env_w; c->ds.env_h = keep.env_h;
    bind_scene_pointers(c);
    c->ds.wide = root->ds.wide ? c->d_wide.as<float>() : nullptr;
    c->ds.top = root->ds.top ? c->d_top.as<float>() : nullptr;
    c->ds.clearance = root->ds.clearance ? c->d_grid.as<float>() : nullptr;
    c->n_lights = root->n_lights; c->n_delta_lights = root->n_delta_lights; c->host_lights = root->host_lights; c->has_microfacet = root->has_microfacet;
    c->have_scene = true;
  }
  return RRT_OK;
}

static int multi_render(rrt_ctx* ctx, const rrt_params* p, float* hdr_rgb, int32_t* sample_counts, void* d_rgbn_user, void* d_counts_user, rrt_stats* stats) {
  const size_t G = ctx->subs.size();
  rrt_ctx* root = ctx->subs[0];
  const size_t npix = (size_t)root->W * root->H;
  if (!root->have_scene || !root->have_cam || root->W <= 0) FAIL(RRT_E_STATE, "rrt_render: scene, camera and frame must be set first");
  for (size_t k = 0; k < G; ++k) { rrt_ctx* c = ctx->subs[k]; CK(cudaSetDevice(c->device)); CK(c->d_out_rgbn.ensure(16 * npix)); }
  std::vector<int> rcs(G, 0);
  std::vector<rrt_stats> sts(G);
  std::vector<std::thread> th;
  for (size_t k = 0; k < G; ++k)
    th.emplace_back([&, k]() { rrt_ctx* c = ctx->subs[k]; rcs[k] = render_impl(c, p, nullptr, nullptr, c->d_out_rgbn.as<float4>(), nullptr, stats ? &sts[k] : nullptr); });
  for (auto& t : th) t.join();
  for (size_t k = 0; k < G; ++k) if (rcs[k]) { ctx->err = ctx->subs[k]->err; return rcs[k]; }
  /* the frame's one collective */
  CK(cudaSetDevice(root->device));
  ctx->events_used = 0;
  cudaEvent_t e0 = next_event(ctx), e1 = next_event(ctx);
  CK(cudaEventRecord(e0, root->stream));
  float4* dst = d_rgbn_user ? (float4*)d_rgbn_user : root->d_out_rgbn.as<float4>();
  NCK(g_nccl.GroupStart());
  for (size_t k = 0; k < G; ++k) {
    rrt_ctx* c = ctx->subs[k];
    NCK(g_nccl.Reduce(c->d_out_rgbn.p, k == 0 ? (void*)dst : c->d_out_rgbn.p, 4 * npix, ncclFloat, ncclSum, 0, (ncclComm_t)ctx->comms[k], c->stream));
  }
  NCK(g_nccl.GroupEnd());
  CK(cudaSetDevice(root->device));
  float* d_rgb = nullptr; int* d_cnt = (int*)d_counts_user;
  if (!d_rgbn_user) {
    CK(root->d_out_rgb.ensure(12 * npix)); CK(root->d_out_cnt.ensure(4 * npix));
    d_rgb = root->d_out_rgb.as<float>(); d_cnt = root->d_out_cnt.as<int>();
  }
  if (d_rgb || d_cnt) k_split_rgbn<<<root->sm_count * 8, 256, 0, root->stream>>>(npix, dst, d_rgb, d_cnt);
  CK(cudaEventRecord(e1, root->stream));
  if (hdr_rgb) CK(cudaMemcpyAsync(hdr_rgb, d_rgb, 12 * npix, cudaMemcpyDeviceToHost, root->stream));
  if (sample_counts) CK(cudaMemcpyAsync(sample_counts, d_cnt, 4 * npix, cudaMemcpyDeviceToHost, root->stream));
  for (size_t k = 0; k < G; ++k) { CK(cudaSetDevice(ctx->subs[k]->device)); CK(cudaStreamSynchronize(ctx->subs[k]->stream)); }
  CK(cudaSetDevice(root->device));
  CK(cudaGetLastError());
  if (stats) {
    memset(stats, 0, sizeof *stats);
    float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
    for (size_t k = 0; k < G; ++k) {
      const rrt_stats& a = sts[k];
      stats->seconds = std::max(stats->seconds, a.seconds); stats->trace_seconds = std::max(stats->trace_seconds, a.trace_seconds);
      stats->paths += a.paths; stats->rays += a.rays; stats->geodesic_steps += a.geodesic_steps; stats->node_visits += a.node_visits;
      stats->tri_tests += a.tri_tests; stats->sphere_tests += a.sphere_tests; stats->kernel_launches += a.kernel_launches;
      stats->trace_launches += a.trace_launches; stats->algorithmic_bytes += a.algorithmic_bytes;
    }
    stats->seconds += ms * 1e-3;      /* slowest GPU + the reduce and the split on GPU 0 */
    stats->kernel_launches += 1;
  }
  return RRT_OK;
}
