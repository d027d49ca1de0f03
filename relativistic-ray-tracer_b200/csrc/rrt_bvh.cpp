// rrt_bvh.cpp — see rrt_bvh.h.  Host code, runs once per rrt_upload_scene.
#include "rrt_bvh.h"

#include <math.h>
#include <cmath>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

namespace rrt {
namespace {

struct Box { double lo[3], hi[3]; };
struct TNode { Box bb; int l, r, start, count; };
static const int kMaxSahDepth = 38;   /* < RRT_STACK (rrt_trace.cuh), checked again at upload */
struct Tree { std::vector<TNode> nodes; std::vector<uint32_t> order; int leaves = 0, depth = 0; };

inline void box_reset(Box& b) { for (int k = 0; k < 3; ++k) { b.lo[k] = INFINITY; b.hi[k] = -INFINITY; } }
inline void box_expand(Box& b, const Box& o) {
  for (int k = 0; k < 3; ++k) { b.lo[k] = fmin(b.lo[k], o.lo[k]); b.hi[k] = fmax(b.hi[k], o.hi[k]); }
}
inline double box_area(const Box& b) {
  double x = b.hi[0] - b.lo[0], y = b.hi[1] - b.lo[1], z = b.hi[2] - b.lo[2];
  if (!(x >= 0 && y >= 0 && z >= 0)) return 0.0;
  return 2.0 * (x * y + y * z + z * x);
}

static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
#define TICK(name) do { if (getenv("RRT_BVH_TIMING")) { double t_ = now_s(); fprintf(stderr, "[bvh] %-12s %.3f s\n", name, t_ - t0_); t0_ = t_; } } while (0)
struct Builder {
  const rrt_scene* s;
  std::vector<Box> pbox;        // Primitive::get_bbox (reference src/static_scene/triangle.cpp:11-19, sphere.h:30-32)
  std::vector<double> pcen;     // 3 per prim: BBox::centroid() = (min + max) / 2, i.e. 0.5 * (min + max)  (src/bbox.h:98-100)
  std::vector<uint32_t> tmp;    // partition scratch; a call on prims[base, base + n) uses tmp[base, base + n), so disjoint
                                // subtrees can be built concurrently
  uint32_t max_leaf = 4;
  struct Task { int node; uint32_t* prims; uint32_t n, base; int depth; };
  std::vector<Task>* tasks = nullptr;
  uint32_t grain = 0;

  /* The top of the tree is built on the calling thread; subtrees of at most `grain` primitives are handed to worker
   * threads, each into its own node array, and spliced under their placeholder afterwards.  The result is the tree the
   * serial recursion builds (same splits, same primitive order); only the node array order differs, and the
   * flattening below renumbers nodes in depth-first preorder anyway. */
  template <class F>
  void build_parallel(Tree& t, F build, uint32_t* prims, uint32_t n) {
    unsigned hw = std::thread::hardware_concurrency();
    unsigned nthreads = hw < 1 ? 1 : hw > 32 ? 32 : hw;
    if (n < 65536 || nthreads == 1) { (this->*build)(t, prims, n, 0, 0); return; }
    std::vector<Task> list;
    tasks = &list; grain = std::max<uint32_t>(8192, n / (8 * nthreads));
    double t0_ = now_s();
    (this->*build)(t, prims, n, 0, 0);
    tasks = nullptr;
    TICK("  top");
    std::vector<Tree> sub(list.size());
    std::atomic<size_t> next{0};
    auto work = [&]() {
      for (size_t i; (i = next.fetch_add(1)) < list.size();) {
        const Task& k = list[i];
        sub[i].nodes.reserve(2 * (size_t)k.n / (max_leaf ? max_leaf : 1) + 16);
        (this->*build)(sub[i], k.prims, k.n, k.base, k.depth);
      }
    };
    std::vector<std::thread> pool;
    for (unsigned w = 1; w < nthreads && w < list.size(); ++w) pool.emplace_back(work);
    work();
    for (auto& th : pool) th.join();
    TICK("  subtrees");
    for (size_t i = 0; i < list.size(); ++i) {
      const std::vector<TNode>& L = sub[i].nodes;
      const int me = list[i].node, off = (int)t.nodes.size();
      auto map = [&](int j) { return j < 0 ? j : j == 0 ? me : off + j - 1; };
      for (size_t j = 0; j < L.size(); ++j) {
        TNode nd = L[j]; nd.l = map(nd.l); nd.r = map(nd.r);
        if (j == 0) t.nodes[me] = nd; else t.nodes.push_back(nd);
      }
      t.leaves += sub[i].leaves;
      if (sub[i].depth > t.depth) t.depth = sub[i].depth;
    }
  }

  void prepare() {
    uint32_t n = s->n_prims;
    pbox.resize(n); pcen.resize(3 * (size_t)n); tmp.resize(n);
    for (uint32_t p = 0; p < n; ++p) {
      Box b;
      if (s->prim_kind[p] == 0) {
        const double* q = s->tri_pos + 9 * (size_t)s->prim_index[p];
        for (int k = 0; k < 3; ++k) b.lo[k] = b.hi[k] = q[k];
        for (int v = 1; v < 3; ++v)
          for (int k = 0; k < 3; ++k) { b.lo[k] = fmin(b.lo[k], q[3 * v + k]); b.hi[k] = fmax(b.hi[k], q[3 * v + k]); }
      } else {
        const double* c = s->sph + 4 * (size_t)s->prim_index[p];
        for (int k = 0; k < 3; ++k) { b.lo[k] = c[k] - c[3]; b.hi[k] = c[k] + c[3]; }
      }
      pbox[p] = b;
      const double half = 1.0 / 2;
      for (int k = 0; k < 3; ++k) pcen[3 * (size_t)p + k] = half * (b.lo[k] + b.hi[k]);
    }
  }

  Box bounds(const uint32_t* prims, uint32_t n) const {
    Box b; box_reset(b);
    for (uint32_t i = 0; i < n; ++i) box_expand(b, pbox[prims[i]]);
    return b;
  }

  // construct_bvh (reference src/bvh.cpp:49-96).  prims[0..n) is reordered in place (stable partition, as the
  // reference's push_back into left/right vectors is).  The reference recurses; this walks the same tree with an explicit stack, nodes numbered
  // in the same depth-first preorder: the midpoint split has no depth bound (boxes that all contain one point give a chain as deep as the scene
  // is large — 50 000 such triangles overflowed the thread's stack here, as they would the reference's), and a scene is input, not code.
  int build_ref(Tree& t, uint32_t* prims0, uint32_t n0, uint32_t base0, int depth0) {
    struct Frame { int parent; bool right; uint32_t* prims; uint32_t n, base; int depth; };
    std::vector<Frame> todo;
    todo.push_back(Frame{-1, false, prims0, n0, base0, depth0});
    const int first = (int)t.nodes.size();
    while (!todo.empty()) {
      const Frame f = todo.back(); todo.pop_back();
      uint32_t* prims = f.prims; const uint32_t n = f.n, base = f.base; const int depth = f.depth;
      const int me = (int)t.nodes.size();
      t.nodes.push_back(TNode());
      if (f.parent >= 0) { if (f.right) t.nodes[f.parent].r = me; else t.nodes[f.parent].l = me; }
      t.nodes[me].l = t.nodes[me].r = -1; t.nodes[me].start = t.nodes[me].count = 0;
      if (tasks && n <= grain && depth > 0) {   /* subtree left to a worker thread (build_parallel) */
        tasks->push_back(Task{me, prims, n, base, depth});
        continue;
      }
      Box bb = bounds(prims, n);
      t.nodes[me].bb = bb;
      if (depth > t.depth) t.depth = depth;
      if (n <= 4) {
        t.nodes[me].start = (int)base; t.nodes[me].count = (int)n; t.leaves++;
        continue;
      }
      double ex = bb.hi[0] - bb.lo[0], ey = bb.hi[1] - bb.lo[1], ez = bb.hi[2] - bb.lo[2];
      int axis = (ex > ey && ex > ez) ? 0 : (ey > ex && ey > ez) ? 1 : 2;
      const double half = 1.0 / 2;
      double mid = half * (bb.lo[axis] + bb.hi[axis]);
      uint32_t nl = 0, nr = 0;
      for (uint32_t i = 0; i < n; ++i) {
        if (pcen[3 * (size_t)prims[i] + axis] < mid) prims[nl++] = prims[i];   // nl <= i: safe in place
        else tmp[base + nr++] = prims[i];
      }
      if (nl == 0 || nr == 0) {
        // both vectors are rebuilt from the ORIGINAL order: first half left, rest right; the in-place pass above kept
        // the original order in whichever side received everything
        if (nl == 0) memcpy(prims, tmp.data() + base, sizeof(uint32_t) * n);
        nl = n / 2; nr = n - nl;
      } else {
        memcpy(prims + nl, tmp.data() + base, sizeof(uint32_t) * nr);
      }
      todo.push_back(Frame{me, true, prims + nl, nr, base + nl, depth + 1});   // popped second: the left subtree is numbered first
      todo.push_back(Frame{me, false, prims, nl, base, depth + 1});
    }
    return first;
  }

  // binned SAH (16 bins per axis); leaves of <= 4 primitives like the reference
  int build_sah(Tree& t, uint32_t* prims, uint32_t n, uint32_t base, int depth) {
    int me = (int)t.nodes.size();
    t.nodes.push_back(TNode());
    if (tasks && n <= grain && depth > 0) {   /* subtree left to a worker thread (build_parallel) */
      t.nodes[me].l = t.nodes[me].r = -1; t.nodes[me].start = t.nodes[me].count = 0;
      tasks->push_back(Task{me, prims, n, base, depth});
      return me;
    }
    Box bb = bounds(prims, n);
    t.nodes[me].bb = bb; t.nodes[me].l = t.nodes[me].r = -1; t.nodes[me].start = t.nodes[me].count = 0;
    if (depth > t.depth) t.depth = depth;
    if (n <= max_leaf) {
      t.nodes[me].start = (int)base; t.nodes[me].count = (int)n; t.leaves++;
      return me;
    }
    const int NB = 16;
    double clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint32_t i = 0; i < n; ++i)
      for (int k = 0; k < 3; ++k) { double c = pcen[3 * (size_t)prims[i] + k]; clo[k] = fmin(clo[k], c); chi[k] = fmax(chi[k], c); }
    {
      /* depth budget: the device traversal stack holds RRT_STACK (40) entries, one per level.  Once the levels a
       * balanced subtree of n primitives still needs would no longer fit under kMaxSahDepth, split at the object median
       * of the widest axis from here down, which bounds the depth by construction for any n < 2^28. */
      int need = 0;
      for (uint64_t cap = max_leaf ? max_leaf : 1; cap < n; cap <<= 1) ++need;
      if (depth + need + 1 >= kMaxSahDepth) {
        int axis = 0;
        for (int k = 1; k < 3; ++k) if (chi[k] - clo[k] > chi[axis] - clo[axis]) axis = k;
        uint32_t h = n / 2;
        const double* pc = pcen.data();
        std::nth_element(prims, prims + h, prims + n, [pc, axis](uint32_t a, uint32_t b) {
          double ca = pc[3 * (size_t)a + axis], cb = pc[3 * (size_t)b + axis];
          return ca < cb || (ca == cb && a < b);
        });
        int l = build_sah(t, prims, h, base, depth + 1);
        int r = build_sah(t, prims + h, n - h, base + h, depth + 1);
        t.nodes[me].l = l; t.nodes[me].r = r;
        return me;
      }
    }
    double best = INFINITY; int best_axis = -1, best_bin = -1;
    /* one pass over the primitives fills the bins of all three axes; large nodes (the serial top of the tree) split
     * the pass over threads and merge the partial bins, which gives the same boxes and counts in any order */
    struct Bins { Box bb[3][NB]; uint32_t cnt[3][NB]; };
    double scale3[3]; bool use[3];
    for (int k = 0; k < 3; ++k) { double w = chi[k] - clo[k]; use[k] = w > 0; scale3[k] = use[k] ? NB / w : 0.0; }
    auto fill = [&](Bins& B, uint32_t i0, uint32_t i1) {
      for (int k = 0; k < 3; ++k) for (int q = 0; q < NB; ++q) { box_reset(B.bb[k][q]); B.cnt[k][q] = 0; }
      for (uint32_t i = i0; i < i1; ++i) {
        const uint32_t p = prims[i];
        const Box& pb = pbox[p];
        for (int k = 0; k < 3; ++k) {
          if (!use[k]) continue;
          int q = (int)((pcen[3 * (size_t)p + k] - clo[k]) * scale3[k]);
          if (q >= NB) q = NB - 1;
          if (q < 0) q = 0;
          box_expand(B.bb[k][q], pb); B.cnt[k][q]++;
        }
      }
    };
    Bins bins;
    unsigned parts = n >= 65536 ? std::min(16u, std::max(1u, std::thread::hardware_concurrency())) : 1;
    if (parts > 1) {
      std::vector<Bins> part(parts);
      std::vector<std::thread> th;
      for (unsigned w = 0; w < parts; ++w)
        th.emplace_back([&, w]() { fill(part[w], (uint32_t)((uint64_t)n * w / parts), (uint32_t)((uint64_t)n * (w + 1) / parts)); });
      for (auto& x : th) x.join();
      bins = part[0];
      for (unsigned w = 1; w < parts; ++w)
        for (int k = 0; k < 3; ++k) for (int q = 0; q < NB; ++q) { box_expand(bins.bb[k][q], part[w].bb[k][q]); bins.cnt[k][q] += part[w].cnt[k][q]; }
    } else {
      fill(bins, 0, n);
    }
    for (int axis = 0; axis < 3; ++axis) {
      if (!use[axis]) continue;
      const Box* bbx = bins.bb[axis]; const uint32_t* cnt = bins.cnt[axis];
      double ra[NB]; uint32_t rc[NB]; Box acc; box_reset(acc); uint32_t c = 0;
      for (int b = NB - 1; b > 0; --b) { box_expand(acc, bbx[b]); c += cnt[b]; ra[b] = box_area(acc); rc[b] = c; }
      box_reset(acc); c = 0;
      for (int b = 0; b < NB - 1; ++b) {
        box_expand(acc, bbx[b]); c += cnt[b];
        if (c == 0 || rc[b + 1] == 0) continue;
        double cost = box_area(acc) * c + ra[b + 1] * rc[b + 1];
        if (cost < best) { best = cost; best_axis = axis; best_bin = b; }
      }
    }
    uint32_t nl = 0, nr = 0;
    if (best_axis >= 0) {
      double w = chi[best_axis] - clo[best_axis], scale = NB / w;
      for (uint32_t i = 0; i < n; ++i) {
        int b = (int)((pcen[3 * (size_t)prims[i] + best_axis] - clo[best_axis]) * scale);
        if (b >= NB) b = NB - 1;
        if (b < 0) b = 0;
        if (b <= best_bin) prims[nl++] = prims[i]; else tmp[base + nr++] = prims[i];
      }
      memcpy(prims + nl, tmp.data() + base, sizeof(uint32_t) * nr);
    }
    if (nl == 0 || nr == 0) { nl = n / 2; nr = n - nl; }
    int l = build_sah(t, prims, nl, base, depth + 1);
    int r = build_sah(t, prims + nl, nr, base + nl, depth + 1);
    t.nodes[me].l = l; t.nodes[me].r = r;
    return me;
  }
};

inline float round_down(double x) { float f = (float)x; if ((double)f > x) f = nextafterf(f, -INFINITY); return f; }
inline float round_up(double x) { float f = (float)x; if ((double)f < x) f = nextafterf(f, INFINITY); return f; }

inline int32_t as_i32(float f) { int32_t i; memcpy(&i, &f, 4); return i; }
inline float as_f32(int32_t i) { float f; memcpy(&f, &i, 4); return f; }

}  // namespace

float round_down_f(double x) { return round_down(x); }
float round_up_f(double x) { return round_up(x); }

int build_bvh(const rrt_scene* s, int topology, int max_leaf, FlatBvh* out, int width) {
  double t0_ = now_s();
  *out = FlatBvh();
  uint32_t n = s->n_prims;
  out->dfs_rank.assign(n, 0);
  out->prim_mat.resize(n);
  for (uint32_t p = 0; p < n; ++p) out->prim_mat[p] = s->prim_kind[p] == 0 ? s->tri_mat[s->prim_index[p]] : s->sph_mat[s->prim_index[p]];
  if (n == 0) { out->root = -1; /* ~(-1) = 0: an empty leaf */ return 0; }
  Builder b; b.s = s; b.prepare();
  TICK("prepare");
  b.max_leaf = (uint32_t)(max_leaf < 1 ? 1 : max_leaf > 7 ? 7 : max_leaf);

  Tree ref; ref.order.resize(n);
  for (uint32_t i = 0; i < n; ++i) ref.order[i] = i;
  ref.nodes.reserve(n);
  b.build_parallel(ref, &Builder::build_ref, ref.order.data(), n);
  for (uint32_t i = 0; i < n; ++i) out->dfs_rank[ref.order[i]] = i;
  TICK("ref tree");
  out->ref_nodes = (int)ref.nodes.size(); out->ref_leaves = ref.leaves; out->ref_depth = ref.depth;

  Tree own;
  const Tree* t = &ref;
  if (topology == 1) {
    own.order.resize(n);
    for (uint32_t i = 0; i < n; ++i) own.order[i] = i;
    own.nodes.reserve(n);
    b.build_parallel(own, &Builder::build_sah, own.order.data(), n);
    t = &own;
  }

  TICK("sah tree");
  // leaf records in the tree's leaf order
  out->leaf.assign(10 * (size_t)n, 0.0);
  for (uint32_t i = 0; i < n; ++i) {
    uint32_t p = t->order[i];
    double* r = &out->leaf[10 * (size_t)i];
    uint32_t kind = s->prim_kind[p];
    if (kind == 0) {
      const double* q = s->tri_pos + 9 * (size_t)s->prim_index[p];
      for (int k = 0; k < 3; ++k) { r[k] = q[k]; r[3 + k] = q[3 + k] - q[k]; r[6 + k] = q[6 + k] - q[k]; }
    } else {
      const double* c = s->sph + 4 * (size_t)s->prim_index[p];
      r[0] = c[0]; r[1] = c[1]; r[2] = c[2]; r[3] = c[3] * c[3]; /* Sphere::r2 = r * r (reference src/static_scene/sphere.h:27) */
    }
    uint32_t w[2] = {p, (out->dfs_rank[p] << 1) | kind};
    memcpy(&r[9], w, 8);
  }

  TICK("leaf records");
  // flatten: interior nodes in depth-first preorder
  auto child_code = [&](int ni, std::vector<int>& remap) -> int32_t {
    const TNode& c = t->nodes[ni];
    if (c.l < 0) return ~(int32_t)(((uint32_t)c.start << 3) | (uint32_t)c.count);
    return remap[ni];
  };
  std::vector<int> remap(t->nodes.size(), -1);
  int ninner = 0;
  {  // interior nodes numbered in depth-first preorder (left subtree first): parents sit next to their near descendants
    std::vector<int> st; st.push_back(0);
    while (!st.empty()) {
      int ni = st.back(); st.pop_back();
      if (t->nodes[ni].l < 0) continue;
      remap[ni] = ninner++;
      st.push_back(t->nodes[ni].r); st.push_back(t->nodes[ni].l);
    }
  }
  out->n_nodes = ninner;
  out->nodes.assign(16 * (size_t)(ninner > 0 ? ninner : 1), 0.f);
  for (size_t i = 0; i < t->nodes.size(); ++i) {
    const TNode& nd = t->nodes[i];
    if (nd.l < 0) continue;
    float* q = &out->nodes[16 * (size_t)remap[i]];
    const Box& L = t->nodes[nd.l].bb; const Box& R = t->nodes[nd.r].bb;
    q[0] = round_down(L.lo[0]); q[1] = round_down(L.lo[1]); q[2] = round_down(L.lo[2]); q[3] = round_up(L.hi[0]);
    q[4] = round_up(L.hi[1]); q[5] = round_up(L.hi[2]); q[6] = round_down(R.lo[0]); q[7] = round_down(R.lo[1]);
    q[8] = round_down(R.lo[2]); q[9] = round_up(R.hi[0]); q[10] = round_up(R.hi[1]); q[11] = round_up(R.hi[2]);
    q[12] = as_f32(child_code(nd.l, remap)); q[13] = as_f32(child_code(nd.r, remap)); q[14] = 0.f; q[15] = 0.f;
  }
  out->root = child_code(0, remap);
  for (int k = 0; k < 3; ++k) { out->root_lo[k] = t->nodes[0].bb.lo[k]; out->root_hi[k] = t->nodes[0].bb.hi[k]; }
  (void)as_i32;
  TICK("flatten");

  // ---- wide form: every interior node adopts grandchildren (largest box first) until it has `width` children.
  // Fewer, fatter nodes: a traversal makes fewer loop iterations with the same instruction stream in every lane, which
  // is what a warp that walks 32 different chords in lockstep wants.  Leaves are untouched (same records, same order).
  out->width = (width == 4 || width == 8) ? width : 2;
  if (out->width > 2) {
    const int W = out->width;
    struct Pending { int wide_index; int bin_node; int depth; };
    std::vector<Pending> todo;
    auto leaf_code = [&](const TNode& c) { return ~(int32_t)(((uint32_t)c.start << 3) | (uint32_t)c.count); };
    if (t->nodes[0].l < 0) { out->wide_root = leaf_code(t->nodes[0]); out->n_wide = 0; out->wide.assign(8 * (size_t)W, 0.f); }
    else {
      out->wide_root = 0; out->n_wide = 1;
      todo.push_back({0, 0, 1});
      size_t head = 0;
      std::vector<std::vector<int>> kids;   // per wide node: binary-node indices of its children
      kids.push_back({});
      // breadth-first numbering keeps the top of the tree (visited by every chord) in a few cache lines
      while (head < todo.size()) {
        Pending cur = todo[head++];
        if (cur.depth > out->wide_depth) out->wide_depth = cur.depth;
        std::vector<int> ch = {t->nodes[cur.bin_node].l, t->nodes[cur.bin_node].r};
        while ((int)ch.size() < W) {
          int pick = -1; double best_area = -1.0;
          for (size_t k = 0; k < ch.size(); ++k) {
            const TNode& c = t->nodes[ch[k]];
            if (c.l < 0) continue;
            double a = box_area(c.bb);
            if (a > best_area) { best_area = a; pick = (int)k; }
          }
          if (pick < 0) break;
          int b = ch[pick];
          ch[pick] = t->nodes[b].l;
          ch.push_back(t->nodes[b].r);
        }
        kids[cur.wide_index] = ch;
        for (int c : ch)
          if (t->nodes[c].l >= 0) { todo.push_back({out->n_wide++, c, cur.depth + 1}); kids.push_back({}); }
      }
      out->wide.assign(8 * (size_t)W * out->n_wide, 0.f);
      // second pass: child codes need the numbering above
      std::vector<int> wide_of(t->nodes.size(), -1);
      for (const Pending& pnd : todo) wide_of[pnd.bin_node] = pnd.wide_index;
      for (const Pending& pnd : todo) {
        float* q = &out->wide[8 * (size_t)W * pnd.wide_index];
        const std::vector<int>& ch = kids[pnd.wide_index];
        for (int k = 0; k < W; ++k) {
          if (k < (int)ch.size()) {
            const TNode& c = t->nodes[ch[k]];
            q[0 * W + k] = round_down(c.bb.lo[0]); q[1 * W + k] = round_down(c.bb.lo[1]); q[2 * W + k] = round_down(c.bb.lo[2]);
            q[3 * W + k] = round_up(c.bb.hi[0]); q[4 * W + k] = round_up(c.bb.hi[1]); q[5 * W + k] = round_up(c.bb.hi[2]);
            q[6 * W + k] = as_f32(c.l < 0 ? leaf_code(c) : wide_of[ch[k]]);
          } else {
            q[0 * W + k] = q[1 * W + k] = q[2 * W + k] = INFINITY;
            q[3 * W + k] = q[4 * W + k] = q[5 * W + k] = -INFINITY;
            q[6 * W + k] = as_f32(kEmptyChild);
          }
          q[7 * W + k] = 0.f;
        }
      }
    }
    TICK("collapse");
  }
  return 0;
}

}  // namespace rrt

namespace rrt {

// Clearance grid.  Cells overlapped by a primitive's bounding box are "occupied".  For every cell c the value stored is
//   clearance(c) = max(0, min over occupied c' of |centre(c) - centre(c')| - |cell diagonal|) * (1 - 1e-4) - 1e-4 * min cell side
// Any point p of cell c and any point q of a primitive (which lies inside some occupied c') satisfy
//   |p - q| >= |centre(c) - centre(c')| - half diagonal - half diagonal,
// so a segment of length < clearance(c) that starts in c (or within 1e-4 cell sides of it: the device locates the cell in
// float) cannot reach a primitive.  The minimum over occupied cells is an exact anisotropic distance transform done as
// three 1-D min-plus passes.
void build_clearance_grid(const rrt_scene* s, int res, FlatBvh* out) {
  out->clearance.clear(); out->grid_n[0] = out->grid_n[1] = out->grid_n[2] = 0;
  if (res <= 0 || s->n_prims == 0) return;
  double lo[3], hi[3], ext[3], emax = 0;
  for (int k = 0; k < 3; ++k) { lo[k] = out->root_lo[k]; hi[k] = out->root_hi[k]; }
  for (int k = 0; k < 3; ++k) {
    double pad = 1e-3 * fmax(hi[k] - lo[k], 1e-6) + 1e-9;   // keeps primitives on the box faces strictly inside
    lo[k] -= pad; hi[k] += pad; ext[k] = hi[k] - lo[k]; emax = fmax(emax, ext[k]);
  }
  if (!(emax > 0) || !std::isfinite(emax)) return;
  int n[3]; double h[3];
  for (int k = 0; k < 3; ++k) {
    n[k] = (int)ceil(res * ext[k] / emax); if (n[k] < 1) n[k] = 1; if (n[k] > res) n[k] = res;
    h[k] = ext[k] / n[k];
  }
  const size_t N = (size_t)n[0] * n[1] * n[2];
  const double BIG = 1e300;
  std::vector<double> d2(N, BIG);   // squared centre distance to the nearest occupied cell
  auto cell = [&](double x, int k) { int c = (int)floor((x - lo[k]) / h[k]); return c < 0 ? 0 : c >= n[k] ? n[k] - 1 : c; };
  for (uint32_t p = 0; p < s->n_prims; ++p) {
    double blo[3], bhi[3];
    if (s->prim_kind[p] == 0) {
      const double* q = s->tri_pos + 9 * (size_t)s->prim_index[p];
      for (int k = 0; k < 3; ++k) { blo[k] = fmin(q[k], fmin(q[3 + k], q[6 + k])); bhi[k] = fmax(q[k], fmax(q[3 + k], q[6 + k])); }
    } else {
      const double* c = s->sph + 4 * (size_t)s->prim_index[p];
      for (int k = 0; k < 3; ++k) { blo[k] = c[k] - fabs(c[3]); bhi[k] = c[k] + fabs(c[3]); }
    }
    int c0[3], c1[3];
    for (int k = 0; k < 3; ++k) {   // one extra cell of margin on each side absorbs the floor() rounding at cell faces
      c0[k] = cell(blo[k], k) - 1; c1[k] = cell(bhi[k], k) + 1;
      if (c0[k] < 0) c0[k] = 0;
      if (c1[k] >= n[k]) c1[k] = n[k] - 1;
    }
    for (int z = c0[2]; z <= c1[2]; ++z)
      for (int y = c0[1]; y <= c1[1]; ++y)
        for (int x = c0[0]; x <= c1[0]; ++x) d2[((size_t)z * n[1] + y) * n[0] + x] = 0.0;
  }
  // separable min-plus passes: d2(i) = min_j d2(j) + ((i - j) h)^2
  std::vector<double> line, res_line;
  for (int axis = 0; axis < 3; ++axis) {
    const int L = n[axis];
    const size_t stride = axis == 0 ? 1 : axis == 1 ? (size_t)n[0] : (size_t)n[0] * n[1];
    line.assign(L, 0.0); res_line.assign(L, 0.0);
    const int A = n[(axis + 1) % 3], B = n[(axis + 2) % 3];
    const size_t sa = (axis + 1) % 3 == 0 ? 1 : (axis + 1) % 3 == 1 ? (size_t)n[0] : (size_t)n[0] * n[1];
    const size_t sb = (axis + 2) % 3 == 0 ? 1 : (axis + 2) % 3 == 1 ? (size_t)n[0] : (size_t)n[0] * n[1];
    for (int a = 0; a < A; ++a)
      for (int b = 0; b < B; ++b) {
        const size_t base = a * sa + b * sb;
        for (int i = 0; i < L; ++i) line[i] = d2[base + i * stride];
        for (int i = 0; i < L; ++i) {
          double best = BIG;
          for (int j = 0; j < L; ++j) {
            if (line[j] >= BIG) continue;
            double dd = (i - j) * h[axis];
            double v = line[j] + dd * dd;
            if (v < best) best = v;
          }
          res_line[i] = best;
        }
        for (int i = 0; i < L; ++i) d2[base + i * stride] = res_line[i];
      }
  }
  const double diag = sqrt(h[0] * h[0] + h[1] * h[1] + h[2] * h[2]);
  const double hmin = fmin(h[0], fmin(h[1], h[2]));
  out->clearance.resize(N);
  for (size_t i = 0; i < N; ++i) {
    double c = d2[i] >= BIG ? 0.0 : (sqrt(d2[i]) - diag) * (1.0 - 1e-4) - 1e-4 * hmin;
    out->clearance[i] = c > 0 ? round_down((float)c > c ? c * (1 - 1e-6) : c) : 0.f;
  }
  for (int k = 0; k < 3; ++k) { out->grid_n[k] = n[k]; out->grid_lo[k] = lo[k]; out->grid_inv_h[k] = 1.0 / h[k]; }
}

}  // namespace rrt
