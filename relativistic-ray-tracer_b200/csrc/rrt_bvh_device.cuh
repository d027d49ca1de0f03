// rrt_bvh_device.cuh — BVH construction ON THE DEVICE (RRT_OPT_BVH_BUILDER).
//
// Replaces, for large scenes, the host build of rrt_bvh.cpp (BVHAccel::BVHAccel / construct_bvh, reference
// src/bvh.cpp:14-19,49-96; 1 M triangles: 1.5-3 s on the host cores) with two device passes over the primitives:
//
//  1. the reference's OWN tree shape, needed only for its depth-first primitive order (equal-t ties go to the primitive the
//     reference visits last, rrt_bvh.h): construct_bvh is a recursive STABLE partition of the primitive list by
//     `centroid[axis] < midpoint of the node box`, with index halving when a side stays empty and leaves of <= 4.  Here it
//     runs level by level over all segments at once: per level one pass of ordered-integer atomicMin/Max for the segment
//     boxes (doubles; min / max are exact, so the order of the atomics is irrelevant), one flag pass, one exclusive scan
//     (cub) and one scatter; the final position of a primitive IS its depth-first rank.  Same doubles, same comparisons as
//     the host restatement => ranks identical bit for bit (tests/test_gpu_parity.py::test_device_bvh_build_*).
//  2. the tree the kernels walk: an LBVH (Karras 2012) over 30-bit Morton codes of the box centroids, leaves of one
//     primitive, boxes refit bottom-up from the primitives' double boxes rounded OUTWARD to float, written straight into the
//     64-byte node records and 80-byte leaf records of rrt_bvh.h.  Primitives whose box spans more than 1/8 of the scene on
//     some axis (the Cornell box's walls) get a key bit of their own, so they form a subtree beside the root instead of
//     inflating the boxes of every Morton cell they happen to fall in.
//
// Any valid BVH gives the same results (the exact leaf tests and the rank tie-break decide), so the LBVH changes speed, not
// images.  Measured on B200 it is walked 0.5-3 % more slowly than the host's binned-SAH tree and is ready ~35x sooner (C5, 1 M
// triangles: rrt_upload_scene 942 -> 27 ms), so it is the default from 4096 primitives up (RRT_OPT_BVH_BUILDER).
#pragma once
#include <cub/cub.cuh>
#include "rrt_trace.cuh"

namespace rrtdev {

__device__ __forceinline__ unsigned long long enc_f64(double x) {   /* order-preserving map double -> uint64 */
  unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double dec_f64(unsigned long long e) {
  unsigned long long b = (e >> 63) ? (e & 0x7fffffffffffffffull) : ~e;
  return __longlong_as_double((long long)b);
}

/* Primitive::get_bbox (reference src/static_scene/triangle.cpp:11-19, sphere.h:30-32) and BBox::centroid (src/bbox.h:98-100) */
__global__ void k_prim_bounds(uint32_t n, const uint8_t* __restrict__ kind, const uint32_t* __restrict__ index, const double* __restrict__ tri_pos,
                              const double* __restrict__ sph, double* __restrict__ plo, double* __restrict__ phi, double* __restrict__ cen) {
  uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  double lo[3], hi[3];
  if (kind[p] == 0) {
    const double* q = tri_pos + 9 * (size_t)index[p];
#pragma unroll
    for (int k = 0; k < 3; ++k) { lo[k] = fmin(q[k], fmin(q[3 + k], q[6 + k])); hi[k] = fmax(q[k], fmax(q[3 + k], q[6 + k])); }
  } else {
    const double* c = sph + 4 * (size_t)index[p];
#pragma unroll
    for (int k = 0; k < 3; ++k) { lo[k] = c[k] - c[3]; hi[k] = c[k] + c[3]; }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    plo[(size_t)k * n + p] = lo[k]; phi[(size_t)k * n + p] = hi[k];
    cen[(size_t)k * n + p] = 0.5 * (lo[k] + hi[k]);
  }
}

/* ---------------------------------------------------------------- 1. reference depth-first order
 * position i of `order` lies in the segment [sstart[i], send[i]); a segment of more than 4 primitives is still to be split.
 * Per-segment data lives at the index of the segment's first position. */
__global__ void k_ref_init(uint32_t n, uint32_t* order, uint32_t* sstart, uint32_t* send) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { order[i] = i; sstart[i] = 0; send[i] = n; }
}
__global__ void k_ref_reset(uint32_t n, const uint32_t* __restrict__ sstart, const uint32_t* __restrict__ send, unsigned long long* __restrict__ segbox) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || sstart[i] != i || send[i] - i <= 4) return;
#pragma unroll
  for (int k = 0; k < 3; ++k) { segbox[(size_t)k * n + i] = ~0ull; segbox[(size_t)(3 + k) * n + i] = 0ull; }
}
__global__ void k_ref_accum(uint32_t n, const uint32_t* __restrict__ order, const uint32_t* __restrict__ sstart, const uint32_t* __restrict__ send,
                            const double* __restrict__ plo, const double* __restrict__ phi, unsigned long long* __restrict__ segbox) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool act = i < n && send[i] - sstart[i] > 4;
  const uint32_t s = act ? sstart[i] : 0xffffffffu;
  unsigned long long v[6];
  if (act) {
    const uint32_t p = order[i];
#pragma unroll
    for (int k = 0; k < 3; ++k) { v[k] = enc_f64(plo[(size_t)k * n + p]); v[3 + k] = enc_f64(phi[(size_t)k * n + p]); }
  } else {
#pragma unroll
    for (int k = 0; k < 3; ++k) { v[k] = ~0ull; v[3 + k] = 0ull; }
  }
  /* a warp that lies inside one segment (every warp of the upper levels) reduces first: one set of atomics per warp */
  const uint32_t s0 = __shfl_sync(0xffffffffu, s, 0);
  if (__all_sync(0xffffffffu, s == s0)) {
    if (s0 == 0xffffffffu) return;
#pragma unroll
    for (int k = 0; k < 6; ++k)
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        unsigned long long w = __shfl_xor_sync(0xffffffffu, v[k], o);
        v[k] = k < 3 ? (w < v[k] ? w : v[k]) : (w > v[k] ? w : v[k]);
      }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
      for (int k = 0; k < 3; ++k) { atomicMin(segbox + (size_t)k * n + s0, v[k]); atomicMax(segbox + (size_t)(3 + k) * n + s0, v[3 + k]); }
    }
  } else if (act) {
#pragma unroll
    for (int k = 0; k < 3; ++k) { atomicMin(segbox + (size_t)k * n + s, v[k]); atomicMax(segbox + (size_t)(3 + k) * n + s, v[3 + k]); }
  }
}
/* the split rule of construct_bvh (reference src/bvh.cpp:66-84): strictly largest extent picks the axis (else z), the
 * midpoint of the node box is the threshold, `centroid < midpoint` goes left */
__global__ void k_ref_flag(uint32_t n, const uint32_t* __restrict__ order, const uint32_t* __restrict__ sstart, const uint32_t* __restrict__ send,
                           const unsigned long long* __restrict__ segbox, const double* __restrict__ cen, uint32_t* __restrict__ is_left) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > n) return;
  if (i == n) { is_left[n] = 0; return; }
  const uint32_t s = sstart[i];
  if (send[i] - s <= 4) { is_left[i] = 0; return; }
  double lo[3], hi[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) { lo[k] = dec_f64(segbox[(size_t)k * n + s]); hi[k] = dec_f64(segbox[(size_t)(3 + k) * n + s]); }
  const double ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
  const int axis = (ex > ey && ex > ez) ? 0 : (ey > ex && ey > ez) ? 1 : 2;
  const double mid = 0.5 * (lo[axis] + hi[axis]);
  is_left[i] = cen[(size_t)axis * n + order[i]] < mid ? 1u : 0u;
}
__global__ void k_ref_scatter(uint32_t n, const uint32_t* __restrict__ order, const uint32_t* __restrict__ sstart, const uint32_t* __restrict__ send,
                              const uint32_t* __restrict__ is_left, const uint32_t* __restrict__ pre, uint32_t* __restrict__ order_o,
                              uint32_t* __restrict__ sstart_o, uint32_t* __restrict__ send_o, int* __restrict__ any_active, int* __restrict__ depth_stats) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t s = sstart[i], e = send[i], cnt = e - s;
  if (cnt <= 4) { order_o[i] = order[i]; sstart_o[i] = s; send_o[i] = e; return; }
  uint32_t nl = pre[e] - pre[s], pos;
  if (nl == 0 || nl == cnt) { pos = i; nl = cnt / 2; }           /* a side stayed empty: original order, first half left */
  else {
    const uint32_t L = pre[i] - pre[s];                           /* lefts before me in my segment */
    pos = is_left[i] ? s + L : s + nl + ((i - s) - L);
  }
  const bool left = pos < s + nl;
  const uint32_t ns = left ? s : s + nl, ne = left ? s + nl : e;
  order_o[pos] = order[i]; sstart_o[pos] = ns; send_o[pos] = ne;
  if (pos == ns) {                                                /* first position of a new segment: one node of the reference tree */
    if (ne - ns > 4) *any_active = 1;
    atomicAdd(depth_stats, 1);
    if (ne - ns <= 4) atomicAdd(depth_stats + 1, 1);              /* ... a leaf */
  }
}
__global__ void k_ref_rank(uint32_t n, const uint32_t* __restrict__ order, uint32_t* __restrict__ rank) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) rank[order[i]] = i;
}

/* ---------------------------------------------------------------- 2. LBVH */
__device__ __forceinline__ uint32_t spread10(uint32_t v) {   /* 10 bits -> every third bit */
  v &= 1023u;
  v = (v | (v << 16)) & 0x030000FFu; v = (v | (v << 8)) & 0x0300F00Fu; v = (v | (v << 4)) & 0x030C30C3u; v = (v | (v << 2)) & 0x09249249u;
  return v;
}
struct Box6 { double lo[3], hi[3]; };
__global__ void k_morton(uint32_t n, const __grid_constant__ Box6 sb, const double* __restrict__ plo, const double* __restrict__ phi, const double* __restrict__ cen,
                         unsigned long long* __restrict__ key, uint32_t* __restrict__ val) {
  uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  uint32_t q[3]; bool large = false;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const double ext = sb.hi[k] - sb.lo[k];
    double t = ext > 0 ? (cen[(size_t)k * n + p] - sb.lo[k]) / ext : 0.0;
    t = t < 0 ? 0 : t > 1 ? 1 : t;
    if (!(t == t)) t = 0;
    q[k] = min(1023u, (uint32_t)(t * 1024.0));
    if ((phi[(size_t)k * n + p] - plo[(size_t)k * n + p]) > 0.125 * ext) large = true;
  }
  const unsigned long long m = ((unsigned long long)spread10(q[0]) << 2) | ((unsigned long long)spread10(q[1]) << 1) | (unsigned long long)spread10(q[2]);
  key[p] = ((unsigned long long)(large ? 1 : 0) << 62) | (m << 32) | (unsigned long long)p;   /* low word makes keys unique */
  val[p] = p;
}
__device__ __forceinline__ int lbvh_delta(const unsigned long long* __restrict__ key, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  return __clzll((long long)(key[i] ^ key[j]));
}
/* Karras, "Maximizing Parallelism in the Construction of BVHs, Octrees, and k-d Trees" (HPG 2012), algorithm of section 4;
 * internal node i in [0, n - 2], leaves are sorted positions; child code >= 0 internal, < 0 leaf ~position */
__global__ void k_lbvh_topology(int n, const unsigned long long* __restrict__ key, int* __restrict__ child, int* __restrict__ parent_int, int* __restrict__ parent_leaf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const int d = lbvh_delta(key, n, i, i + 1) - lbvh_delta(key, n, i, i - 1) >= 0 ? 1 : -1;
  const int dmin = lbvh_delta(key, n, i, i - d);
  int lmax = 2;
  while (lbvh_delta(key, n, i, i + lmax * d) > dmin) lmax *= 2;
  int l = 0;
  for (int t = lmax / 2; t >= 1; t /= 2) if (lbvh_delta(key, n, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = lbvh_delta(key, n, i, j);
  int s = 0, t = l;
  do {
    t = (t + 1) >> 1;
    if (lbvh_delta(key, n, i, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  const int g = i + s * d + min(d, 0);
  const int lo = min(i, j), hi = max(i, j);
  const int cl = lo == g ? ~g : g, cr = hi == g + 1 ? ~(g + 1) : g + 1;
  child[2 * i] = cl; child[2 * i + 1] = cr;
  if (cl < 0) parent_leaf[g] = i; else parent_int[g] = i;
  if (cr < 0) parent_leaf[g + 1] = i; else parent_int[g + 1] = i;
  if (i == 0) parent_int[0] = -1;
}
/* bottom-up: the second thread to arrive at a node computes its box (union of float boxes: exact) and depth */
__global__ void k_lbvh_refit(int n, const uint32_t* __restrict__ sorted_prim, const double* __restrict__ plo, const double* __restrict__ phi,
                             const int* __restrict__ child, const int* __restrict__ parent_int, const int* __restrict__ parent_leaf,
                             float* __restrict__ leafbox, float* __restrict__ nodebox, int* __restrict__ nodedepth, int* __restrict__ visits) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t p = sorted_prim[i];
  float b[6];
#pragma unroll
  for (int k = 0; k < 3; ++k) { b[k] = __double2float_rd(plo[(size_t)k * n + p]); b[3 + k] = __double2float_ru(phi[(size_t)k * n + p]); }
#pragma unroll
  for (int k = 0; k < 6; ++k) leafbox[6 * (size_t)i + k] = b[k];
  __threadfence();
  int node = parent_leaf[i];
  while (node >= 0) {
    if (atomicAdd(visits + node, 1) == 0) return;    /* first arrival: the sibling subtree is not ready yet */
    __threadfence();
    float bb[6]; int dep = 0;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const int code = child[2 * node + c];
      const volatile float* src = code < 0 ? leafbox + 6 * (size_t)(~code) : nodebox + 6 * (size_t)code;
      const int dc = code < 0 ? 0 : ((const volatile int*)nodedepth)[code];
      dep = max(dep, dc);
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        bb[k] = c == 0 ? src[k] : fminf(bb[k], src[k]);
        bb[3 + k] = c == 0 ? src[3 + k] : fmaxf(bb[3 + k], src[3 + k]);
      }
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) nodebox[6 * (size_t)node + k] = bb[k];
    nodedepth[node] = dep + 1;
    __threadfence();
    node = parent_int[node];
  }
}
/* 64-byte node records (rrt_bvh.h): both children's boxes + the two child codes; leaf code ~((first << 3) | count) */
__global__ void k_lbvh_emit(int n, const int* __restrict__ child, const float* __restrict__ leafbox, const float* __restrict__ nodebox, float4* __restrict__ nodes) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const int cl = child[2 * i], cr = child[2 * i + 1];
  const float* L = cl < 0 ? leafbox + 6 * (size_t)(~cl) : nodebox + 6 * (size_t)cl;
  const float* R = cr < 0 ? leafbox + 6 * (size_t)(~cr) : nodebox + 6 * (size_t)cr;
  const int kl = cl < 0 ? ~(int)((((uint32_t)~cl) << 3) | 1u) : cl, kr = cr < 0 ? ~(int)((((uint32_t)~cr) << 3) | 1u) : cr;
  nodes[4 * (size_t)i + 0] = make_float4(L[0], L[1], L[2], L[3]);
  nodes[4 * (size_t)i + 1] = make_float4(L[4], L[5], R[0], R[1]);
  nodes[4 * (size_t)i + 2] = make_float4(R[2], R[3], R[4], R[5]);
  nodes[4 * (size_t)i + 3] = make_float4(__int_as_float(kl), __int_as_float(kr), 0.f, 0.f);
}
/* 80-byte leaf records in the tree's leaf order (rrt_bvh.h): the reference's doubles, primitive id, rank, kind */
__global__ void k_leaf_records(uint32_t n, const uint32_t* __restrict__ sorted_prim, const uint8_t* __restrict__ kind, const uint32_t* __restrict__ index,
                               const double* __restrict__ tri_pos, const double* __restrict__ sph, const uint32_t* __restrict__ rank, double* __restrict__ leaf) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t p = sorted_prim[i], kd = kind[p];
  double* r = leaf + 10 * (size_t)i;
  if (kd == 0) {
    const double* q = tri_pos + 9 * (size_t)index[p];
#pragma unroll
    for (int k = 0; k < 3; ++k) { r[k] = q[k]; r[3 + k] = q[3 + k] - q[k]; r[6 + k] = q[6 + k] - q[k]; }
  } else {
    const double* c = sph + 4 * (size_t)index[p];
    r[0] = c[0]; r[1] = c[1]; r[2] = c[2]; r[3] = c[3] * c[3];
    r[4] = r[5] = r[6] = r[7] = r[8] = 0.0;
  }
  const unsigned long long tag = (unsigned long long)p | ((unsigned long long)((rank[p] << 1) | kd) << 32);
  r[9] = __longlong_as_double((long long)tag);
}

}  // namespace rrtdev
