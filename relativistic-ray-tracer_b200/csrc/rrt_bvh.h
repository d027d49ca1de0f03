// rrt_bvh.h — host-side BVH construction and flattening for the device traversal.
//
// Replaces BVHAccel::BVHAccel / construct_bvh (reference src/bvh.cpp:14-19,49-96) and the heap-allocated
// BVHNode tree (reference src/bvh.h:24-41).  Two things come out of the build:
//   1. the reference's own tree shape (midpoint split on the strictly largest axis, leaves of <= 4 primitives,
//      index halving when a side is empty) — needed because the reference keeps the LAST primitive in its
//      depth-first order among equal-t hits (src/static_scene/triangle.cpp:43-45, src/bvh.cpp:120-135), so every
//      primitive carries its depth-first rank and the device breaks ties toward the higher rank;
//   2. the flattened arrays the kernels read: 64-byte interior nodes holding BOTH children's boxes (outward
//      rounded floats) and 80-byte leaf records holding the reference's doubles.
#pragma once
#include <stdint.h>
#include <vector>
#include "../../include/rrt_b200.h"

namespace rrt {

// 64-byte node = 4 x float4:
//   q0 = (L.min.x, L.min.y, L.min.z, L.max.x)   q1 = (L.max.y, L.max.z, R.min.x, R.min.y)
//   q2 = (R.min.z, R.max.x, R.max.y, R.max.z)   q3 = (left, right, 0, 0) as int32
// child >= 0: interior node index; child < 0: leaf, ~child = (first record << 3) | record count.
// 80-byte leaf record = 9 doubles + 2 x uint32:
//   triangle: p0, e1 = p1 - p0, e2 = p2 - p0     sphere: centre, r*r, 0...
//   word0 = primitive index in reference order; word1 = (depth-first rank << 1) | kind (0 triangle, 1 sphere)
struct FlatBvh {
  std::vector<float> nodes;       // 16 floats per node
  std::vector<double> leaf;       // 10 doubles per record
  std::vector<uint32_t> dfs_rank; // [n_prims]
  std::vector<int32_t> prim_mat;  // [n_prims] material of each primitive
  int32_t root = -1;              // child code of the root (0 if interior, leaf code if the whole scene is one leaf)
  double root_lo[3] = {0, 0, 0}, root_hi[3] = {0, 0, 0};
  int n_nodes = 0;                // interior nodes in the flattened array
  // wide form (see collapse in rrt_bvh.cpp): the binary tree collapsed into nodes of up to `width` children.
  //   node = 8 * width floats: lo.x[width] lo.y[] lo.z[] hi.x[] hi.y[] hi.z[] (outward rounded), child code[width] as int32, pad[width]
  //   child code as above; an unused slot holds kEmptyChild and an inverted box
  std::vector<float> wide;        // 8 * width floats per node (256 B for width 8)
  int width = 2;                  // 2 = `nodes` is what the device reads; 4 / 8 = `wide`
  int n_wide = 0;
  int wide_depth = 0;             // levels of wide interior nodes
  int32_t wide_root = -1;
  // clearance grid (see build_clearance_grid): per cell, a lower bound on the distance from ANY point of the cell to ANY primitive
  std::vector<float> clearance;   // [gz][gy][gx]
  int grid_n[3] = {0, 0, 0};
  double grid_lo[3] = {0, 0, 0}, grid_inv_h[3] = {0, 0, 0};
  int ref_nodes = 0, ref_leaves = 0, ref_depth = 0; // shape of the reference tree (nodes incl. leaves)
  bool device_built = false;      // built by rrt_bvh_device.cuh: nodes / leaf live only on the device
  int device_depth = 0;           // its interior levels
};

static const int32_t kEmptyChild = (int32_t)0x80000000;   // not a valid leaf code: n_prims < 2^28

// topology: 0 = reference shape (midpoint split, leaves <= 4), 1 = binned SAH with leaves of <= max_leaf primitives
// (own shape; ranks still come from the reference shape).  max_leaf is clamped to 1..7 and ignored for topology 0.
// width: 2 keeps the binary nodes; 4 or 8 also fills the wide form.
int build_bvh(const rrt_scene* s, int topology, int max_leaf, FlatBvh* out, int width = 2);
// Empty-space grid over the scene box, res cells along the longest axis (0 disables).  The device skips the BVH for a chord
// whose length is below the clearance of the cell its origin lies in: no primitive can be within reach, so the
// result is unchanged.
void build_clearance_grid(const rrt_scene* s, int res, FlatBvh* out);
// outward float rounding used for every box the device sees
float round_down_f(double x);
float round_up_f(double x);

}  // namespace rrt
