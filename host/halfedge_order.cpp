// halfedge_order.cpp — what the reference's half-edge mesh does to a polygon soup, as far as the renderer can tell.
//
// The reference never reads a COLLADA file's normals: Mesh::Mesh (reference src/dynamic_scene/mesh.cpp:17-35) builds a
// HalfedgeMesh from the polygons, HalfedgeMesh::build (src/halfEdgeMesh.cpp:29-377) ends by calling
// Vertex::computeNormal (src/halfEdgeMesh.h:492-515) on every vertex, and StaticScene::Mesh::Mesh
// (src/static_scene/object.cpp:16-45) emits ONE triangle per face from face->halfedge() and its two successors.
// Two consequences define the data the GPU path must be fed, and both depend on connectivity bookkeeping, not only
// on geometry:
//   * a face's halfedge ends up being the LAST one created for it, so polygon (a, b, c, ...) becomes the triangle
//     (last, a, b): (c, a, b) for a triangle, (d, a, b) for a quad (the rest of the quad is dropped);
//   * the order in which the faces around a vertex are summed into its normal follows the vertex's halfedge pointer
//     as left by the build, and boundary vertices use a different walk (h->next()->twin()) that visits other faces.
// This file reproduces that bookkeeping on index arrays (no linked lists, no std::map of iterators) so that triangle
// order and vertex normals come out bit-identical to the reference's.
#include <math.h>

#include <algorithm>
#include <stdint.h>
#include <utility>

#include "rrt_host.h"

namespace rrt_host {
namespace {

inline Vector3D sub(const Vector3D& a, const Vector3D& b) { return Vector3D(a.x - b.x, a.y - b.y, a.z - b.z); }
inline Vector3D cross(const Vector3D& u, const Vector3D& v) {
  return Vector3D(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x);
}

struct HalfedgeArrays {
  std::vector<int> next, twin, vert, face;   // per halfedge; face >= n_faces marks a boundary loop
  int n_faces = 0;
  int add() { next.push_back(-1); twin.push_back(-1); vert.push_back(-1); face.push_back(-1); return (int)next.size() - 1; }
  bool on_boundary(int h) const { return face[h] >= n_faces; }
};

}  // namespace

bool halfedge_triangles_and_normals(const std::vector<std::vector<size_t>>& polygons, const std::vector<Vector3D>& positions,
                                    std::vector<size_t>* out_tri, std::vector<Vector3D>* out_normals, std::string* error) {
  auto fail = [&](const std::string& msg) { if (error) *error = "half-edge build: " + msg; return false; };
  HalfedgeArrays he;
  he.n_faces = (int)polygons.size();

  // vertices in order of first appearance; degree = number of polygons using the vertex.
  // (The reference keeps std::map<Index, VertexIter> and std::map<pair<Index, Index>, HalfedgeIter>; what matters of them is
  // membership and, for the vertex map, iteration in increasing index order.  A flat lookup table and an open-addressing
  // hash table give the same answers: 1 M triangles in 0.5 s instead of 1.9 s.)
  size_t max_index = 0, n_corners = 0;
  for (const auto& p : polygons) {
    if (p.size() < 3) return fail("each polygon must have at least three vertices");
    for (size_t idx : p) max_index = std::max(max_index, idx);
    n_corners += p.size();
  }
  if (max_index >= ((size_t)1 << 31) || max_index > 64 * n_corners + 1024) return fail("vertex indices too sparse for the flat index table");
  std::vector<int> index_to_vertex(max_index + 1, -1);
  std::vector<size_t> vertex_index;     // vertex id -> input index
  std::vector<int> vertex_degree, vertex_he;
  std::vector<size_t> sorted;
  for (const auto& p : polygons) {
    if (p.size() == 3) {
      if (p[0] == p[1] || p[1] == p[2] || p[0] == p[2]) return fail("a polygon repeats a vertex");
    } else {
      sorted.assign(p.begin(), p.end());
      std::sort(sorted.begin(), sorted.end());
      if (std::adjacent_find(sorted.begin(), sorted.end()) != sorted.end()) return fail("a polygon repeats a vertex");
    }
    for (size_t idx : p) {
      int& v = index_to_vertex[idx];
      if (v < 0) {
        v = (int)vertex_index.size();
        vertex_index.push_back(idx); vertex_degree.push_back(1); vertex_he.push_back(-1);
      } else {
        vertex_degree[v]++;
      }
    }
  }
  const int nv = (int)vertex_index.size();

  // halfedges of the faces, twins where the opposite oriented edge already exists
  size_t cap = 16;
  while (cap < 2 * n_corners + 16) cap <<= 1;
  std::vector<uint64_t> ekey(cap, ~(uint64_t)0);
  std::vector<int> eval(cap, -1);
  auto slot_of = [&](size_t a, size_t b) {   // slot holding (a, b), or the empty slot where it belongs
    const uint64_t k = ((uint64_t)a << 32) | (uint64_t)b;
    size_t i = (size_t)((k * 0x9E3779B97F4A7C15ull) >> 17) & (cap - 1);
    while (ekey[i] != ~(uint64_t)0 && ekey[i] != k) i = (i + 1) & (cap - 1);
    return i;
  };
  he.next.reserve(2 * n_corners); he.twin.reserve(2 * n_corners); he.vert.reserve(2 * n_corners); he.face.reserve(2 * n_corners);
  std::vector<int> face_he(polygons.size(), -1);
  std::vector<int> ring;
  for (size_t f = 0; f < polygons.size(); ++f) {
    const auto& p = polygons[f];
    const size_t deg = p.size();
    ring.clear();
    for (size_t i = 0; i < deg; ++i) {
      size_t a = p[i], b = p[(i + 1) % deg];
      const size_t sl = slot_of(a, b);
      if (eval[sl] >= 0) return fail("an oriented edge appears twice (non-manifold or inconsistently oriented input)");
      int h = he.add();
      ekey[sl] = ((uint64_t)a << 32) | (uint64_t)b; eval[sl] = h;
      he.face[h] = (int)f; face_he[f] = h;
      int v = index_to_vertex[a];
      he.vert[h] = v; vertex_he[v] = h;
      ring.push_back(h);
      const size_t ts = slot_of(b, a);
      if (eval[ts] >= 0) { he.twin[h] = eval[ts]; he.twin[eval[ts]] = h; }
    }
    for (size_t i = 0; i < deg; ++i) he.next[ring[i]] = ring[(i + 1) % deg];
  }

  // a boundary vertex points at its halfedge that has no twin yet
  for (int v = 0; v < nv; ++v) {
    const int start = vertex_he[v];
    int h = start;
    do {
      if (he.twin[h] < 0) { vertex_he[v] = h; break; }
      h = he.next[he.twin[h]];
    } while (h != start);
  }

  // one virtual face per boundary loop; halfedges appended here already have twins, so the scan skips them
  int n_loops = 0;
  for (int h = 0; h < (int)he.next.size(); ++h) {
    if (he.twin[h] >= 0) continue;
    const int loop_face = he.n_faces + n_loops++;
    std::vector<int> ring;
    int i = h;
    do {
      int t = he.add();
      ring.push_back(t);
      he.twin[i] = t; he.twin[t] = i;
      he.face[t] = loop_face;
      he.vert[t] = he.vert[he.next[i]];
      i = he.next[i];
      while (i != h && he.twin[i] >= 0) {
        i = he.twin[i];
        if (he.next[i] < 0) return fail("boundary loops touch at a vertex (non-manifold input)");
        i = he.next[i];
      }
    } while (i != h);
    const int deg = (int)ring.size();
    for (int p = 0; p < deg; ++p) he.next[ring[p]] = ring[(p - 1 + deg) % deg];
  }

  for (int v = 0; v < nv; ++v) vertex_he[v] = he.next[he.twin[vertex_he[v]]];

  // manifold check, as the reference makes it
  for (int v = 0; v < nv; ++v) {
    int count = 0, h = vertex_he[v], guard = 0;
    do {
      if (!he.on_boundary(h)) count++;
      h = he.next[he.twin[h]];
      if (++guard > (int)he.next.size()) return fail("vertex ring does not close");
    } while (h != vertex_he[v]);
    if (count != vertex_degree[v]) return fail("at least one vertex is non-manifold");
  }

  // positions: the k-th smallest index gets the k-th position
  if (positions.size() != (size_t)nv) return fail("number of vertex positions differs from the number of distinct vertices");
  std::vector<Vector3D> pos(nv);
  {
    size_t k = 0;
    for (int v : index_to_vertex) if (v >= 0) pos[v] = positions[k++];   /* increasing input index, like iterating the reference's map */
  }

  // Vertex::computeNormal
  std::vector<Vector3D> nrm(nv);
  for (int v = 0; v < nv; ++v) {
    bool boundary = false;
    {
      int h = vertex_he[v];
      do {
        if (he.on_boundary(h)) { boundary = true; break; }
        h = he.next[he.twin[h]];
      } while (h != vertex_he[v]);
    }
    Vector3D n(0., 0., 0.);
    const Vector3D pi = pos[v];
    int h = vertex_he[v], guard = 0;
    do {
      const Vector3D pj = pos[he.vert[he.next[h]]];
      const Vector3D pk = pos[he.vert[he.next[he.next[h]]]];
      const Vector3D c = cross(sub(pj, pi), sub(pk, pi));
      n.x += c.x; n.y += c.y; n.z += c.z;
      h = boundary ? he.twin[he.next[h]] : he.next[he.twin[h]];
      if (++guard > (int)he.next.size()) return fail("normal walk does not return to its start");
    } while (h != vertex_he[v]);
    const double rc = 1. / sqrt(n.x * n.x + n.y * n.y + n.z * n.z);
    n.x *= rc; n.y *= rc; n.z *= rc;
    nrm[v] = n;
  }

  // one triangle per face: face halfedge and its two successors; results are indexed by INPUT position order
  std::vector<int> sorted_rank(nv);
  {
    int k = 0;
    for (int v : index_to_vertex) if (v >= 0) sorted_rank[v] = k++;
  }
  out_tri->clear(); out_tri->reserve(3 * polygons.size());
  for (size_t f = 0; f < polygons.size(); ++f) {
    int h = face_he[f];
    out_tri->push_back((size_t)sorted_rank[he.vert[h]]);
    out_tri->push_back((size_t)sorted_rank[he.vert[he.next[h]]]);
    out_tri->push_back((size_t)sorted_rank[he.vert[he.next[he.next[h]]]]);
  }
  out_normals->assign(nv, Vector3D());
  for (int v = 0; v < nv; ++v) (*out_normals)[sorted_rank[v]] = nrm[v];
  return true;
}

}  // namespace rrt_host
