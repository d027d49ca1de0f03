// scene_io.cpp — the flattened scene container and its "RRTSCN01" file form (layout: DESIGN.md section 3; the same
// files are written by oracle/ref_probe.cpp from the reference's own objects and read by relativistic-ray-tracer_b200/scene.py).
#include <stdio.h>
#include <string.h>

#include "rrt_host.h"

namespace rrt_host {

rrt_scene FlatScene::view() const {
  rrt_scene s;
  memset(&s, 0, sizeof s);
  s.n_prims = (uint32_t)prim_kind.size(); s.prim_kind = prim_kind.data(); s.prim_index = prim_index.data();
  s.n_tris = (uint32_t)tri_mat.size(); s.tri_pos = tri_pos.data(); s.tri_nrm = tri_nrm.data(); s.tri_mat = tri_mat.data();
  s.n_spheres = (uint32_t)sph_mat.size(); s.sph = sph.data(); s.sph_mat = sph_mat.data();
  s.n_mats = (uint32_t)mats.size(); s.mats = mats.data();
  s.n_lights = (uint32_t)lights.size(); s.lights = lights.data();
  return s;
}

namespace {
template <class T>
bool put(FILE* f, const std::vector<T>& v) {
  size_t bytes = sizeof(T) * v.size(), pad = (8 - bytes % 8) % 8;
  const char z[8] = {0};
  if (bytes && fwrite(v.data(), 1, bytes, f) != bytes) return false;
  return fwrite(z, 1, pad, f) == pad;
}
template <class T>
bool take(FILE* f, std::vector<T>* v, size_t n) {
  /* the count comes from the file: nothing is allocated that the rest of the file cannot fill */
  const long here = ftell(f);
  if (here < 0 || fseek(f, 0, SEEK_END) != 0) return false;
  const long end = ftell(f);
  if (end < here || fseek(f, here, SEEK_SET) != 0) return false;
  if (n > (size_t)(end - here) / sizeof(T)) return false;
  v->resize(n);
  size_t bytes = sizeof(T) * n, pad = (8 - bytes % 8) % 8;
  char z[8];
  if (bytes && fread(v->data(), 1, bytes, f) != bytes) return false;
  return fread(z, 1, pad, f) == pad;
}
}  // namespace

bool save_rrtscene(const std::string& filename, const FlatScene& s, const Camera& c, std::string* error) {
  FILE* f = fopen(filename.c_str(), "wb");
  if (!f) { if (error) *error = "cannot write " + filename; return false; }
  uint32_t hdr[8] = {(uint32_t)s.prim_kind.size(), (uint32_t)s.tri_mat.size(), (uint32_t)s.sph_mat.size(), (uint32_t)s.mats.size(),
                     (uint32_t)s.lights.size(), 0, 0, 0};
  double camv[24] = {c.pos.x, c.pos.y, c.pos.z, c.c2w[0].x, c.c2w[0].y, c.c2w[0].z, c.c2w[1].x, c.c2w[1].y, c.c2w[1].z,
                     c.c2w[2].x, c.c2w[2].y, c.c2w[2].z, c.hFov, c.vFov, c.nClip, c.fClip, c.lensRadius, c.focalDistance,
                     (double)c.screenW, (double)c.screenH, c.screenDist, 0, 0, 0};
  bool ok = fwrite("RRTSCN01", 1, 8, f) == 8 && fwrite(hdr, 4, 8, f) == 8 && fwrite(camv, 8, 24, f) == 24 && put(f, s.prim_kind) &&
            put(f, s.prim_index) && put(f, s.tri_pos) && put(f, s.tri_nrm) && put(f, s.tri_mat) && put(f, s.sph) && put(f, s.sph_mat) &&
            put(f, s.mats) && put(f, s.lights);
  ok = (fclose(f) == 0) && ok;
  if (!ok && error) *error = "short write to " + filename;
  return ok;
}

bool load_rrtscene(const std::string& filename, FlatScene* s, Camera* c, std::string* error) {
  auto fail = [&](const char* m) { if (error) *error = filename + ": " + m; return false; };
  FILE* f = fopen(filename.c_str(), "rb");
  if (!f) return fail("cannot open");
  char magic[8]; uint32_t hdr[8]; double camv[24];
  bool ok = fread(magic, 1, 8, f) == 8 && !memcmp(magic, "RRTSCN01", 8) && fread(hdr, 4, 8, f) == 8 && fread(camv, 8, 24, f) == 24;
  if (!ok) { fclose(f); return fail("not an RRTSCN01 scene file"); }
  *s = FlatScene();
  ok = take(f, &s->prim_kind, hdr[0]) && take(f, &s->prim_index, hdr[0]) && take(f, &s->tri_pos, 9 * (size_t)hdr[1]) &&
       take(f, &s->tri_nrm, 9 * (size_t)hdr[1]) && take(f, &s->tri_mat, hdr[1]) && take(f, &s->sph, 4 * (size_t)hdr[2]) &&
       take(f, &s->sph_mat, hdr[2]) && take(f, &s->mats, hdr[3]) && take(f, &s->lights, hdr[4]);
  fclose(f);
  if (!ok) return fail("truncated scene file");
  *c = Camera();
  c->pos = Vector3D(camv[0], camv[1], camv[2]);
  for (int j = 0; j < 3; ++j) c->c2w[j] = Vector3D(camv[3 + 3 * j], camv[4 + 3 * j], camv[5 + 3 * j]);
  c->hFov = camv[12]; c->vFov = camv[13]; c->nClip = camv[14]; c->fClip = camv[15]; c->lensRadius = camv[16]; c->focalDistance = camv[17];
  c->screenW = (size_t)camv[18]; c->screenH = (size_t)camv[19]; c->screenDist = camv[20];
  c->ar = c->screenH ? (double)c->screenW / (double)c->screenH : 1.0;
  return true;
}

}  // namespace rrt_host
