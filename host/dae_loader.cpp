// dae_loader.cpp — COLLADA (.dae) scenes with the reference loader's semantics, straight into the flattened scene the
// C-ABI takes.
//
// What is reproduced (all of it host-side, none of it hot):
//   ColladaParser::load / parse_node / parse_camera / parse_light / parse_sphere / parse_polymesh / parse_material
//       reference src/collada/collada.cpp:131-225, 234-430, 432-477, 479-575, 577-602, 604-850, 852-936
//   Application::load, init_sphere, init_polymesh, init_light        src/application.cpp:219-345
//   DynamicScene lights -> StaticScene lights                          src/dynamic_scene/*_light.h, src/static_scene/light.cpp:10-78
//   Camera::configure / place                                          src/camera.cpp:22-54 (in host/camera.cpp)
//   half-edge vertex order and vertex normals                          host/halfedge_order.cpp
// including the quirks that shape the data: Z_UP / X_UP scenes enter through a fixed axis-swap matrix; a node's
// transform is its <matrix> only (the reference's <rotate>/<translate>/<scale> branches multiply uninitialised
// matrices; here they compose from the identity); vertex positions are parsed as FLOAT and widened; the camera's view
// direction is the node transform applied to (0,0,-1,1) INCLUDING the translation, normalised; file normals and
// texture coordinates are ignored; a polygon with more than three corners yields a single triangle.
// The arithmetic follows the reference's operation order so that the flattened scene equals, bit for bit, the one the
// reference's own objects hold (tests/test_host_cpp.py compares with the dumps under tests/golden/).
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <sstream>

#include "rrt_host.h"
#include "xml_lite.h"

namespace rrt_host {
namespace {

struct Vec4 { double x, y, z, w; };
struct Mat4 {              // m[i][j] = row i, column j
  double m[4][4];
  static Mat4 identity() { Mat4 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = i == j ? 1. : 0.; return r; }
};
Mat4 mul(const Mat4& A, const Mat4& B) {
  Mat4 C;
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      double c = 0.;
      for (int k = 0; k < 4; ++k) c += A.m[i][k] * B.m[k][j];
      C.m[i][j] = c;
    }
  return C;
}
Vec4 mul(const Mat4& A, const Vec4& v) {   // sum of scaled columns, left to right
  Vec4 r;
  r.x = v.x * A.m[0][0] + v.y * A.m[0][1] + v.z * A.m[0][2] + v.w * A.m[0][3];
  r.y = v.x * A.m[1][0] + v.y * A.m[1][1] + v.z * A.m[1][2] + v.w * A.m[1][3];
  r.z = v.x * A.m[2][0] + v.y * A.m[2][1] + v.z * A.m[2][2] + v.w * A.m[2][3];
  r.w = v.x * A.m[3][0] + v.y * A.m[3][1] + v.z * A.m[3][2] + v.w * A.m[3][3];
  return r;
}
inline Vector3D to3D(const Vec4& v) { return Vector3D(v.x, v.y, v.z); }
inline Vector3D project3D(const Vec4& v) { double iw = 1.0 / v.w; return Vector3D(v.x * iw, v.y * iw, v.z * iw); }
inline Vec4 point(const Vector3D& p) { return Vec4{p.x, p.y, p.z, 1.}; }
inline Vector3D sub(const Vector3D& a, const Vector3D& b) { return Vector3D(a.x - b.x, a.y - b.y, a.z - b.z); }
inline Vector3D neg(const Vector3D& a) { return Vector3D(-a.x, -a.y, -a.z); }
inline Vector3D cross(const Vector3D& u, const Vector3D& v) { return Vector3D(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x); }
inline double norm(const Vector3D& v) { return sqrt(v.x * v.x + v.y * v.y + v.z * v.z); }
inline Vector3D unit(const Vector3D& v) { double r = 1. / sqrt(v.x * v.x + v.y * v.y + v.z * v.z); return Vector3D(r * v.x, r * v.y, r * v.z); }
inline void normalize(Vector3D& v) { double c = 1. / norm(v); v.x *= c; v.y *= c; v.z *= c; }

const double kPI = 3.14159265358979323;
inline double radians(double deg) { return deg * (kPI / 180); }
inline double degrees(double rad) { return rad * (180 / kPI); }

struct Spectrum { float r = 0, g = 0, b = 0; };
Spectrum spectrum_from(const char* text) {
  Spectrum s;
  std::stringstream ss(text ? text : "");
  ss >> s.r; ss >> s.g; ss >> s.b;
  return s;
}

enum NodeKind { K_NONE, K_CAMERA, K_LIGHT, K_SPHERE, K_MESH };
enum LightKind { L_NONE, L_AMBIENT, L_DIRECTIONAL, L_AREA, L_POINT, L_SPOT };

struct NodeRec {
  NodeKind kind = K_NONE;
  Mat4 transform = Mat4::identity();
  // camera
  float hFov = 0, vFov = 0, nClip = 0, fClip = 0;
  // light
  LightKind light = L_NONE;
  Spectrum spectrum;
  // sphere
  float radius = 0;
  // mesh
  std::vector<Vector3D> vertices;
  std::vector<std::vector<size_t>> polygons;
  // material (sphere / mesh)
  bool has_material = false;
  rrt_material material;
};

struct Loader {
  std::map<std::string, const XmlNode*> sources;
  Mat4 transform = Mat4::identity();
  std::vector<NodeRec> nodes;
  std::string err;

  // "#id" -> "id"; an empty attribute has nothing after its terminator
  static const char* ref_id(const char* s) { return (s && *s) ? s + 1 : ""; }

  void uri_load(const XmlNode* e) {
    if (const char* id = e->attr("id")) sources[id] = e;
    for (const auto& c : e->children) uri_load(c.get());
  }
  const XmlNode* uri_find(const std::string& id) const {
    auto it = sources.find(id);
    return it == sources.end() ? nullptr : it->second;
  }
  // child-by-child descent, then one level of url indirection on the element reached
  const XmlNode* get_element(const XmlNode* xml, const std::string& query) const {
    const XmlNode* e = xml->path(query);
    if (e)
      if (const char* url = e->attr("url")) e = uri_find(ref_id(url));
    return e;
  }
  const XmlNode* technique_common(const XmlNode* xml) const {
    if (const XmlNode* prof = xml->child("profile_COMMON"))
      for (const XmlNode* t = prof->child("technique"); t; t = t->next("technique")) {
        const char* sid = t->attr("sid");
        if (sid && !strcmp(sid, "common")) return t;
      }
    return xml->child("technique_common");
  }
  const XmlNode* technique_CGL(const XmlNode* xml) const {
    for (const XmlNode* t = get_element(xml, "extra/technique"); t; t = t->next("technique")) {
      const char* prof = t->attr("profile");
      if (prof && !strcmp(prof, "CGL")) return t;
    }
    return nullptr;
  }
  static const char* text_of(const XmlNode* e) { return e && e->has_text ? e->text.c_str() : nullptr; }
  bool fail(const std::string& m) { if (err.empty()) err = m; return false; }

  bool parse_material(const XmlNode* xml, rrt_material* out) {
    memset(out, 0, sizeof *out);
    auto set3 = [](float* d, const Spectrum& s) { d[0] = s.r; d[1] = s.g; d[2] = s.b; };
    auto diffuse = [&](const Spectrum& s) { memset(out, 0, sizeof *out); out->type = RRT_MAT_DIFFUSE; set3(out->a, s); };
    const XmlNode* effect = get_element(xml, "instance_effect");
    if (!effect) return fail(std::string("no target effects found for material: ") + (xml->attr("id") ? xml->attr("id") : "?"));
    const XmlNode* common = technique_common(effect);
    const XmlNode* cgl = technique_CGL(effect);
    Spectrum grey; grey.r = grey.g = grey.b = .5f;
    if (cgl) {
      bool any = false;
      for (const XmlNode* b = cgl->child(); b; b = b->next()) {
        const std::string& type = b->name;
        auto spec = [&](const char* q) { return spectrum_from(text_of(get_element(b, q))); };
        auto num = [&](const char* q) { const char* t = text_of(get_element(b, q)); return (float)atof(t ? t : "0"); };
        if (type == "emission") { memset(out, 0, sizeof *out); out->type = RRT_MAT_EMISSION; set3(out->a, spec("radiance")); any = true; }
        else if (type == "mirror") { memset(out, 0, sizeof *out); out->type = RRT_MAT_MIRROR; set3(out->a, spec("reflectance")); any = true; }
        else if (type == "microfacet") {
          memset(out, 0, sizeof *out); out->type = RRT_MAT_MICROFACET;
          out->alpha = num("alpha"); set3(out->a, spec("eta")); set3(out->b, spec("k")); any = true;
        } else if (type == "refraction") {
          memset(out, 0, sizeof *out); out->type = RRT_MAT_REFRACTION; set3(out->b, spec("transmittance")); out->ior = num("ior"); any = true;
        } else if (type == "glass") {
          memset(out, 0, sizeof *out); out->type = RRT_MAT_GLASS;
          set3(out->b, spec("transmittance")); set3(out->a, spec("reflectance")); out->ior = num("ior"); any = true;
        }
      }
      if (!any) return fail("material with a CGL technique but no known BSDF element (the reference leaves its BSDF pointer unset)");
    } else if (common) {
      const XmlNode* d = get_element(common, "phong/diffuse/color");
      if (d) diffuse(spectrum_from(text_of(d))); else diffuse(grey);
    } else {
      diffuse(grey);
    }
    return true;
  }

  bool bind_material(const XmlNode* node_xml, NodeRec* rec) {
    const XmlNode* im = get_element(node_xml, "instance_geometry/bind_material/technique_common/instance_material");
    if (!im) return true;
    const char* target = im->attr("target");
    if (!target) return fail("no target material in instance_material");
    const XmlNode* m = uri_find(ref_id(target));
    if (!m) return fail(std::string("invalid target material id: ") + ref_id(target));
    rec->has_material = true;
    return parse_material(m, &rec->material);
  }

  bool parse_camera(const XmlNode* xml, NodeRec* rec) {
    const XmlNode* persp = get_element(xml, "optics/technique_common/perspective");
    if (!persp) return fail("no perspective defined in camera");
    const XmlNode *xf = persp->child("xfov"), *yf = persp->child("yfov"), *zn = persp->child("znear"), *zf = persp->child("zfar");
    rec->hFov = xf ? atof(text_of(xf) ? text_of(xf) : "0") : 50.0f;
    rec->vFov = yf ? atof(text_of(yf) ? text_of(yf) : "0") : 35.0f;
    rec->nClip = zn ? atof(text_of(zn) ? text_of(zn) : "0") : 0.001f;
    rec->fClip = zf ? atof(text_of(zf) ? text_of(zf) : "0") : 1000.0f;
    if (!yf) {
      const XmlNode* ar = get_element(persp, "aspect_ratio");
      if (!ar) return fail("incomplete perspective definition in camera");
      float aspect = atof(text_of(ar) ? text_of(ar) : "0");
      rec->vFov = 2 * degrees(atan(tan(radians(0.5 * rec->hFov)) / aspect));
    }
    return true;
  }

  bool parse_light(const XmlNode* xml, NodeRec* rec) {
    const XmlNode* common = technique_common(xml);
    const XmlNode* cgl = technique_CGL(xml);
    const XmlNode* tech = cgl ? cgl : common;
    if (!tech) return fail("no supported profile defined in light");
    rec->spectrum.r = rec->spectrum.g = rec->spectrum.b = 1.f;
    const XmlNode* l = tech->child();
    if (!l) return true;
    const std::string& type = l->name;
    const XmlNode* color = get_element(l, "color");
    if (type == "ambient") rec->light = L_AMBIENT;
    else if (type == "directional") rec->light = L_DIRECTIONAL;
    else if (type == "area") rec->light = L_AREA;
    else if (type == "point") {
      rec->light = L_POINT;
      if (!(get_element(l, "constant_attenuation") && get_element(l, "linear_attenuation") && get_element(l, "quadratic_attenuation")))
        return fail("incomplete definition of point light");
    } else if (type == "spot") {
      rec->light = L_SPOT;
      if (!(l->child("falloff_angle") && l->child("falloff_exponent") && get_element(l, "constant_attenuation") &&
            get_element(l, "linear_attenuation") && get_element(l, "quadratic_attenuation")))
        return fail("incomplete definition of spot light");
    } else return fail("light type " + type + " is not supported");
    if (!color) return fail("no color definition in " + type + " light");
    rec->spectrum = spectrum_from(text_of(color));
    return true;
  }

  bool parse_sphere(const XmlNode* geom, NodeRec* rec) {
    const XmlNode* tech = technique_CGL(geom);
    if (!tech) return fail("no CGL profile technique in sphere geometry");
    const XmlNode* radius = get_element(tech, "sphere/radius");
    if (!radius) return fail("invalid sphere definition in geometry");
    rec->radius = atof(text_of(radius) ? text_of(radius) : "0");
    return true;
  }

  bool parse_polymesh(const XmlNode* geom, NodeRec* rec) {
    const XmlNode* mesh = geom->child("mesh");
    if (!mesh) return fail("no mesh data defined in geometry");
    std::map<std::string, std::vector<float>> arrays;
    for (const XmlNode* src = mesh->child("source"); src; src = src->next("source")) {
      const char* id = src->attr("id");
      const XmlNode* fa = src->child("float_array");
      if (!id || !fa) continue;
      std::vector<float> floats;
      size_t count = (size_t)fa->int_attr("count");
      const char* p = text_of(fa) ? text_of(fa) : "";
      if (count > strlen(p) + 1) return fail("float_array count exceeds its data");   // a negative or damaged count: no allocation, no endless loop
      floats.reserve(count);
      float last = 0.f;
      for (size_t i = 0; i < count; ++i) {
        char* end = nullptr;
        float f = strtof(p, &end);
        if (end == p) f = last; else { last = f; p = end; }   // a short array repeats the last value, as a failed >> leaves it
        floats.push_back(f);
      }
      arrays[id] = std::move(floats);
    }
    const XmlNode* verts = mesh->child("vertices");
    if (!verts) return fail("no vertices defined in geometry");
    std::string vertices_id = verts->attr("id") ? verts->attr("id") : "";
    std::vector<Vector3D> vertices;
    for (const XmlNode* in = verts->child("input"); in; in = in->next("input")) {
      const char* sem = in->attr("semantic");
      if (!sem || strcmp(sem, "POSITION")) continue;
      const char* source = in->attr("source");
      auto it = arrays.find(ref_id(source));
      if (it == arrays.end()) return fail("undefined input source for POSITION");
      const std::vector<float>& f = it->second;
      for (size_t i = 0; i + 2 < f.size(); i += 3) vertices.push_back(Vector3D(f[i], f[i + 1], f[i + 2]));
    }
    const XmlNode* pl = mesh->child("polylist");
    if (!pl) return true;   // the reference accepts a mesh without polygons (and then builds an empty half-edge mesh)
    bool has_v = false, has_n = false, has_t = false;
    size_t v_off = 0;
    for (const XmlNode* in = pl->child("input"); in; in = in->next("input")) {
      const char* sem = in->attr("semantic");
      const char* source = in->attr("source");
      size_t off = (size_t)in->int_attr("offset");
      if (!sem) continue;
      if (!strcmp(sem, "VERTEX")) {
        has_v = true; v_off = off;
        if (!source || vertices_id != ref_id(source)) return fail("undefined source for VERTEX semantic");
        rec->vertices = vertices;
      } else if (!strcmp(sem, "NORMAL")) has_n = true;
      else if (!strcmp(sem, "TEXCOORD")) has_t = true;
    }
    size_t n_poly = (size_t)pl->int_attr("count");
    size_t stride = (has_v ? 1 : 0) + (has_n ? 1 : 0) + (has_t ? 1 : 0);
    /* the reference indexes p[k * stride + offset] without a check; a VERTEX offset outside the stride reads past the array there */
    if (has_v && v_off >= stride) return fail("VERTEX offset outside the polylist's index stride");
    const XmlNode* vcount = pl->child("vcount");
    if (!vcount) return fail("polygon sizes undefined in geometry");
    const XmlNode* pe = pl->child("p");
    if (!pe) return fail("no index array defined in geometry");
    const size_t vcount_len = strlen(text_of(vcount) ? text_of(vcount) : ""), p_len = strlen(text_of(pe) ? text_of(pe) : "");
    if (n_poly > vcount_len + 1) return fail("polylist count exceeds its vcount data");
    std::vector<size_t> sizes; sizes.reserve(n_poly);
    size_t n_idx = 0;
    {
      const char* p = text_of(vcount) ? text_of(vcount) : "";
      size_t last = 0;
      for (size_t i = 0; i < n_poly; ++i) {
        char* end = nullptr;
        unsigned long long v = strtoull(p, &end, 10);
        if (end == p) v = last; else { last = (size_t)v; p = end; }
        if (v > p_len + 1) return fail("polygon size exceeds the index data");
        sizes.push_back((size_t)v);
        n_idx += (size_t)v * stride;
        if (n_idx > p_len + 1) return fail("polygon sizes exceed the index data");
      }
    }
    std::vector<size_t> indices; indices.reserve(n_idx);
    {
      const char* p = text_of(pe) ? text_of(pe) : "";
      size_t last = 0;
      for (size_t i = 0; i < n_idx; ++i) {
        char* end = nullptr;
        unsigned long long v = strtoull(p, &end, 10);
        if (end == p) v = last; else { last = (size_t)v; p = end; }
        indices.push_back((size_t)v);
      }
    }
    rec->polygons.resize(n_poly);
    if (has_v) {
      size_t k = 0;
      for (size_t i = 0; i < n_poly; ++i)
        for (size_t j = 0; j < sizes[i]; ++j, ++k) {
          const size_t v = indices[k * stride + v_off];
          if (v >= rec->vertices.size()) return fail("polygon refers to a vertex that does not exist");
          rec->polygons[i].push_back(v);
        }
    }
    return true;
  }

  bool parse_node(const XmlNode* xml) {
    NodeRec rec;
    for (const XmlNode* e = xml->child(); e; e = e->next()) {
      const std::string& name = e->name;
      std::stringstream ss(text_of(e) ? text_of(e) : "");
      if (name == "matrix") {
        Mat4 mat = Mat4::identity();
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) ss >> mat.m[i][j];
        rec.transform = mat;
        break;
      }
      if (name == "rotate") {
        Mat4 m = Mat4::identity();
        const char* sid = e->attr("sid");
        char axis = sid && *sid ? sid[strlen(sid) - 1] : ' ';
        if (axis == 'X') { ss >> m.m[1][1]; ss >> m.m[1][2]; ss >> m.m[2][1]; ss >> m.m[2][2]; }
        else if (axis == 'Y') { ss >> m.m[0][0]; ss >> m.m[2][0]; ss >> m.m[0][2]; ss >> m.m[2][2]; }
        else if (axis == 'Z') { ss >> m.m[0][0]; ss >> m.m[0][1]; ss >> m.m[1][0]; ss >> m.m[1][1]; }
        rec.transform = mul(m, rec.transform);
      } else if (name == "translate") {
        Mat4 m = Mat4::identity();
        ss >> m.m[0][3]; ss >> m.m[1][3]; ss >> m.m[2][3];
        rec.transform = mul(m, rec.transform);
      } else if (name == "scale") {
        Mat4 m = Mat4::identity();
        ss >> m.m[0][0]; ss >> m.m[1][1]; ss >> m.m[1][1];   // sic: the reference reads the third value into (1,1) again
        rec.transform = mul(m, rec.transform);
      }
    }
    Mat4 saved = transform;
    rec.transform = mul(transform, rec.transform);
    transform = rec.transform;
    for (const XmlNode* c = get_element(xml, "node"); c; c = c->next("node"))
      if (!parse_node(c)) return false;
    transform = saved;

    const XmlNode* cam = get_element(xml, "instance_camera");
    const XmlNode* light = get_element(xml, "instance_light");
    const XmlNode* geom = get_element(xml, "instance_geometry");
    if (cam) { rec.kind = K_CAMERA; if (!parse_camera(cam, &rec)) return false; }
    else if (light) { rec.kind = K_LIGHT; if (!parse_light(light, &rec)) return false; }
    else if (geom) {
      if (get_element(geom, "mesh")) { rec.kind = K_MESH; if (!parse_polymesh(geom, &rec) || !bind_material(xml, &rec)) return false; }
      else if (get_element(geom, "extra")) { rec.kind = K_SPHERE; if (!parse_sphere(geom, &rec) || !bind_material(xml, &rec)) return false; }
    }
    nodes.push_back(std::move(rec));
    return true;
  }
};

struct Bounds {
  Vector3D lo = Vector3D(INFINITY, INFINITY, INFINITY), hi = Vector3D(-INFINITY, -INFINITY, -INFINITY);
  bool any = false;
  void add(const Vector3D& p) {
    for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); }
    any = true;
  }
};

}  // namespace

bool load_dae(const std::string& filename, FlatScene* scene, Camera* camera, std::string* error, size_t screen_w, size_t screen_h) {
  auto fail = [&](const std::string& m) { if (error) *error = filename + ": " + m; return false; };
  XmlDocument doc;
  if (!doc.load(filename)) return fail(doc.error);
  const XmlNode* root = doc.root->child("COLLADA");
  if (!root) return fail("not a COLLADA file");
  Loader L;
  L.uri_load(root);
  if (const XmlNode* asset = L.get_element(root, "asset")) {
    const XmlNode* up = L.get_element(asset, "up_axis");
    if (!up) return fail("no up direction defined");
    std::string dir = Loader::text_of(up) ? Loader::text_of(up) : "";
    L.transform = Mat4::identity();
    if (dir == "X_UP") {
      L.transform.m[0][0] = 0; L.transform.m[0][1] = 1; L.transform.m[1][0] = 1; L.transform.m[1][1] = 0; L.transform.m[2][2] = -1;
    } else if (dir == "Z_UP") {
      L.transform.m[1][1] = 0; L.transform.m[1][2] = 1; L.transform.m[2][1] = 1; L.transform.m[2][2] = 0; L.transform.m[0][0] = -1;
    } else if (dir != "Y_UP") return fail("invalid up direction");
  }
  const XmlNode* vs = L.get_element(root, "scene/instance_visual_scene");
  if (!vs) return fail("no scene description found");
  for (const XmlNode* n = L.get_element(vs, "node"); n; n = n->next("node"))
    if (!L.parse_node(n)) return fail(L.err);

  // ---- Application::load
  *scene = FlatScene();
  Camera cam;                                     // Application::init's dummy camera
  cam.configure(50, 35, (float)0.01, 100, screen_w, screen_h);
  Vector3D c_dir(0, 0, 0);
  Bounds bb;
  for (const NodeRec& n : L.nodes) {
    const Mat4& T = n.transform;
    switch (n.kind) {
      case K_CAMERA:
        c_dir = unit(to3D(mul(T, Vec4{0, 0, -1, 1})));
        cam.configure(n.hFov, n.vFov, n.nClip, n.fClip, screen_w, screen_h);
        break;
      case K_LIGHT: {
        rrt_light o; memset(&o, 0, sizeof o);
        o.radiance[0] = n.spectrum.r; o.radiance[1] = n.spectrum.g; o.radiance[2] = n.spectrum.b;
        const Vector3D pos0(0, 0, 0), dir0(0, 0, -1), up0(0, 1, 0);   // LightInfo defaults (reference src/collada/light_info.cpp:7-24)
        auto put = [](double* d, const Vector3D& v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; };
        if (n.light == L_AMBIENT) { o.type = RRT_LIGHT_HEMISPHERE; o.is_delta = 0; }
        else if (n.light == L_DIRECTIONAL) {
          Vector3D d = neg(to3D(mul(T, point(dir0)))); normalize(d);
          o.type = RRT_LIGHT_DIRECTIONAL; o.is_delta = 1; put(o.direction, neg(unit(d)));
        } else if (n.light == L_AREA) {
          Vector3D position = to3D(mul(T, point(pos0)));
          Vector3D direction = sub(to3D(mul(T, point(dir0))), position); normalize(direction);
          Vector3D dim_x = sub(to3D(mul(T, point(cross(up0, dir0)))), position);
          Vector3D dim_y = sub(to3D(mul(T, point(up0))), position);
          o.type = RRT_LIGHT_AREA; o.is_delta = 0;
          put(o.position, position); put(o.direction, direction); put(o.dim_x, dim_x); put(o.dim_y, dim_y);
          o.area = (float)(norm(dim_x) * norm(dim_y));
        } else if (n.light == L_POINT) {
          o.type = RRT_LIGHT_POINT; o.is_delta = 1; put(o.position, to3D(mul(T, point(pos0))));
        } else if (n.light == L_SPOT) {
          o.type = RRT_LIGHT_NULL; o.is_delta = 1;     /* StaticScene::SpotLight::sample_L returns no radiance */
        } else break;                                   /* LightType::NONE: Application::init_light returns nullptr */
        scene->lights.push_back(o);
        break;
      }
      case K_SPHERE: {
        Vector3D p = project3D(mul(T, Vec4{0, 0, 0, 1}));
        double scale = norm(to3D(mul(T, Vec4{1, 0, 0, 0})));
        double r = n.radius * scale;
        rrt_material m; memset(&m, 0, sizeof m);
        if (n.has_material) m = n.material; else { m.type = RRT_MAT_DIFFUSE; m.a[0] = m.a[1] = m.a[2] = 0.5f; }
        scene->prim_kind.push_back(1); scene->prim_index.push_back((uint32_t)scene->sph_mat.size());
        scene->sph.push_back(p.x); scene->sph.push_back(p.y); scene->sph.push_back(p.z); scene->sph.push_back(r);
        scene->sph_mat.push_back((int32_t)scene->mats.size());
        scene->mats.push_back(m);
        bb.add(Vector3D(p.x - r, p.y - r, p.z - r)); bb.add(Vector3D(p.x + r, p.y + r, p.z + r));
        break;
      }
      case K_MESH: {
        std::vector<Vector3D> v = n.vertices;
        for (auto& q : v) q = project3D(mul(T, point(q)));
        std::vector<size_t> tri; std::vector<Vector3D> nrm; std::string herr;
        if (!halfedge_triangles_and_normals(n.polygons, v, &tri, &nrm, &herr)) return fail(herr);
        rrt_material m; memset(&m, 0, sizeof m);
        if (n.has_material) m = n.material; else { m.type = RRT_MAT_DIFFUSE; m.a[0] = m.a[1] = m.a[2] = 0.5f; }
        const int32_t mi = (int32_t)scene->mats.size();
        if (!tri.empty()) scene->mats.push_back(m);     /* a material is numbered when its first primitive is met */
        for (size_t t = 0; t + 2 < tri.size(); t += 3) {
          scene->prim_kind.push_back(0); scene->prim_index.push_back((uint32_t)scene->tri_mat.size());
          scene->tri_mat.push_back(mi);
          for (int k = 0; k < 3; ++k) {
            const Vector3D& p = v[tri[t + k]]; const Vector3D& nn = nrm[tri[t + k]];
            scene->tri_pos.push_back(p.x); scene->tri_pos.push_back(p.y); scene->tri_pos.push_back(p.z);
            scene->tri_nrm.push_back(nn.x); scene->tri_nrm.push_back(nn.y); scene->tri_nrm.push_back(nn.z);
          }
        }
        for (const auto& q : v) bb.add(q);
        break;
      }
      default: break;
    }
  }
  if (bb.any) {
    Vector3D target(0.5 * (bb.lo.x + bb.hi.x), 0.5 * (bb.lo.y + bb.hi.y), 0.5 * (bb.lo.z + bb.hi.z));
    Vector3D extent = sub(bb.hi, bb.lo);
    double canonical = norm(extent) / 2 * 1.5;
    cam.place(target, acos(c_dir.y), atan2(c_dir.x, c_dir.z), canonical * 2, canonical / 10.0, canonical * 20.0);
  }
  *camera = cam;
  return true;
}

}  // namespace rrt_host
