// camera.cpp — Camera bookkeeping of the reference (src/camera.cpp:22-119, 138-169) for the host side: field of view
// against the frame, orbit placement, settings files.  Ray generation itself is on the device (rrt_generate_rays /
// k_generate); this class only produces the numbers rrt_set_camera takes.
#include <math.h>

#include <algorithm>
#include <fstream>

#include "rrt_host.h"

namespace rrt_host {
namespace {
const double kPI = 3.14159265358979323;
const float kEpsF = 0.00001f;
inline double radians(double deg) { return deg * (kPI / 180); }
inline double degrees(double rad) { return rad * (180 / kPI); }
inline Vector3D cross(const Vector3D& u, const Vector3D& v) { return Vector3D(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x); }
inline void normalize(Vector3D& v) { double c = 1. / sqrt(v.x * v.x + v.y * v.y + v.z * v.z); v.x *= c; v.y *= c; v.z *= c; }
inline Vector3D unit(const Vector3D& v) { double r = 1. / sqrt(v.x * v.x + v.y * v.y + v.z * v.z); return Vector3D(r * v.x, r * v.y, r * v.z); }
}  // namespace

void Camera::configure(double hFov_deg, double vFov_deg, double nClip_, double fClip_, size_t w, size_t h) {
  screenW = w; screenH = h;
  nClip = nClip_; fClip = fClip_; hFov = hFov_deg; vFov = vFov_deg;
  double ar1 = tan(radians(hFov) / 2) / tan(radians(vFov) / 2);
  ar = static_cast<double>(screenW) / screenH;
  if (ar1 < ar) hFov = 2 * degrees(atan(tan(radians(vFov) / 2) * ar));        // frame wider than the camera's aspect: widen the horizontal angle
  else if (ar1 > ar) vFov = 2 * degrees(atan(tan(radians(hFov) / 2) / ar));   // frame taller: widen the vertical angle
  screenDist = ((double)screenH) / (2.0 * tan(radians(vFov) / 2));
}

void Camera::place(const Vector3D& target, double phi_, double theta_, double r_, double minR_, double maxR_) {
  double rr = std::min(std::max(r_, minR_), maxR_);
  double ph = (sin(phi_) == 0) ? (phi_ + kEpsF) : phi_;
  targetPos = target; phi = ph; theta = theta_; r = rr; minR = minR_; maxR = maxR_;
  compute_position();
}

void Camera::set_screen_size(size_t w, size_t h) {
  screenW = w; screenH = h;
  ar = 1.0 * screenW / screenH;
  hFov = 2 * degrees(atan(((double)screenW) / (2 * screenDist)));
  vFov = 2 * degrees(atan(((double)screenH) / (2 * screenDist)));
}

void Camera::compute_position() {
  double sinPhi = sin(phi);
  if (sinPhi == 0) { phi += kEpsF; sinPhi = sin(phi); }
  const Vector3D dirToCamera(r * sinPhi * sin(theta), r * cos(phi), r * sinPhi * cos(theta));
  pos = Vector3D(targetPos.x + dirToCamera.x, targetPos.y + dirToCamera.y, targetPos.z + dirToCamera.z);
  Vector3D upVec(0, sinPhi > 0 ? 1 : -1, 0);
  Vector3D screenXDir = cross(upVec, dirToCamera); normalize(screenXDir);
  Vector3D screenYDir = cross(dirToCamera, screenXDir); normalize(screenYDir);
  c2w[0] = screenXDir; c2w[1] = screenYDir; c2w[2] = unit(dirToCamera);
}

bool Camera::dump_settings(const std::string& filename) const {
  std::ofstream file(filename);
  if (!file.is_open()) return false;
  file << hFov << " " << vFov << " " << ar << " " << nClip << " " << fClip << std::endl;
  for (int i = 0; i < 3; ++i) file << pos[i] << " ";
  for (int i = 0; i < 3; ++i) file << targetPos[i] << " ";
  file << std::endl;
  file << phi << " " << theta << " " << r << " " << minR << " " << maxR << std::endl;
  for (int i = 0; i < 9; ++i) file << c2w[i % 3][i / 3] << " ";     // c2w(i/3, i%3): row i/3, column i%3
  file << std::endl;
  file << screenW << " " << screenH << " " << screenDist << std::endl;
  file << focalDistance << " " << lensRadius << std::endl;
  return true;
}

bool Camera::load_settings(const std::string& filename) {
  std::ifstream file(filename);
  if (!file.is_open()) return false;
  file >> hFov >> vFov >> ar >> nClip >> fClip;
  for (int i = 0; i < 3; ++i) file >> pos[i];
  for (int i = 0; i < 3; ++i) file >> targetPos[i];
  file >> phi >> theta >> r >> minR >> maxR;
  for (int i = 0; i < 9; ++i) file >> c2w[i % 3][i / 3];
  file >> screenW >> screenH >> screenDist;
  file >> focalDistance >> lensRadius;
  return !file.fail();
}

rrt_camera Camera::view() const {
  rrt_camera c;
  for (int k = 0; k < 3; ++k) c.pos[k] = pos[k];
  for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) c.c2w[3 * j + i] = c2w[j][i];
  c.hFov = hFov; c.vFov = vFov; c.nClip = nClip; c.fClip = fClip; c.lensRadius = lensRadius; c.focalDistance = focalDistance;
  return c;
}

}  // namespace rrt_host
