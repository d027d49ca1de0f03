// png_write.cpp — 8-bit RGBA PNG with stored (uncompressed) deflate blocks.  The reference writes its frames through
// lodepng (src/pathtracer.cpp:676); the image has no compression library, and the pixels are what matters.
#include <stdio.h>

#include <vector>

#include "rrt_host.h"

namespace rrt_host {
namespace {
uint32_t crc_table[256];
bool crc_ready = false;
uint32_t crc32(const uint8_t* p, size_t n, uint32_t c = 0xFFFFFFFFu) {
  if (!crc_ready) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t v = i;
      for (int k = 0; k < 8; ++k) v = (v & 1) ? 0xEDB88320u ^ (v >> 1) : v >> 1;
      crc_table[i] = v;
    }
    crc_ready = true;
  }
  for (size_t i = 0; i < n; ++i) c = crc_table[(c ^ p[i]) & 0xFF] ^ (c >> 8);
  return c;
}
void be32(std::vector<uint8_t>& o, uint32_t v) { o.push_back(v >> 24); o.push_back(v >> 16); o.push_back(v >> 8); o.push_back(v); }
void chunk(std::vector<uint8_t>& out, const char* tag, const std::vector<uint8_t>& data) {
  be32(out, (uint32_t)data.size());
  size_t start = out.size();
  out.insert(out.end(), tag, tag + 4);
  out.insert(out.end(), data.begin(), data.end());
  be32(out, crc32(out.data() + start, out.size() - start) ^ 0xFFFFFFFFu);
}
}  // namespace

bool write_png(const std::string& filename, const uint32_t* rgba, size_t w, size_t h) {
  std::vector<uint8_t> raw;
  raw.reserve(h * (4 * w + 1));
  for (size_t y = 0; y < h; ++y) {
    raw.push_back(0);   // filter: none
    const uint8_t* row = (const uint8_t*)(rgba + y * w);   // little-endian 0xAABBGGRR == bytes R,G,B,A
    raw.insert(raw.end(), row, row + 4 * w);
  }
  std::vector<uint8_t> z;
  z.push_back(0x78); z.push_back(0x01);
  uint32_t a = 1, b = 0;
  for (uint8_t v : raw) { a = (a + v) % 65521u; b = (b + a) % 65521u; }
  size_t pos = 0;
  do {
    size_t n = raw.size() - pos < 65535 ? raw.size() - pos : 65535;
    z.push_back(pos + n == raw.size() ? 1 : 0);
    z.push_back(n & 0xFF); z.push_back(n >> 8); z.push_back(~n & 0xFF); z.push_back((~n >> 8) & 0xFF);
    z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
    pos += n;
  } while (pos < raw.size());
  be32(z, (b << 16) | a);
  std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'};
  std::vector<uint8_t> ihdr;
  be32(ihdr, (uint32_t)w); be32(ihdr, (uint32_t)h);
  ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
  chunk(out, "IHDR", ihdr);
  chunk(out, "IDAT", z);
  chunk(out, "IEND", std::vector<uint8_t>());
  FILE* f = fopen(filename.c_str(), "wb");
  if (!f) return false;
  bool ok = fwrite(out.data(), 1, out.size(), f) == out.size();
  return (fclose(f) == 0) && ok;
}

}  // namespace rrt_host
