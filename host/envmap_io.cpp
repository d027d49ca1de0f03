// envmap_io.cpp — environment maps for -e (reference src/main.cpp:42-79 reads them through tinyexr).
// Two containers:
//   * OpenEXR scanline files with compression NONE, RLE, ZIPS or ZIP (zlib) and HALF / FLOAT / UINT channels R, G, B; the wavelet
//     and lossy codecs (PIZ, PXR24, B44, DWA) and tiled / deep / multi-part files are reported, not guessed at;
//   * PFM ("PF", little- or big-endian, bottom row first).
// Output: [h][w][3] floats, row 0 = top, the layout rrt_set_envmap takes.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>

#include <zlib.h>

#include "rrt_host.h"

namespace rrt_host {
namespace {

float half_to_float(uint16_t h) {
  uint32_t sign = (uint32_t)(h >> 15) << 31, exp = (h >> 10) & 0x1F, man = h & 0x3FF, bits;
  if (exp == 0) {
    if (man == 0) bits = sign;
    else {   // subnormal half: renormalise
      int e = -1;
      do { ++e; man <<= 1; } while (!(man & 0x400));
      bits = sign | (uint32_t)(127 - 15 - e) << 23 | (man & 0x3FF) << 13;
    }
  } else if (exp == 31) bits = sign | 0x7F800000u | man << 13;
  else bits = sign | (exp + 127 - 15) << 23 | man << 13;
  float f; memcpy(&f, &bits, 4);
  return f;
}

bool read_file(const std::string& path, std::vector<uint8_t>* out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
  out->resize(n > 0 ? (size_t)n : 0);
  bool ok = n >= 0 && fread(out->data(), 1, out->size(), f) == out->size();
  fclose(f);
  return ok;
}

bool load_pfm(const std::vector<uint8_t>& b, int* w, int* h, std::vector<float>* rgb, std::string* err) {
  int channels = b[1] == 'F' ? 3 : 1;
  size_t pos = 2;
  auto token = [&]() {
    while (pos < b.size() && (b[pos] == ' ' || b[pos] == '\n' || b[pos] == '\r' || b[pos] == '\t')) ++pos;
    size_t s = pos;
    while (pos < b.size() && !(b[pos] == ' ' || b[pos] == '\n' || b[pos] == '\r' || b[pos] == '\t')) ++pos;
    return std::string((const char*)b.data() + s, pos - s);
  };
  *w = atoi(token().c_str()); *h = atoi(token().c_str());
  double scale = atof(token().c_str());
  ++pos;   // the single whitespace after the scale
  if (*w <= 0 || *h <= 0 || scale == 0) { *err = "malformed PFM header"; return false; }
  size_t need = (size_t)*w * *h * channels * 4;
  if (b.size() < pos + need) { *err = "truncated PFM data"; return false; }
  const bool big = scale > 0;
  rgb->assign((size_t)*w * *h * 3, 0.f);
  for (int y = 0; y < *h; ++y)
    for (int x = 0; x < *w; ++x)
      for (int c = 0; c < 3; ++c) {
        const uint8_t* q = b.data() + pos + 4 * (((size_t)(*h - 1 - y) * *w + x) * channels + (channels == 3 ? c : 0));
        uint8_t t[4] = {q[0], q[1], q[2], q[3]};
        if (big) { t[0] = q[3]; t[1] = q[2]; t[2] = q[1]; t[3] = q[0]; }
        float f; memcpy(&f, t, 4);
        (*rgb)[((size_t)y * *w + x) * 3 + c] = f;
      }
  return true;
}

// OpenEXR's byte shuffling behind RLE / ZIPS / ZIP (ImfZip.cpp, ImfRle.cpp): the block is split into even and odd bytes, then
// delta-coded; decoding undoes the delta (t[i] += t[i-1] - 128) and re-interleaves the two halves.
void exr_unpredict_and_interleave(std::vector<uint8_t>& t, std::vector<uint8_t>* out) {
  for (size_t i = 1; i < t.size(); ++i) t[i] = (uint8_t)(t[i - 1] + t[i] - 128);
  out->resize(t.size());
  const size_t half = (t.size() + 1) / 2;
  for (size_t i = 0, a = 0, b = half; i < t.size(); ++i) (*out)[i] = (i & 1) ? t[b++] : t[a++];
}
bool exr_rle_decode(const uint8_t* in, size_t n, size_t expect, std::vector<uint8_t>* out) {
  out->clear(); out->reserve(expect);
  size_t i = 0;
  while (i < n) {
    const int8_t c = (int8_t)in[i++];
    if (c < 0) {
      const size_t k = (size_t)(-(int)c);
      if (i + k > n || out->size() + k > expect) return false;
      out->insert(out->end(), in + i, in + i + k); i += k;
    } else {
      const size_t k = (size_t)c + 1;
      if (i >= n || out->size() + k > expect) return false;
      out->insert(out->end(), k, in[i++]);
    }
  }
  return out->size() == expect;
}

struct Channel { std::string name; int type; };   // 0 uint, 1 half, 2 float

bool load_exr(const std::vector<uint8_t>& b, int* w, int* h, std::vector<float>* rgb, std::string* err) {
  size_t pos = 8;
  uint32_t version; memcpy(&version, b.data() + 4, 4);
  if ((version & 0xFF) != 2) { *err = "unsupported OpenEXR version"; return false; }
  if (version & 0x1E00) { *err = "tiled / deep / multi-part OpenEXR files are not supported (scanline only)"; return false; }
  std::vector<Channel> chans;
  int compression = -1, dw[4] = {0, 0, -1, -1};
  auto cstr = [&](std::string* s) {
    size_t e = pos;
    while (e < b.size() && b[e]) ++e;
    if (e >= b.size()) return false;
    s->assign((const char*)b.data() + pos, e - pos); pos = e + 1;
    return true;
  };
  for (;;) {
    std::string name, type;
    if (!cstr(&name)) { *err = "truncated OpenEXR header"; return false; }
    if (name.empty()) break;
    if (!cstr(&type) || pos + 4 > b.size()) { *err = "truncated OpenEXR header"; return false; }
    int32_t size; memcpy(&size, b.data() + pos, 4); pos += 4;
    if (size < 0 || pos + (size_t)size > b.size()) { *err = "truncated OpenEXR attribute"; return false; }
    const uint8_t* v = b.data() + pos;
    if (name == "channels") {
      size_t p = 0;
      while (p < (size_t)size && v[p]) {
        size_t e = p;
        while (e < (size_t)size && v[e]) ++e;                    // the name must end inside the attribute
        if (e >= (size_t)size || e + 1 + 16 > (size_t)size) { *err = "truncated OpenEXR channel list"; return false; }
        Channel c; c.name.assign((const char*)v + p, e - p); p = e + 1;
        int32_t t, xs, ys; memcpy(&t, v + p, 4); memcpy(&xs, v + p + 8, 4); memcpy(&ys, v + p + 12, 4);
        c.type = t; p += 16;   // pixel type, pLinear + reserved, xSampling, ySampling
        if (xs != 1 || ys != 1) { *err = "sub-sampled OpenEXR channels are not supported"; return false; }
        chans.push_back(c);
      }
    } else if (name == "compression") { if (size < 1) { *err = "truncated OpenEXR attribute"; return false; } compression = v[0]; }
    else if (name == "dataWindow") { if (size < 16) { *err = "truncated OpenEXR attribute"; return false; } memcpy(dw, v, 16); }
    pos += (size_t)size;
  }
  if (compression < 0 || compression > 3) {
    static const char* names[] = {"NONE", "RLE", "ZIPS", "ZIP", "PIZ", "PXR24", "B44", "B44A", "DWAA", "DWAB"};
    *err = std::string("OpenEXR compression ") + (compression >= 0 && compression < 10 ? names[compression] : "(unknown)") +
           " is not supported (NONE, RLE, ZIPS and ZIP are): re-save the map with one of those, or as .pfm";
    return false;
  }
  const int64_t w64 = (int64_t)dw[2] - dw[0] + 1, h64 = (int64_t)dw[3] - dw[1] + 1;
  if (w64 <= 0 || h64 <= 0 || chans.empty()) { *err = "OpenEXR header without channels / data window"; return false; }
  if (w64 > 65536 || h64 > 65536 || w64 * h64 >= ((int64_t)1 << 31)) { *err = "OpenEXR data window too large (rrt_set_envmap takes width * height < 2^31)"; return false; }
  *w = (int)w64; *h = (int)h64;
  std::map<std::string, int> slot = {{"R", 0}, {"G", 1}, {"B", 2}};
  size_t row_bytes = 0;
  for (const Channel& c : chans) {
    if (c.type < 0 || c.type > 2) { *err = "unknown OpenEXR pixel type"; return false; }
    row_bytes += (size_t)*w * (c.type == 1 ? 2 : 4);
  }
  const int lines_per_chunk = compression == 3 ? 16 : 1;
  const int n_chunks = (*h + lines_per_chunk - 1) / lines_per_chunk;
  if (pos + 8 * (size_t)n_chunks > b.size()) { *err = "truncated OpenEXR offset table"; return false; }
  rgb->assign((size_t)*w * *h * 3, 0.f);
  std::vector<uint8_t> tmp, block;
  for (int chunk = 0; chunk < n_chunks; ++chunk) {
    uint64_t off; memcpy(&off, b.data() + pos + 8 * (size_t)chunk, 8);
    if (off + 8 > b.size()) { *err = "truncated OpenEXR chunk"; return false; }
    int32_t y, size; memcpy(&y, b.data() + off, 4); memcpy(&size, b.data() + off + 4, 4);
    const int row0 = y - dw[1];
    if (row0 < 0 || row0 >= *h) { *err = "OpenEXR scan line outside the data window"; return false; }
    const int lines = std::min(lines_per_chunk, *h - row0);
    const size_t expect = row_bytes * (size_t)lines;
    if (size < 0 || off + 8 + (size_t)size > b.size()) { *err = "truncated OpenEXR chunk"; return false; }
    const uint8_t* q = b.data() + off + 8;
    if (compression != 0 && (size_t)size < expect) {      // a chunk that did not shrink is stored raw
      if (compression == 1) {
        if (!exr_rle_decode(q, (size_t)size, expect, &tmp)) { *err = "corrupt RLE data in OpenEXR chunk"; return false; }
      } else {
        tmp.resize(expect);
        uLongf got = (uLongf)expect;
        if (uncompress(tmp.data(), &got, q, (uLong)size) != Z_OK || got != expect) { *err = "corrupt zlib data in OpenEXR chunk"; return false; }
      }
      exr_unpredict_and_interleave(tmp, &block);
      q = block.data();
    } else if ((size_t)size != expect) { *err = "OpenEXR chunk of unexpected size"; return false; }
    for (int l = 0; l < lines; ++l) {
      const int row = row0 + l;
      for (const Channel& c : chans) {    // per scan line, channels in the header's (alphabetical) order
        auto it = slot.find(c.name);
        for (int x = 0; x < *w; ++x) {
          float f;
          if (c.type == 1) { uint16_t hv; memcpy(&hv, q, 2); q += 2; f = half_to_float(hv); }
          else if (c.type == 2) { memcpy(&f, q, 4); q += 4; }
          else { uint32_t u; memcpy(&u, q, 4); q += 4; f = (float)u; }
          if (it != slot.end()) (*rgb)[((size_t)row * *w + x) * 3 + it->second] = f;
        }
      }
    }
  }
  return true;
}

}  // namespace

bool load_envmap(const std::string& filename, int* width, int* height, std::vector<float>* rgb, std::string* error) {
  std::vector<uint8_t> b;
  std::string err;
  if (!read_file(filename, &b)) { if (error) *error = "cannot read " + filename; return false; }
  bool ok;
  if (b.size() >= 8 && b[0] == 0x76 && b[1] == 0x2f && b[2] == 0x31 && b[3] == 0x01) ok = load_exr(b, width, height, rgb, &err);
  else if (b.size() >= 3 && b[0] == 'P' && (b[1] == 'F' || b[1] == 'f')) ok = load_pfm(b, width, height, rgb, &err);
  else { err = "neither an OpenEXR nor a PFM file"; ok = false; }
  if (!ok && error) *error = filename + ": " + err;
  return ok;
}

}  // namespace rrt_host
