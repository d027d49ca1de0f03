// envmap_io.cpp — environment maps for -e (reference src/main.cpp:42-79 reads them through tinyexr).
// Two containers:
//   * OpenEXR scanline files with compression NONE, RLE, ZIPS, ZIP (zlib) or PIZ (wavelet + Huffman, below) and HALF / FLOAT / UINT channels
//     R, G, B — a superset of what the reference's tinyexr reads (NONE, ZIPS, ZIP, PIZ); the lossy codecs (PXR24, B44, DWA) and tiled / deep /
//     multi-part files are reported, not guessed at;
//   * PFM ("PF", little- or big-endian, bottom row first).
// Output: [h][w][3] floats, row 0 = top, the layout rrt_set_envmap takes.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <memory>

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
  if (*w <= 0 || *h <= 0 || scale == 0 || scale != scale) { *err = "malformed PFM header"; return false; }
  if (*w > 65536 || *h > 65536 || (int64_t)*w * *h >= ((int64_t)1 << 31)) { *err = "PFM map too large (rrt_set_envmap takes width * height < 2^31)"; return false; }
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

// ---------------------------------------------------------------------------------------------------------------- PIZ chunks
// The codec most HDR probes are saved with, and the one the reference's tinyexr reads beside ZIP (reference CGL/include/CGL/tinyexr.h,
// DecompressPiz).  As OpenEXR publishes it (ImfPizCompressor, ImfHuf, ImfWav): the 16-bit words of a chunk of up to 32 scan lines — per channel
// a plane of lines x width x (1 | 2) words — are renumbered densely (a bitmap says which of the 65536 values occur), every plane is
// Haar-wavelet transformed, and all words are Huffman coded as one sequence with a run marker.  Decoding runs the three steps backwards.
// A chunk is untrusted input: every read goes through a bounded cursor, every table index is checked.
namespace piz {

struct Bits {   // MSB-first bit cursor over [p, p + nbytes), never past `limit` bits
  const uint8_t* p; size_t nbytes; uint64_t pos, limit;
  uint64_t left() const { return limit - pos; }
  uint64_t peek(int n) const {   // the next n bits, 1 <= n <= 57, zero-padded past the last byte
    const size_t byte = (size_t)(pos >> 3);
    uint64_t acc = 0;
    for (size_t k = 0; k < 8; ++k) acc = (acc << 8) | (byte + k < nbytes ? p[byte + k] : 0);
    acc <<= (pos & 7);           // drop what has been read of the first byte: at least 57 bits remain at the top
    return acc >> (64 - n);
  }
  void skip(int n) { pos += (uint64_t)n; }
};

const int kMaxLen = 58, kSymbols = 65537, kFast = 12;

struct Code {
  uint8_t len[kSymbols];
  uint64_t base[kMaxLen + 1]; uint32_t count[kMaxLen + 1], offset[kMaxLen + 1];
  std::vector<uint32_t> syms;                       // symbols in (length, symbol) order: canonical codes count up along it
  std::vector<uint32_t> fast;                       // kFast-bit prefix -> symbol << 8 | length, 0 = longer code (or none)
};

// code lengths of symbols im .. iM, 6 bits each; 59 .. 62 = 2 .. 5 zero lengths, 63 + 8 bits = 6 .. 261 zero lengths
bool read_lengths(Bits* in, uint32_t im, uint32_t iM, Code* c) {
  memset(c->len, 0, sizeof c->len);
  for (uint32_t s = im; s <= iM;) {
    if (in->left() < 6) return false;
    const uint32_t l = (uint32_t)in->peek(6); in->skip(6);
    if (l >= 59) {
      uint32_t run = l - 59 + 2;
      if (l == 63) { if (in->left() < 8) return false; run = (uint32_t)in->peek(8) + 6; in->skip(8); }
      if (s + run > iM + 1) return false;
      s += run;
    } else c->len[s++] = (uint8_t)l;
  }
  return true;
}
// canonical assignment: the longest codes take the numerically smallest values, and within a length codes count up with the symbol
bool build_code(Code* c) {
  memset(c->count, 0, sizeof c->count);
  for (int s = 0; s < kSymbols; ++s) c->count[c->len[s]]++;
  c->count[0] = 0;
  uint64_t next = 0; uint32_t off = 0;
  for (int l = kMaxLen; l >= 1; --l) { c->base[l] = next; next = (next + c->count[l]) >> 1; }
  for (int l = 1; l <= kMaxLen; ++l) {
    c->offset[l] = off; off += c->count[l];
    if (c->count[l] && c->base[l] + c->count[l] > ((uint64_t)1 << l)) return false;   // more codes than l bits hold
  }
  if (off < 2) return false;                         // at least one value and the run marker
  c->syms.assign(off, 0);
  std::vector<uint32_t> fill(c->offset, c->offset + kMaxLen + 1);
  for (int s = 0; s < kSymbols; ++s) if (c->len[s]) c->syms[fill[c->len[s]]++] = (uint32_t)s;
  c->fast.assign((size_t)1 << kFast, 0);
  for (int l = 1; l <= kFast; ++l)
    for (uint32_t k = 0; k < c->count[l]; ++k) {
      const uint64_t code = c->base[l] + k;
      const size_t lo = (size_t)(code << (kFast - l)), hi = (size_t)((code + 1) << (kFast - l));
      for (size_t e = lo; e < hi; ++e) {
        if (c->fast[e]) return false;                // two codes share a prefix: not a prefix code
        c->fast[e] = (c->syms[c->offset[l] + k] << 8) | (uint32_t)l;
      }
    }
  return true;
}
bool next_symbol(const Code& c, Bits* in, uint32_t* sym) {
  const uint32_t f = c.fast[(size_t)in->peek(kFast)];
  if (f) {
    const int l = (int)(f & 0xFF);
    if ((uint64_t)l > in->left()) return false;
    in->skip(l); *sym = f >> 8;
    return true;
  }
  const uint64_t window = in->peek(57);              // codes longer than 57 bits would need 2^57 symbols below them
  for (int l = kFast + 1; l <= 57; ++l) {
    if (!c.count[l]) continue;
    const uint64_t code = window >> (57 - l);
    if (code >= c.base[l] && code - c.base[l] < c.count[l]) {
      if ((uint64_t)l > in->left()) return false;
      in->skip(l); *sym = c.syms[c.offset[l] + (size_t)(code - c.base[l])];
      return true;
    }
  }
  return false;
}
// the Huffman stream of a chunk -> n words
bool huffman_decode(const uint8_t* src, size_t size, uint16_t* out, size_t n, Code* code) {
  if (n == 0) return true;
  if (size < 20) return false;
  uint32_t im, iM, nbits;
  memcpy(&im, src, 4); memcpy(&iM, src + 4, 4); memcpy(&nbits, src + 12, 4);
  if (im >= (uint32_t)kSymbols || iM >= (uint32_t)kSymbols || im > iM) return false;
  Bits tb; tb.p = src + 20; tb.nbytes = size - 20; tb.pos = 0; tb.limit = 8 * (uint64_t)(size - 20);
  if (!read_lengths(&tb, im, iM, code)) return false;
  const size_t table_bytes = (size_t)((tb.pos + 7) >> 3);   // the data starts at the next byte
  if (!build_code(code) || !code->len[iM]) return false;
  Bits in; in.p = src + 20 + table_bytes; in.nbytes = size - 20 - table_bytes; in.pos = 0; in.limit = nbits;
  if (in.limit > 8 * (uint64_t)in.nbytes) return false;
  size_t k = 0;
  while (k < n) {
    uint32_t s;
    if (!next_symbol(*code, &in, &s)) return false;
    if (s == iM) {                                    // run marker: the previous word, `count` more times
      if (in.left() < 8 || k == 0) return false;
      const size_t run = (size_t)in.peek(8); in.skip(8);
      if (k + run > n) return false;
      for (size_t r = 0; r < run; ++r, ++k) out[k] = out[k - 1];
    } else {
      if (s > 0xFFFF) return false;
      out[k++] = (uint16_t)s;
    }
  }
  return true;
}

// inverse Haar steps: (average, difference) -> the two values.  14-bit form: plain integers; 16-bit form: modulo 2^16
inline void pair14(uint16_t l, uint16_t h, uint16_t* a, uint16_t* b) {
  const int ls = (int16_t)l, hs = (int16_t)h;
  const int ai = ls + (hs & 1) + (hs >> 1);
  *a = (uint16_t)ai; *b = (uint16_t)(ai - hs);
}
inline void pair16(uint16_t l, uint16_t h, uint16_t* a, uint16_t* b) {
  const int bb = ((int)l - ((int)h >> 1)) & 0xFFFF;
  *b = (uint16_t)bb; *a = (uint16_t)(((int)h + bb - 0x8000) & 0xFFFF);
}
// one plane: element (x, y) at a[x * ox + y * oy]; levels from the coarsest (largest power of two <= min(nx, ny)) down to 1
template <void (*PAIR)(uint16_t, uint16_t, uint16_t*, uint16_t*)>
void unwavelet(uint16_t* a, int nx, size_t ox, int ny, size_t oy) {
  int p = 1;
  while (p <= std::min(nx, ny)) p <<= 1;
  p >>= 1;
  for (int p2 = p, q = p >> 1; q >= 1; p2 = q, q >>= 1) {
    int y = 0;
    for (; y + p2 <= ny; y += p2) {
      int x = 0;
      for (; x + p2 <= nx; x += p2) {
        uint16_t* v00 = a + x * ox + y * oy; uint16_t* v01 = v00 + q * ox; uint16_t* v10 = v00 + q * oy; uint16_t* v11 = v10 + q * ox;
        uint16_t i00, i01, i10, i11;
        PAIR(*v00, *v10, &i00, &i10); PAIR(*v01, *v11, &i01, &i11);
        PAIR(i00, i01, v00, v01); PAIR(i10, i11, v10, v11);
      }
      if (nx & q) {                                   // a last column without a partner in x
        uint16_t* v00 = a + x * ox + y * oy; uint16_t* v10 = v00 + q * oy;
        uint16_t i00;
        PAIR(*v00, *v10, &i00, v10); *v00 = i00;
      }
    }
    if (ny & q)                                       // a last line without a partner in y
      for (int x = 0; x + p2 <= nx; x += p2) {
        uint16_t* v00 = a + x * ox + y * oy; uint16_t* v01 = v00 + q * ox;
        uint16_t i00;
        PAIR(*v00, *v01, &i00, v01); *v00 = i00;
      }
  }
}

// one chunk -> `expect` bytes in the uncompressed chunk layout (per scan line, per channel, the channel's row)
bool decode_chunk(const uint8_t* src, size_t size, const std::vector<Channel>& chans, int w, int lines, size_t expect, Code* code,
                  std::vector<uint8_t>* out, std::string* err) {
  const size_t nwords = expect / 2;
  if (size < 4) { *err = "truncated PIZ chunk"; return false; }
  uint16_t lo, hi;
  memcpy(&lo, src, 2); memcpy(&hi, src + 2, 2);
  size_t pos = 4;
  std::vector<uint8_t> bitmap(8192, 0);
  if (hi >= 8192) { *err = "corrupt PIZ chunk (value bitmap)"; return false; }
  if (lo <= hi) {
    const size_t nb = (size_t)hi - lo + 1;
    if (pos + nb > size) { *err = "truncated PIZ chunk"; return false; }
    memcpy(bitmap.data() + lo, src + pos, nb); pos += nb;
  }
  std::vector<uint16_t> value(65536, 0);             // dense number -> value; zero is always number 0
  uint32_t n_values = 0;
  for (uint32_t v = 0; v < 65536; ++v)
    if (v == 0 || (bitmap[v >> 3] & (1u << (v & 7)))) value[n_values++] = (uint16_t)v;
  const uint32_t max_number = n_values - 1;
  if (pos + 4 > size) { *err = "truncated PIZ chunk"; return false; }
  int32_t hlen; memcpy(&hlen, src + pos, 4); pos += 4;
  if (hlen < 0 || pos + (size_t)hlen > size) { *err = "truncated PIZ chunk"; return false; }
  std::vector<uint16_t> words(nwords);
  if (!huffman_decode(src + pos, (size_t)hlen, words.data(), nwords, code)) { *err = "corrupt Huffman data in PIZ chunk"; return false; }
  size_t start = 0;
  for (const Channel& c : chans) {
    const size_t sz = c.type == 1 ? 1 : 2;
    for (size_t j = 0; j < sz; ++j) {
      if (max_number < (1u << 14)) unwavelet<pair14>(words.data() + start + j, w, sz, lines, (size_t)w * sz);
      else unwavelet<pair16>(words.data() + start + j, w, sz, lines, (size_t)w * sz);
    }
    start += (size_t)w * lines * sz;
  }
  for (uint16_t& x : words) x = value[x];            // numbers past the last one map to 0, as in the format's own table
  out->resize(expect);
  uint8_t* q = out->data();
  std::vector<size_t> at(chans.size());
  start = 0;
  for (size_t k = 0; k < chans.size(); ++k) { at[k] = start; start += (size_t)w * lines * (chans[k].type == 1 ? 1 : 2); }
  for (int l = 0; l < lines; ++l)
    for (size_t k = 0; k < chans.size(); ++k) {
      const size_t n = (size_t)w * (chans[k].type == 1 ? 1 : 2);
      memcpy(q, words.data() + at[k], 2 * n); q += 2 * n; at[k] += n;
    }
  return true;
}

}  // namespace piz

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
  if (compression < 0 || compression > 4) {
    static const char* names[] = {"NONE", "RLE", "ZIPS", "ZIP", "PIZ", "PXR24", "B44", "B44A", "DWAA", "DWAB"};
    *err = std::string("OpenEXR compression ") + (compression >= 0 && compression < 10 ? names[compression] : "(unknown)") +
           " is not supported (NONE, RLE, ZIPS, ZIP and PIZ are): re-save the map with one of those, or as .pfm";
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
  const int lines_per_chunk = compression == 4 ? 32 : compression == 3 ? 16 : 1;
  std::unique_ptr<piz::Code> piz_code(compression == 4 ? new piz::Code() : nullptr);
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
    // a chunk that did not shrink is stored raw (same size); a PIZ chunk LARGER than its data can only be codec output (writers without the
    // raw fall-back exist, and the reference's reader decodes whatever it finds)
    if (compression != 0 && ((size_t)size < expect || (compression == 4 && (size_t)size > expect))) {
      if (compression == 4) {
        if (!piz::decode_chunk(q, (size_t)size, chans, *w, lines, expect, piz_code.get(), &block, err)) return false;
        q = block.data();
      } else if (compression == 1) {
        if (!exr_rle_decode(q, (size_t)size, expect, &tmp)) { *err = "corrupt RLE data in OpenEXR chunk"; return false; }
      } else {
        tmp.resize(expect);
        uLongf got = (uLongf)expect;
        if (uncompress(tmp.data(), &got, q, (uLong)size) != Z_OK || got != expect) { *err = "corrupt zlib data in OpenEXR chunk"; return false; }
      }
      if (compression != 4) { exr_unpredict_and_interleave(tmp, &block); q = block.data(); }
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
