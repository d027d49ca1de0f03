// exr_probe.cpp — TEST INFRASTRUCTURE (oracle side).  Reads an OpenEXR file the way the reference's `-e map.exr` does — the calls of
// load_exr(), reference src/main.cpp:42-79, on the reference's vendored tinyexr (CGL/include/CGL/tinyexr.h, compiled from where it lies) —
// and dumps what the reference would put into its HDRImageBuffer: int32 width, int32 height, then width * height float RGB triples.
// Used by tests/golden/make_golden.py to pin host/envmap_io.cpp's PIZ reader against the reference's decoder.
// build: oracle/build_ref.sh  ->  oracle/_ref/exr_probe          usage: exr_probe in.exr out.bin
#define TINYEXR_IMPLEMENTATION
#include "CGL/tinyexr.h"
#include <cstdio>
#include <cstdint>
#include <vector>

int main(int argc, char** argv) {
  if (argc != 3) { fprintf(stderr, "usage: exr_probe in.exr out.bin\n"); return 2; }
  const char* err = nullptr;
  EXRImage exr;
  InitEXRImage(&exr);
  if (ParseMultiChannelEXRHeaderFromFile(&exr, argv[1], &err) != 0) { fprintf(stderr, "parse: %s\n", err ? err : "?"); return 1; }
  for (int i = 0; i < exr.num_channels; i++)
    if (exr.pixel_types[i] == TINYEXR_PIXELTYPE_HALF) exr.requested_pixel_types[i] = TINYEXR_PIXELTYPE_FLOAT;
  if (LoadMultiChannelEXRFromFile(&exr, argv[1], &err) != 0) { fprintf(stderr, "load: %s\n", err ? err : "?"); return 1; }
  if (exr.num_channels < 3) { fprintf(stderr, "fewer than 3 channels\n"); return 1; }
  const float* r = (const float*)exr.images[2];   // channels are stored alphabetically: B, G, R
  const float* g = (const float*)exr.images[1];
  const float* b = (const float*)exr.images[0];
  FILE* f = fopen(argv[2], "wb");
  if (!f) { fprintf(stderr, "cannot write %s\n", argv[2]); return 1; }
  int32_t wh[2] = {exr.width, exr.height};
  fwrite(wh, 4, 2, f);
  std::vector<float> row((size_t)exr.width * 3);
  for (int y = 0; y < exr.height; ++y) {
    for (int x = 0; x < exr.width; ++x) {
      const size_t i = (size_t)y * exr.width + x;
      row[3 * x] = r[i]; row[3 * x + 1] = g[i]; row[3 * x + 2] = b[i];
    }
    fwrite(row.data(), 4, row.size(), f);
  }
  fclose(f);
  return 0;
}
