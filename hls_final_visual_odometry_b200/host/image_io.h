// Frame readers for the host shim (SURVEY.md row F4, I/O adapters): the reference's raw 8-bit `.dat` frames
// (src/demo.cpp:82-110 reads width*height bytes per file) and KITTI-style PNG files, decoded on the host with zlib
// (the reference links libpng but never reads a PNG: CMakeLists.txt:32).  Header-only; link with -lz.
//
// PNG support: non-interlaced, bit depth 8 or 16, colour types 0 (grey: KITTI image_0 / image_1), 2 (RGB), 4 (grey + alpha),
// 6 (RGBA); colour is reduced to grey with the integer luma (77 R + 150 G + 29 B) >> 8, 16-bit samples keep their high byte.
// Anything else (palette, interlace, depth < 8) is refused with a message, never guessed.
#pragma once
#include <zlib.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace vo_io {

inline uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

inline bool read_file(const std::string& path, std::vector<uint8_t>& out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  const long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  out.resize(n > 0 ? (size_t)n : 0);
  const size_t got = out.empty() ? 0 : fread(out.data(), 1, out.size(), f);
  fclose(f);
  return got == out.size();
}

// PNG filter type 4 predictor
inline int paeth(int a, int b, int c) {
  const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// Decodes a PNG held in memory into 8-bit grey, row-major, tightly packed.  err receives a one-line reason on failure.
inline bool decode_png(const std::vector<uint8_t>& file, std::vector<uint8_t>& grey, int32_t* width, int32_t* height,
                       std::string* err) {
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  auto fail = [&](const char* m) { if (err) *err = m; return false; };
  if (file.size() < 8 + 25 || memcmp(file.data(), sig, 8) != 0) return fail("not a PNG file");
  uint32_t w = 0, h = 0;
  int depth = 0, ctype = 0;
  std::vector<uint8_t> idat;
  bool have_ihdr = false, have_iend = false;
  for (size_t pos = 8; pos + 12 <= file.size();) {
    const uint32_t len = be32(&file[pos]);
    const uint8_t* type = &file[pos + 4];
    if (pos + 12 + (size_t)len > file.size()) return fail("truncated PNG chunk");
    const uint8_t* data = &file[pos + 8];
    if (crc32(crc32(0L, type, 4), data, len) != be32(data + len)) return fail("PNG chunk CRC mismatch");
    if (!memcmp(type, "IHDR", 4)) {
      if (len != 13) return fail("bad IHDR");
      w = be32(data); h = be32(data + 4); depth = data[8]; ctype = data[9];
      if (data[10] != 0 || data[11] != 0) return fail("unknown PNG compression / filter method");
      if (data[12] != 0) return fail("interlaced PNG files are not supported");
      have_ihdr = true;
    } else if (!memcmp(type, "IDAT", 4)) {
      idat.insert(idat.end(), data, data + len);
    } else if (!memcmp(type, "IEND", 4)) {
      have_iend = true;
      break;
    }
    pos += 12 + (size_t)len;
  }
  if (!have_ihdr || !have_iend || idat.empty()) return fail("PNG without IHDR / IDAT / IEND");
  if (w == 0 || h == 0 || w > 32768 || h > 32768) return fail("PNG dimensions out of range");
  if (depth != 8 && depth != 16) return fail("PNG bit depth must be 8 or 16");
  int channels = 0;
  switch (ctype) {
    case 0: channels = 1; break;
    case 2: channels = 3; break;
    case 4: channels = 2; break;
    case 6: channels = 4; break;
    default: return fail("PNG colour type not supported (palette?)");
  }
  const size_t bpp = (size_t)channels * depth / 8, stride = bpp * w;
  std::vector<uint8_t> raw((stride + 1) * h);
  uLongf raw_len = (uLongf)raw.size();
  if (uncompress(raw.data(), &raw_len, idat.data(), (uLong)idat.size()) != Z_OK || raw_len != raw.size())
    return fail("PNG image data does not inflate to the declared size");
  // undo the per-row filters in place (rows are raw[y * (stride + 1) + 1 ...])
  std::vector<uint8_t> zero(stride, 0);
  for (uint32_t y = 0; y < h; y++) {
    uint8_t* row = &raw[(size_t)y * (stride + 1) + 1];
    const uint8_t* up = y ? row - (stride + 1) : zero.data();
    const int ft = row[-1];
    for (size_t x = 0; x < stride; x++) {
      const int a = x >= bpp ? row[x - bpp] : 0, b = up[x], c = x >= bpp ? up[x - bpp] : 0;
      int pred = 0;
      switch (ft) {
        case 0: pred = 0; break;
        case 1: pred = a; break;
        case 2: pred = b; break;
        case 3: pred = (a + b) >> 1; break;
        case 4: pred = paeth(a, b, c); break;
        default: return fail("unknown PNG row filter");
      }
      row[x] = (uint8_t)(row[x] + pred);
    }
  }
  grey.resize((size_t)w * h);
  const size_t step = depth / 8;  // 16-bit samples are big-endian: the first byte is the high one
  for (uint32_t y = 0; y < h; y++) {
    const uint8_t* row = &raw[(size_t)y * (stride + 1) + 1];
    uint8_t* dst = &grey[(size_t)y * w];
    if (channels <= 2) {
      for (uint32_t x = 0; x < w; x++) dst[x] = row[x * bpp];
    } else {
      for (uint32_t x = 0; x < w; x++) {
        const uint8_t* p = row + x * bpp;
        dst[x] = (uint8_t)((77 * p[0] + 150 * p[step] + 29 * p[2 * step]) >> 8);
      }
    }
  }
  *width = (int32_t)w; *height = (int32_t)h;
  return true;
}

// Reads one frame.  `.png` (any case) -> decode_png; anything else -> raw 8-bit frame of exactly width*height bytes as
// src/demo.cpp:96-110 reads it (width / height must then be given).  On success *width / *height hold the frame size.
inline bool read_frame(const std::string& path, std::vector<uint8_t>& grey, int32_t* width, int32_t* height, std::string* err) {
  std::vector<uint8_t> file;
  if (!read_file(path, file)) { if (err) *err = "cannot read " + path; return false; }
  const size_t n = path.size();
  const bool png = n >= 4 && path[n - 4] == '.' && (path[n - 3] | 32) == 'p' && (path[n - 2] | 32) == 'n' && (path[n - 1] | 32) == 'g';
  if (png) return decode_png(file, grey, width, height, err);
  if (*width <= 0 || *height <= 0 || file.size() != (size_t)*width * (size_t)*height) {
    if (err) *err = "raw frame " + path + " does not hold width*height bytes";
    return false;
  }
  grey.swap(file);
  return true;
}

}  // namespace vo_io
