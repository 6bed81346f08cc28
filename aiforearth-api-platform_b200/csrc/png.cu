// PNG payloads ("JPEG/PNG -> uint8 HWC").  PNG is lossless, so the whole decode is
// byte shuffling with no arithmetic to accelerate: chunk walk, zlib inflate of the IDAT stream and the
// five scan-line filters of the PNG specification (section 9) run on the SUBMITTING thread's host
// core, like the JPEG entropy decode, and the pixels then take the ordinary submit path (pinned slot
// -> H2D -> K1).  Output equals `Image.open(...).convert("RGB")` of Pillow 12.2 bit for bit: grey is
// replicated, palette entries are looked up, alpha is dropped.
//
// Supported: every image type of the PNG specification - colour types 0 (grey; 1, 2, 4, 8, 16 bits), 2 (RGB; 8, 16),
// 3 (palette; 1, 2, 4, 8), 4 (grey + alpha; 8, 16), 6 (RGBA; 8, 16), plain or Adam7-interlaced.  16-bit samples are
// reduced as Pillow reduces them (high byte; 16-bit grey without alpha is Pillow's "I;16", which convert("RGB") clips
// at 255).  Anything else is not a PNG -> AI4E_ERR_UNSUPPORTED (the shell answers 415).
// Reference call site: the decode inside the callback at APIs/1.0/base-py/ai4e_service.py:203; the
// content gate that admits image MIME types is ai4e_service.py:127-129.
#include "../../include/ai4e_b200.h"
#include "common.cuh"
#include "kernels.h"

#include <zlib.h>

#include <cstring>
#include <new>
#include <vector>

namespace ai4e {

namespace {

inline uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

inline int paeth(int a, int b, int c) {
  const int p = a + b - c;
  const int pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

}  // namespace

// Headers only when rgb == nullptr (sizes the output); otherwise decodes into rgb[height][width][3].
static int png_host_decode_impl(const uint8_t* data, uint64_t len, uint8_t* rgb, uint64_t capacity, int* out_h, int* out_w);

int png_host_decode(const uint8_t* data, uint64_t len, uint8_t* rgb, uint64_t capacity, int* out_h, int* out_w) {
  try {  // no exception may cross the C ABI: a payload that makes an allocation fail is an error, not a crash
    return png_host_decode_impl(data, len, rgb, capacity, out_h, out_w);
  } catch (const std::bad_alloc&) {
    set_last_error("png: out of memory while decoding");
    return AI4E_ERR_NOMEM;
  }
}

static int png_host_decode_impl(const uint8_t* data, uint64_t len, uint8_t* rgb, uint64_t capacity, int* out_h, int* out_w) {
  static const uint8_t kSig[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1a, '\n'};
  AI4E_REQUIRE(data && out_h && out_w, "png: null pointer");
  AI4E_REQUIRE(len >= 8 + 25 && memcmp(data, kSig, 8) == 0, "png: not a PNG stream (bad signature)");
  uint64_t pos = 8;
  int width = 0, height = 0, depth = 0, ctype = -1, interlace = 0;
  uint8_t palette[256 * 3];
  int pal_n = 0;
  std::vector<uint8_t> idat;
  bool have_ihdr = false, have_iend = false;
  while (pos + 12 <= len && !have_iend) {
    const uint32_t n = be32(data + pos);
    const uint8_t* type = data + pos + 4;
    AI4E_REQUIRE((uint64_t)n <= len - pos - 12, "png: chunk runs past the end of the stream");
    const uint8_t* body = data + pos + 8;
    if (!memcmp(type, "IHDR", 4)) {
      AI4E_REQUIRE(n == 13 && !have_ihdr, "png: bad IHDR");
      width = (int)be32(body); height = (int)be32(body + 4);
      depth = body[8]; ctype = body[9]; interlace = body[12];
      AI4E_REQUIRE(width > 0 && height > 0 && width <= 65535 && height <= 65535 && body[10] == 0 && body[11] == 0,
                   "png: bad IHDR fields");
      have_ihdr = true;
      // the combinations the PNG specification allows (table 11.1): grey 1/2/4/8/16, palette 1/2/4/8, the rest 8/16
      const bool ok = ctype == 0 ? (depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16)
                    : ctype == 3 ? (depth == 1 || depth == 2 || depth == 4 || depth == 8)
                    : (ctype == 2 || ctype == 4 || ctype == 6) ? (depth == 8 || depth == 16) : false;
      if (!ok || interlace > 1) {
        set_last_error("png: bit depth %d with colour type %d (interlace method %d) is not a PNG image type", depth, ctype, interlace);
        return AI4E_ERR_UNSUPPORTED;
      }
      *out_h = height; *out_w = width;
      if (!rgb) return AI4E_OK;
      AI4E_REQUIRE(capacity >= (uint64_t)width * height * 3, "png: output needs %llu bytes, capacity %llu",
                   (unsigned long long)width * height * 3, (unsigned long long)capacity);
    } else {
      AI4E_REQUIRE(have_ihdr, "png: chunk before IHDR");
      if (!memcmp(type, "PLTE", 4)) {
        AI4E_REQUIRE(n % 3 == 0 && n <= 768, "png: bad PLTE");
        memcpy(palette, body, n);
        pal_n = (int)(n / 3);
      } else if (!memcmp(type, "IDAT", 4)) {
        idat.insert(idat.end(), body, body + n);
      } else if (!memcmp(type, "IEND", 4)) {
        have_iend = true;
      }  // ancillary chunks (tRNS, gAMA, ...) do not change what convert("RGB") returns
    }
    pos += 12 + (uint64_t)n;
  }
  AI4E_REQUIRE(have_ihdr, "png: no IHDR");
  AI4E_REQUIRE(!idat.empty() && idat.size() < (1ull << 31), "png: no image data (or more than 2 GiB of it)");
  AI4E_REQUIRE(ctype != 3 || pal_n > 0, "png: palette image without PLTE");

  const int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : 4;
  const int bits = depth * channels;                       // bits per pixel
  const size_t fbpp = bits >= 8 ? (size_t)bits / 8 : 1;    // the filters' "corresponding byte of the pixel to the left"
  auto row_bytes = [&](int w) { return ((size_t)w * bits + 7) / 8; };
  // Adam7 (PNG specification section 8.2): seven reduced images, each a complete filtered image of its own
  static const int kX0[7] = {0, 4, 0, 2, 0, 1, 0}, kY0[7] = {0, 0, 4, 0, 2, 0, 1}, kDx[7] = {8, 8, 4, 4, 2, 2, 1}, kDy[7] = {8, 8, 8, 4, 4, 2, 2};
  const int n_pass = interlace ? 7 : 1;
  size_t total = 0;
  for (int ps = 0; ps < n_pass; ++ps) {
    const int pw = interlace ? (width - kX0[ps] + kDx[ps] - 1) / kDx[ps] : width;
    const int ph = interlace ? (height - kY0[ps] + kDy[ps] - 1) / kDy[ps] : height;
    if (pw > 0 && ph > 0) total += (row_bytes(pw) + 1) * (size_t)ph;
  }
  std::vector<uint8_t> raw(total);
  z_stream zs;
  memset(&zs, 0, sizeof(zs));
  AI4E_REQUIRE(inflateInit(&zs) == Z_OK, "png: zlib init failed");
  zs.next_in = idat.data();
  zs.avail_in = (uInt)idat.size();
  zs.next_out = raw.data();
  zs.avail_out = (uInt)raw.size();
  const int zrc = inflate(&zs, Z_FINISH);
  const size_t got = raw.size() - zs.avail_out;
  inflateEnd(&zs);
  AI4E_REQUIRE((zrc == Z_STREAM_END || zrc == Z_OK || zrc == Z_BUF_ERROR) && got == raw.size(),
               "png: corrupt or truncated image data (zlib %d, %zu of %zu bytes)", zrc, got, raw.size());

  // one pixel of an un-filtered line as Pillow's convert("RGB") renders it: grey replicated (1/2/4-bit grey scaled to
  // 0..255), palette entries looked up (an index past the palette reads zeros), alpha dropped, 16-bit samples reduced to
  // their high byte - except 16-bit grey WITHOUT alpha, which Pillow opens as "I;16" and converts by clipping at 255
  const int sub_mask = (1 << depth) - 1, sub_scale = depth < 8 ? 255 / sub_mask : 1;
  auto pixel = [&](const uint8_t* cur, int j, uint8_t* dst) {
    if (depth == 8) {
      switch (ctype) {
        case 2: dst[0] = cur[3 * j]; dst[1] = cur[3 * j + 1]; dst[2] = cur[3 * j + 2]; return;
        case 6: dst[0] = cur[4 * j]; dst[1] = cur[4 * j + 1]; dst[2] = cur[4 * j + 2]; return;
        case 0: dst[0] = dst[1] = dst[2] = cur[j]; return;
        case 4: dst[0] = dst[1] = dst[2] = cur[2 * j]; return;
        default: break;
      }
    } else if (depth == 16) {
      const uint8_t* q = cur + (size_t)j * channels * 2;
      if (ctype == 0) { dst[0] = dst[1] = dst[2] = q[0] ? 255 : q[1]; return; }
      if (ctype == 4) { dst[0] = dst[1] = dst[2] = q[0]; return; }
      dst[0] = q[0]; dst[1] = q[2]; dst[2] = q[4];
      return;
    }
    int v = cur[j];  // depth 8 palette, or a 1/2/4-bit sample (most significant bits first)
    if (depth < 8) {
      const int bit = j * depth;
      v = (cur[bit >> 3] >> (8 - depth - (bit & 7))) & sub_mask;
    }
    if (ctype == 0) { dst[0] = dst[1] = dst[2] = (uint8_t)(v * sub_scale); return; }
    if (v < pal_n) { dst[0] = palette[3 * v]; dst[1] = palette[3 * v + 1]; dst[2] = palette[3 * v + 2]; }
    else { dst[0] = dst[1] = dst[2] = 0; }
  };

  // ---- un-filter each (reduced) image in place (a line: filter byte + its bytes), then place its pixels
  std::vector<uint8_t> zero(row_bytes(width), 0);
  uint8_t* line = raw.data();
  for (int ps = 0; ps < n_pass; ++ps) {
    const int x0 = interlace ? kX0[ps] : 0, y0 = interlace ? kY0[ps] : 0, dx = interlace ? kDx[ps] : 1, dy = interlace ? kDy[ps] : 1;
    const int pw = (width - x0 + dx - 1) / dx, ph = (height - y0 + dy - 1) / dy;
    if (pw <= 0 || ph <= 0) continue;
    const size_t stride = row_bytes(pw);
    for (int r = 0; r < ph; ++r, line += stride + 1) {
      const int ft = line[0];
      uint8_t* cur = line + 1;
      const uint8_t* up = r ? cur - (stride + 1) : zero.data();
      switch (ft) {
        case 0: break;
        case 1: for (size_t i = fbpp; i < stride; ++i) cur[i] = (uint8_t)(cur[i] + cur[i - fbpp]); break;
        case 2: for (size_t i = 0; i < stride; ++i) cur[i] = (uint8_t)(cur[i] + up[i]); break;
        case 3:
          for (size_t i = 0; i < stride; ++i) cur[i] = (uint8_t)(cur[i] + (((i >= fbpp ? cur[i - fbpp] : 0) + up[i]) >> 1));
          break;
        case 4:
          for (size_t i = 0; i < stride; ++i) {
            const int a = i >= fbpp ? cur[i - fbpp] : 0, c = i >= fbpp ? up[i - fbpp] : 0;
            cur[i] = (uint8_t)(cur[i] + paeth(a, up[i], c));
          }
          break;
        default: set_last_error("png: unknown filter type %d on line %d of pass %d", ft, r, ps); return AI4E_ERR_INVALID;
      }
      uint8_t* dst = rgb + ((size_t)(y0 + r * dy) * width + x0) * 3;
      if (depth == 8 && ctype == 2 && dx == 1) {
        memcpy(dst, cur, (size_t)pw * 3);
      } else {
        for (int j = 0; j < pw; ++j) pixel(cur, j, dst + (size_t)j * dx * 3);
      }
    }
  }
  return AI4E_OK;
}

}  // namespace ai4e
