// PNG payloads ("JPEG/PNG -> uint8 HWC").  PNG is lossless, so the whole decode is
// byte shuffling with no arithmetic to accelerate: chunk walk, zlib inflate of the IDAT stream and the
// five scan-line filters of the PNG specification (section 9) run on the SUBMITTING thread's host
// core, like the JPEG entropy decode, and the pixels then take the ordinary submit path (pinned slot
// -> H2D -> K1).  Output equals `Image.open(...).convert("RGB")` of Pillow bit for bit: grey is
// replicated, palette entries are looked up, alpha is dropped.
//
// Supported: bit depth 8, colour types 0 (grey), 2 (RGB), 3 (palette), 4 (grey + alpha), 6 (RGBA),
// non-interlaced.  16-bit, 1/2/4-bit and Adam7 files -> AI4E_ERR_UNSUPPORTED (the shell answers 415).
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
      const bool known = ctype == 0 || ctype == 2 || ctype == 3 || ctype == 4 || ctype == 6;
      if (depth != 8 || !known || interlace != 0) {
        set_last_error("png: bit depth %d / colour type %d / interlace %d is not supported (8-bit non-interlaced grey, RGB, "
                       "palette, with or without alpha, are)", depth, ctype, interlace);
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

  const int bpp = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : 4;  // bytes per pixel
  const size_t stride = (size_t)width * bpp;
  std::vector<uint8_t> raw((stride + 1) * (size_t)height);
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

  // ---- un-filter in place (each line: filter byte + stride bytes), then expand to RGB
  std::vector<uint8_t> zero(stride, 0);
  for (int y = 0; y < height; ++y) {
    uint8_t* line = raw.data() + (size_t)y * (stride + 1);
    const int ft = line[0];
    uint8_t* cur = line + 1;
    const uint8_t* up = y ? cur - (stride + 1) : zero.data();
    switch (ft) {
      case 0: break;
      case 1: for (size_t i = bpp; i < stride; ++i) cur[i] = (uint8_t)(cur[i] + cur[i - bpp]); break;
      case 2: for (size_t i = 0; i < stride; ++i) cur[i] = (uint8_t)(cur[i] + up[i]); break;
      case 3:
        for (size_t i = 0; i < stride; ++i) cur[i] = (uint8_t)(cur[i] + (((i >= (size_t)bpp ? cur[i - bpp] : 0) + up[i]) >> 1));
        break;
      case 4:
        for (size_t i = 0; i < stride; ++i) {
          const int a = i >= (size_t)bpp ? cur[i - bpp] : 0, c = i >= (size_t)bpp ? up[i - bpp] : 0;
          cur[i] = (uint8_t)(cur[i] + paeth(a, up[i], c));
        }
        break;
      default: set_last_error("png: unknown filter type %d on line %d", ft, y); return AI4E_ERR_INVALID;
    }
    uint8_t* dst = rgb + (size_t)y * width * 3;
    switch (ctype) {
      case 2: memcpy(dst, cur, (size_t)width * 3); break;
      case 6: for (int x = 0; x < width; ++x) { dst[3 * x] = cur[4 * x]; dst[3 * x + 1] = cur[4 * x + 1]; dst[3 * x + 2] = cur[4 * x + 2]; } break;
      case 0: for (int x = 0; x < width; ++x) dst[3 * x] = dst[3 * x + 1] = dst[3 * x + 2] = cur[x]; break;
      case 4: for (int x = 0; x < width; ++x) dst[3 * x] = dst[3 * x + 1] = dst[3 * x + 2] = cur[2 * x]; break;
      default:
        for (int x = 0; x < width; ++x) {
          const int k = cur[x] < pal_n ? cur[x] : 0;  // an index past the palette reads entry 0 (Pillow pads the palette with zeros; so does this when pal_n < 256)
          if (cur[x] < pal_n) { dst[3 * x] = palette[3 * k]; dst[3 * x + 1] = palette[3 * k + 1]; dst[3 * x + 2] = palette[3 * k + 2]; }
          else { dst[3 * x] = dst[3 * x + 1] = dst[3 * x + 2] = 0; }
        }
    }
  }
  return AI4E_OK;
}

}  // namespace ai4e
