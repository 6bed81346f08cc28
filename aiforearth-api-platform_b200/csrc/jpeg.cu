// K0 - JPEG decode (baseline and progressive), hybrid: entropy decoding on the host, everything else on the GPU.
//
// Requests carry encoded images (the reference gates on image MIME types,
// APIs/1.0/base-py/ai4e_service.py:72,127-129, and leaves decoding to the callback it calls at
// :203 - Pillow / libjpeg-turbo in a PyTorch container).  This file matches libjpeg's output bit
// for bit: Huffman decoding per ITU T.81 F.2.2 (progressive scans per annex G, after jdphuff.c), the accurate integer IDCT (jidctint.c "islow",
// the library default), fancy triangle-filter chroma upsampling (jdsample.c h2v1 / h2v2, edge
// rows and columns replicated as jdmainct.c does) and the fixed-point YCbCr -> RGB tables of
// jdcolor.c.  Oracle: oracle/jpeg_oracle.py, pinned to Pillow on tests/golden/jpeg/.
//
// Why hybrid: a Huffman bit stream is sequential (one symbol's length decides where the next
// starts); a 224 x 224 image is ~25 KB of it, ~0.15 ms on one host core, and the host shell
// already runs one thread per request, so the entropy decode parallelises over requests for free
// and overlaps the GPU.  What the GPU gets per image is 150 KB of int16 coefficients (the same
// bytes as the raw RGB image) and 27 k independent 8-point IDCTs plus a per-pixel filter: two small
// HBM-bound kernels in front of K1.
#include "../../include/ai4e_b200.h"
#include "common.cuh"
#include "kernels.h"

#include <cstring>
#include <vector>

namespace ai4e {

// zigzag position -> natural (row-major) index
static const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                    41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                    30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

namespace {

constexpr int kLookBits = 9;

struct HuffTable {
  bool present = false;
  // canonical decoding (T.81 F.2.2.3) for codes longer than kLookBits
  int32_t mincode[17], maxcode[18], valptr[17];
  uint8_t symbols[256];
  // first kLookBits bits -> (length << 8) | symbol, 0 = longer code
  uint16_t look[1 << kLookBits];
};

bool build_table(HuffTable& t, const uint8_t* counts, const uint8_t* syms, int nsyms) {
  if (nsyms > 256) return false;
  memcpy(t.symbols, syms, nsyms);
  memset(t.look, 0, sizeof(t.look));
  int code = 0, k = 0;
  for (int len = 1; len <= 16; ++len) {
    t.valptr[len] = k;
    t.mincode[len] = code;
    // over-subscribed code lengths would index past the look-up table below: refuse them first
    if (code + counts[len - 1] > (1 << len)) return false;
    for (int i = 0; i < counts[len - 1]; ++i, ++k, ++code) {
      if (len <= kLookBits) {
        const int first = code << (kLookBits - len);
        for (int j = 0; j < (1 << (kLookBits - len)); ++j) t.look[first + j] = (uint16_t)((len << 8) | syms[k]);
      }
    }
    t.maxcode[len] = counts[len - 1] ? code - 1 : -1;
    code <<= 1;
  }
  t.maxcode[17] = 0x7fffffff;
  t.present = true;
  return true;
}

// Bit reader over the entropy-coded segment: removes the stuffed 0x00 after 0xFF, stops at markers
// (feeding zero bits from there on, like libjpeg).
struct BitReader {
  const uint8_t* p;
  const uint8_t* end;
  uint64_t acc = 0;
  int n = 0;
  bool hit_marker = false;

  void fill() {
    while (n <= 56) {
      uint32_t b = 0;
      if (!hit_marker && p < end) {
        b = *p;
        if (b == 0xFF) {
          if (p + 1 < end && p[1] == 0x00) {
            p += 2;
          } else {
            hit_marker = true;  // RSTn / EOI / truncated: leave p on the 0xFF
            b = 0;
          }
        } else {
          ++p;
        }
      } else {
        hit_marker = true;
      }
      acc |= (uint64_t)b << (56 - n);
      n += 8;
    }
  }
  inline uint32_t peek(int k) { return (uint32_t)(acc >> (64 - k)); }
  inline void skip(int k) { acc <<= k; n -= k; }
  inline uint32_t get(int k) {
    const uint32_t v = peek(k);
    skip(k);
    return v;
  }
  void reset_bits() { acc = 0; n = 0; hit_marker = false; }
};

inline int decode_symbol(BitReader& br, const HuffTable& t) {
  if (br.n < 16) br.fill();
  const uint16_t e = t.look[br.peek(kLookBits)];
  if (e) {
    br.skip(e >> 8);
    return e & 0xFF;
  }
  int len = kLookBits + 1;
  int code = (int)br.peek(len);
  while (len <= 16 && code > t.maxcode[len]) {
    ++len;
    code = (int)br.peek(len);
  }
  if (len > 16) return -1;
  br.skip(len);
  return t.symbols[(t.valptr[len] + code - t.mincode[len]) & 0xFF];
}

inline int extend(int v, int s) { return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v; }

inline int rd16(const uint8_t* p) { return (p[0] << 8) | p[1]; }

inline uint32_t getbits(BitReader& br, int k) {
  if (k == 0) return 0;
  if (br.n < k) br.fill();
  return br.get(k);
}

struct ScanSpec {
  int ns = 0;
  int comp[3] = {0, 0, 0};  // frame component index of each scan component
  int ss = 0, se = 63, ah = 0, al = 0;
  bool progressive = false;
};

// One scan into the record.  Sequential scans (SOF0 / SOF1) carry all 64 coefficients of their components;
// progressive scans (SOF2) are one of the four kinds of ITU T.81 annex G, decoded as libjpeg's jdphuff.c
// does (DC first / refinement, AC first pass with end-of-band runs, AC refinement with correction bits).
// A scan of one component walks that component's own blocks in raster order (without the padding to whole
// MCUs); a scan of several walks the frame's MCUs.  Leaves br.p at (or before) the marker that ends it.
template <bool kProgressive>  // (the sequential loop, what nearly every file takes, compiled without the progressive branches)
int decode_scan_t(BitReader& br, const ai4e_jpeg_info& f, const ScanSpec& sc, const int* td, const int* ta,
                  const std::vector<HuffTable>& dc, const std::vector<HuffTable>& ac, int16_t* record) {
  const int mcus_x = f.blocks_w[0] / f.h[0], mcus_y = f.blocks_h[0] / f.v[0];
  const bool inter = sc.ns > 1;
  const int c0 = sc.comp[0];
  const int bw0 = (f.comp_w[c0] + 7) / 8, bh0 = (f.comp_h[c0] + 7) / 8;
  const int total = inter ? mcus_x * mcus_y : bw0 * bh0;
  const int ri = f.restart_interval;
  const int p1 = 1 << sc.al, m1 = -(1 << sc.al);
  int pred[3] = {0, 0, 0};
  int eobrun = 0;
  int until_restart = ri ? ri : total;
  for (int mcu = 0; mcu < total; ++mcu) {
    if (until_restart == 0) {  // expect RSTn: byte-align, skip the marker, reset the predictors and the run
      br.fill();
      const uint8_t* q = br.p;
      // br.p sits on the 0xFF of the marker once the reader has hit it; otherwise search forward
      while (q + 1 < br.end && !(q[0] == 0xFF && q[1] >= 0xD0 && q[1] <= 0xD7)) ++q;
      AI4E_REQUIRE(q + 1 < br.end, "jpeg: restart marker missing");
      br.p = q + 2;
      br.reset_bits();
      pred[0] = pred[1] = pred[2] = 0;
      eobrun = 0;
      until_restart = ri;
    }
    --until_restart;
    const int my = inter ? mcu / mcus_x : 0, mx = inter ? mcu - my * mcus_x : 0;
    for (int si = 0; si < sc.ns; ++si) {
      const int c = sc.comp[si];
      const int nv = inter ? f.v[c] : 1, nh = inter ? f.h[c] : 1;
      for (int by = 0; by < nv; ++by) {
        for (int bx = 0; bx < nh; ++bx) {
          const int64_t brow = inter ? my * f.v[c] + by : mcu / bw0, bcol = inter ? mx * f.h[c] + bx : mcu % bw0;
          int16_t* blk = record + f.coef_off[c] + (brow * f.blocks_w[c] + bcol) * 64;
          if (!kProgressive) {
            int s = decode_symbol(br, dc[td[c]]);
            AI4E_REQUIRE(s >= 0 && s <= 11, "jpeg: corrupt DC code");
            if (s) pred[c] += extend((int)getbits(br, s), s);
            blk[0] = (int16_t)pred[c];
            for (int k = 1; k < 64;) {
              const int rs = decode_symbol(br, ac[ta[c]]);
              AI4E_REQUIRE(rs >= 0, "jpeg: corrupt AC code");
              const int r = rs >> 4;
              s = rs & 15;
              if (s == 0) {
                if (r != 15) break;
                k += 16;
                continue;
              }
              k += r;
              AI4E_REQUIRE(k < 64, "jpeg: AC coefficient index past 63");
              blk[kZigzag[k]] = (int16_t)extend((int)getbits(br, s), s);
              ++k;
            }
          } else if (sc.ss == 0) {
            if (sc.ah == 0) {  // DC, first pass
              const int s = decode_symbol(br, dc[td[c]]);
              AI4E_REQUIRE(s >= 0 && s <= 11, "jpeg: corrupt DC code");
              if (s) pred[c] += extend((int)getbits(br, s), s);
              blk[0] = (int16_t)(pred[c] * p1);
            } else if (getbits(br, 1)) {  // DC refinement: one more bit
              blk[0] = (int16_t)(blk[0] | p1);
            }
          } else if (sc.ah == 0) {  // AC, first pass
            if (eobrun > 0) { --eobrun; continue; }
            for (int k = sc.ss; k <= sc.se; ++k) {
              const int rs = decode_symbol(br, ac[ta[c]]);
              AI4E_REQUIRE(rs >= 0, "jpeg: corrupt AC code");
              const int r = rs >> 4, s = rs & 15;
              if (s) {
                k += r;
                AI4E_REQUIRE(k < 64, "jpeg: AC coefficient index past 63");
                blk[kZigzag[k]] = (int16_t)(extend((int)getbits(br, s), s) * p1);
              } else if (r == 15) {
                k += 15;  // ZRL: sixteen zeros (the loop adds the last)
              } else {
                eobrun = (1 << r) + (int)getbits(br, r) - 1;
                break;
              }
            }
          } else {  // AC refinement
            int k = sc.ss;
            if (eobrun == 0) {
              for (; k <= sc.se; ++k) {
                const int rs = decode_symbol(br, ac[ta[c]]);
                AI4E_REQUIRE(rs >= 0, "jpeg: corrupt AC code");
                int r = rs >> 4, s = rs & 15;
                if (s) {
                  s = getbits(br, 1) ? p1 : m1;  // the new coefficient's sign (its magnitude is 1 << al)
                } else if (r != 15) {
                  eobrun = (1 << r) + (int)getbits(br, r);
                  break;  // end of band: the rest of this block gets correction bits only
                }
                // over already-nonzero coefficients (one correction bit each) and r still-zero ones
                while (k <= sc.se) {
                  int16_t& co = blk[kZigzag[k]];
                  if (co != 0) {
                    if (getbits(br, 1) && (co & p1) == 0) co = (int16_t)(co + (co >= 0 ? p1 : m1));
                  } else if (--r < 0) {
                    break;
                  }
                  ++k;
                }
                if (s) {
                  AI4E_REQUIRE(k < 64, "jpeg: AC coefficient index past 63");
                  blk[kZigzag[k]] = (int16_t)s;
                }
              }
            }
            if (eobrun > 0) {
              for (; k <= sc.se; ++k) {
                int16_t& co = blk[kZigzag[k]];
                if (co != 0 && getbits(br, 1) && (co & p1) == 0) co = (int16_t)(co + (co >= 0 ? p1 : m1));
              }
              --eobrun;
            }
          }
        }
      }
    }
  }
  return AI4E_OK;
}

int decode_scan(BitReader& br, const ai4e_jpeg_info& f, const ScanSpec& sc, const int* td, const int* ta,
                const std::vector<HuffTable>& dc, const std::vector<HuffTable>& ac, int16_t* record) {
  return sc.progressive ? decode_scan_t<true>(br, f, sc, td, ta, dc, ac, record) : decode_scan_t<false>(br, f, sc, td, ta, dc, ac, record);
}

}  // namespace

// headers (+ entropy decode when `record` is given).  Returns AI4E_OK or an error with a message.
int jpeg_host_decode(const uint8_t* data, uint64_t len, int16_t* record, uint64_t capacity, ai4e_jpeg_info* out) {
  AI4E_REQUIRE(data && out, "jpeg: null pointer");
  AI4E_REQUIRE(len >= 4 && data[0] == 0xFF && data[1] == 0xD8, "jpeg: not a JPEG stream (no SOI marker)");
  memset(out, 0, sizeof(*out));
  uint16_t qt[4][64];
  bool have_qt[4] = {false, false, false, false};
  std::vector<HuffTable> dc(4), ac(4);
  int tq[3] = {0, 0, 0}, cid[3] = {0, 0, 0}, td[3] = {0, 0, 0}, ta[3] = {0, 0, 0};
  bool have_frame = false, have_scan = false, progressive = false, done = false;
  bool latched[3] = {false, false, false};   // quantisation table copied into the record (at the component's first scan)
  bool seen[3] = {false, false, false};      // sequential files: the component's scan has been decoded
  int scans = 0;
  uint64_t pos = 2;
  while (pos + 4 <= len && !done) {
    AI4E_REQUIRE(data[pos] == 0xFF, "jpeg: marker expected at byte %llu", (unsigned long long)pos);
    while (pos < len && data[pos] == 0xFF) ++pos;
    AI4E_REQUIRE(pos < len, "jpeg: truncated");
    const int m = data[pos++];
    if (m == 0xD9) break;
    if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
    AI4E_REQUIRE(pos + 2 <= len, "jpeg: truncated");
    const int seg_len = rd16(data + pos);
    AI4E_REQUIRE(seg_len >= 2 && pos + seg_len <= len, "jpeg: bad segment length");
    const uint8_t* seg = data + pos + 2;
    const int n = seg_len - 2;
    pos += seg_len;
    if (m == 0xDB) {
      for (int i = 0; i < n;) {
        const int pq = seg[i] >> 4, t = seg[i] & 15;
        ++i;
        AI4E_REQUIRE(t < 4 && pq < 2 && i + 64 * (pq + 1) <= n, "jpeg: bad quantisation table");
        for (int k = 0; k < 64; ++k) {
          qt[t][kZigzag[k]] = pq ? (uint16_t)rd16(seg + i) : seg[i];
          i += pq + 1;
        }
        have_qt[t] = true;
      }
    } else if (m == 0xC4) {
      for (int i = 0; i < n;) {
        AI4E_REQUIRE(i + 17 <= n, "jpeg: bad Huffman table");
        const int tc = seg[i] >> 4, th = seg[i] & 15;
        int ns = 0;
        for (int k = 0; k < 16; ++k) ns += seg[i + 1 + k];
        AI4E_REQUIRE(tc < 2 && th < 4 && i + 17 + ns <= n, "jpeg: bad Huffman table");
        AI4E_REQUIRE(build_table(tc ? ac[th] : dc[th], seg + i + 1, seg + i + 17, ns), "jpeg: invalid Huffman code lengths");
        i += 17 + ns;
      }
    } else if (m == 0xC0 || m == 0xC1 || m == 0xC2) {
      AI4E_REQUIRE(!have_frame, "jpeg: second frame header");
      progressive = m == 0xC2;
      AI4E_REQUIRE(n >= 6 && seg[0] == 8, "jpeg: only 8-bit samples are supported");
      out->height = rd16(seg + 1);
      out->width = rd16(seg + 3);
      out->components = seg[5];
      if (out->components != 1 && out->components != 3) {
        set_last_error("jpeg: %d components are not supported", out->components);
        return AI4E_ERR_UNSUPPORTED;
      }
      AI4E_REQUIRE(out->width > 0 && out->height > 0 && n >= 6 + 3 * out->components, "jpeg: bad frame header");
      for (int c = 0; c < out->components; ++c) {
        cid[c] = seg[6 + 3 * c];
        out->h[c] = seg[7 + 3 * c] >> 4;
        out->v[c] = seg[7 + 3 * c] & 15;
        tq[c] = seg[8 + 3 * c];
        AI4E_REQUIRE(tq[c] < 4 && out->h[c] >= 1 && out->v[c] >= 1, "jpeg: bad component");
      }
      if (out->components == 1) out->h[0] = out->v[0] = 1;
      out->hmax = out->vmax = 1;
      for (int c = 0; c < out->components; ++c) {
        out->hmax = std::max(out->hmax, out->h[c]);
        out->vmax = std::max(out->vmax, out->v[c]);
      }
      for (int c = 0; c < out->components; ++c) {
        const int hr = out->hmax / out->h[c], vr = out->vmax / out->v[c];
        const bool ok = out->hmax % out->h[c] == 0 && out->vmax % out->v[c] == 0 && out->h[c] <= 4 && out->v[c] <= 2 &&
                        (hr == 1 || hr == 2 || hr == 4) && (vr == 1 || vr == 2) && (c > 0 || (hr == 1 && vr == 1));
        if (!ok) {
          set_last_error("jpeg: sampling %dx%d of component %d (max %dx%d) is not supported (horizontal ratios 1, 2, 4 and vertical "
                         "ratios 1, 2 against a full-resolution first component are: 4:4:4, 4:2:2, 4:2:0, 4:4:0, 4:1:1, 4:1:0)",
                         out->h[c], out->v[c], c, out->hmax, out->vmax);
          return AI4E_ERR_UNSUPPORTED;
        }
      }
      const int mcus_x = (out->width + 8 * out->hmax - 1) / (8 * out->hmax);
      const int mcus_y = (out->height + 8 * out->vmax - 1) / (8 * out->vmax);
      int64_t off = 3 * 64, plane = 0;
      for (int c = 0; c < out->components; ++c) {
        out->blocks_w[c] = mcus_x * out->h[c];
        out->blocks_h[c] = mcus_y * out->v[c];
        out->comp_w[c] = (out->width * out->h[c] + out->hmax - 1) / out->hmax;
        out->comp_h[c] = (out->height * out->v[c] + out->vmax - 1) / out->vmax;
        out->coef_off[c] = off;
        off += (int64_t)out->blocks_w[c] * out->blocks_h[c] * 64;
        plane += (int64_t)out->blocks_w[c] * out->blocks_h[c] * 64;
      }
      out->record_int16 = (off + 63) & ~(int64_t)63;
      out->plane_bytes = (plane + 255) & ~(int64_t)255;
      have_frame = true;
    } else if (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
      set_last_error("jpeg: SOF%d (lossless / hierarchical / arithmetic coding) is not supported, only Huffman-coded "
                     "sequential and progressive JPEG", m - 0xC0);
      return AI4E_ERR_UNSUPPORTED;
    } else if (m == 0xDD) {
      AI4E_REQUIRE(n >= 2, "jpeg: bad DRI");
      out->restart_interval = rd16(seg);
    } else if (m == 0xDA) {
      AI4E_REQUIRE(have_frame, "jpeg: SOS before SOF");
      AI4E_REQUIRE(n >= 1 && seg[0] >= 1 && seg[0] <= out->components && n >= 1 + 2 * seg[0] + 3, "jpeg: bad scan header");
      ScanSpec sc;
      sc.ns = seg[0];
      sc.progressive = progressive;
      bool named[3] = {false, false, false};
      for (int i = 0; i < sc.ns; ++i) {
        int c = -1;
        for (int j = 0; j < out->components; ++j)
          if (cid[j] == seg[1 + 2 * i]) c = j;
        AI4E_REQUIRE(c >= 0, "jpeg: scan names an unknown component");
        AI4E_REQUIRE(!named[c], "jpeg: scan names component %d twice", c);
        named[c] = true;
        sc.comp[i] = c;
        td[c] = seg[2 + 2 * i] >> 4;
        ta[c] = seg[2 + 2 * i] & 15;
        AI4E_REQUIRE(td[c] < 4 && ta[c] < 4, "jpeg: bad Huffman table selector");
      }
      sc.ss = seg[1 + 2 * sc.ns]; sc.se = seg[2 + 2 * sc.ns];
      sc.ah = seg[3 + 2 * sc.ns] >> 4; sc.al = seg[3 + 2 * sc.ns] & 15;
      if (progressive) {
        AI4E_REQUIRE(sc.ss <= sc.se && sc.se <= 63 && sc.ah <= 13 && sc.al <= 13 && (sc.ss > 0 || sc.se == 0) &&
                     (sc.ss == 0 || sc.ns == 1), "jpeg: bad progressive scan parameters (Ss %d, Se %d, Ah %d, Al %d, %d components)",
                     sc.ss, sc.se, sc.ah, sc.al, sc.ns);
      } else {
        AI4E_REQUIRE(sc.ss == 0 && sc.se == 63 && sc.ah == 0 && sc.al == 0, "jpeg: sequential scan with a spectral selection");
        AI4E_REQUIRE(sc.ns == out->components || sc.ns == 1, "jpeg: sequential scan of %d of %d components is not supported",
                     sc.ns, out->components);
      }
      for (int i = 0; i < sc.ns; ++i) {  // the tables this kind of scan reads must exist
        const int c = sc.comp[i];
        const bool need_dc = !progressive || (sc.ss == 0 && sc.ah == 0), need_ac = !progressive || sc.ss > 0;
        AI4E_REQUIRE((!need_dc || dc[td[c]].present) && (!need_ac || ac[ta[c]].present), "jpeg: scan uses an undefined Huffman table");
      }
      have_scan = true;
      if (!record) break;  // headers only: the first scan header is as far as the caller needs
      if (scans == 0) {
        AI4E_REQUIRE(capacity >= (uint64_t)out->record_int16, "jpeg: record needs %lld int16, capacity %llu",
                     (long long)out->record_int16, (unsigned long long)capacity);
        memset(record, 0, (size_t)out->record_int16 * sizeof(int16_t));
      }
      for (int i = 0; i < sc.ns; ++i) {  // a component's quantisation table is the one in force at its first scan
        const int c = sc.comp[i];
        if (latched[c]) continue;
        AI4E_REQUIRE(have_qt[tq[c]], "jpeg: component %d uses an undefined quantisation table", c);
        for (int k = 0; k < 64; ++k) record[c * 64 + k] = (int16_t)qt[tq[c]][k];
        latched[c] = true;
      }
      BitReader br;
      br.p = data + pos;
      br.end = data + len;
      const int rc = decode_scan(br, *out, sc, td, ta, dc, ac, record);
      if (rc) return rc;
      ++scans;
      AI4E_REQUIRE(scans <= 256, "jpeg: too many scans");
      // the next marker: the reader stopped on it if it ran into it, otherwise it lies ahead
      const uint8_t* q = br.p;
      while (q + 1 < br.end && !(q[0] == 0xFF && q[1] != 0x00 && q[1] != 0xFF && !(q[1] >= 0xD0 && q[1] <= 0xD7))) ++q;
      pos = (uint64_t)(q - data);
      if (!progressive) {
        bool all = true;
        for (int i = 0; i < sc.ns; ++i) seen[sc.comp[i]] = true;
        for (int c = 0; c < out->components; ++c) all = all && seen[c];
        done = all;  // sequential: every component has had its scan
      }
    }
  }
  AI4E_REQUIRE(have_frame, "jpeg: no frame header");
  if (!record) return AI4E_OK;
  AI4E_REQUIRE(have_scan && scans > 0, "jpeg: no scan");
  for (int c = 0; c < out->components; ++c) AI4E_REQUIRE(latched[c], "jpeg: component %d has no scan", c);
  if (out->components == 1)  // the record always carries three table slots
    for (int c = 1; c < 3; ++c) memcpy(record + c * 64, record, 64 * sizeof(int16_t));
  return AI4E_OK;
}


// ----------------------------------------------------------------------------------------
// device side
// ----------------------------------------------------------------------------------------
struct JpegGeom {
  int comps, width, height;
  int blocks_w[3], blocks_h[3], comp_w[3], comp_h[3];
  int hr[3], vr[3];          // upsampling ratio of each component
  int block_begin[4];        // prefix sums of blocks per component
  long long coef_off[3], record_int16;
  long long plane_off[3], plane_bytes;
};

// jidctint.c constants (CONST_BITS = 13, PASS1_BITS = 2)
#define JFIX_0_298631336 2446
#define JFIX_0_390180644 3196
#define JFIX_0_541196100 4433
#define JFIX_0_765366865 6270
#define JFIX_0_899976223 7373
#define JFIX_1_175875602 9633
#define JFIX_1_501321110 12299
#define JFIX_1_847759065 15137
#define JFIX_1_961570560 16069
#define JFIX_2_053119869 16819
#define JFIX_2_562915447 20995
#define JFIX_3_072711026 25172

// one 8-point pass of jpeg_idct_islow, in place, results descaled by `shift`
template <int kShift>
__device__ __forceinline__ void idct_pass(int& i0, int& i1, int& i2, int& i3, int& i4, int& i5, int& i6, int& i7) {
  int z1 = (i2 + i6) * JFIX_0_541196100;
  const int tmp2 = z1 + i6 * (-JFIX_1_847759065);
  const int tmp3 = z1 + i2 * JFIX_0_765366865;
  const int tmp0 = (i0 + i4) << 13, tmp1 = (i0 - i4) << 13;
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  int t0 = i7, t1 = i5, t2 = i3, t3 = i1;
  z1 = t0 + t3;
  int z2 = t1 + t2, z3 = t0 + t2, z4 = t1 + t3;
  const int z5 = (z3 + z4) * JFIX_1_175875602;
  t0 *= JFIX_0_298631336; t1 *= JFIX_2_053119869; t2 *= JFIX_3_072711026; t3 *= JFIX_1_501321110;
  z1 *= -JFIX_0_899976223; z2 *= -JFIX_2_562915447;
  z3 = z3 * (-JFIX_1_961570560) + z5;
  z4 = z4 * (-JFIX_0_390180644) + z5;
  t0 += z1 + z3; t1 += z2 + z4; t2 += z2 + z3; t3 += z1 + z4;
  constexpr int kRound = 1 << (kShift - 1);
  i0 = (tmp10 + t3 + kRound) >> kShift; i7 = (tmp10 - t3 + kRound) >> kShift;
  i1 = (tmp11 + t2 + kRound) >> kShift; i6 = (tmp11 - t2 + kRound) >> kShift;
  i2 = (tmp12 + t1 + kRound) >> kShift; i5 = (tmp12 - t1 + kRound) >> kShift;
  i3 = (tmp13 + t0 + kRound) >> kShift; i4 = (tmp13 - t0 + kRound) >> kShift;
}

// range_limit[x & 1023] of jdmaster.c's table (centred on 128)
__device__ __forceinline__ uint32_t range_limit(int x) {
  const int i = x & 1023;
  return i < 128 ? i + 128 : (i < 512 ? 255 : (i < 896 ? 0 : i - 896));
}

// one thread = one 8x8 block: dequantise, IDCT, write 8 rows of 8 samples into the component plane
__global__ void __launch_bounds__(128)
jpeg_idct_kernel(const int16_t* __restrict__ records, uint8_t* __restrict__ planes, const JpegGeom g, int batch) {
  const int total = g.block_begin[g.comps];
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= (long long)batch * total) return;
  const int img = (int)(idx / total);
  int b = (int)(idx - (long long)img * total);
  int c = 0;
  while (c + 1 < g.comps && b >= g.block_begin[c + 1]) ++c;
  b -= g.block_begin[c];
  const int by = b / g.blocks_w[c], bx = b - by * g.blocks_w[c];
  const int16_t* rec = records + (long long)img * g.record_int16;
  const uint4* src = reinterpret_cast<const uint4*>(rec + g.coef_off[c] + (long long)b * 64);
  const uint4* qsrc = reinterpret_cast<const uint4*>(rec + c * 64);
  int v[64];
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    const uint4 cw = __ldg(src + r), qw = __ldg(qsrc + r);
    const uint32_t cu[4] = {cw.x, cw.y, cw.z, cw.w}, qu[4] = {qw.x, qw.y, qw.z, qw.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      v[r * 8 + 2 * k] = (int)(int16_t)(cu[k] & 0xFFFF) * (int)(qu[k] & 0xFFFF);
      v[r * 8 + 2 * k + 1] = (int)(int16_t)(cu[k] >> 16) * (int)(qu[k] >> 16);
    }
  }
#pragma unroll
  for (int col = 0; col < 8; ++col)  // pass 1: columns, scaled up by 2^PASS1_BITS
    idct_pass<13 - 2>(v[col], v[8 + col], v[16 + col], v[24 + col], v[32 + col], v[40 + col], v[48 + col], v[56 + col]);
  uint8_t* dst = planes + (long long)img * g.plane_bytes + g.plane_off[c] + ((long long)by * 8 * g.blocks_w[c] + bx) * 8;
#pragma unroll
  for (int r = 0; r < 8; ++r) {  // pass 2: rows
    idct_pass<13 + 2 + 3>(v[r * 8], v[r * 8 + 1], v[r * 8 + 2], v[r * 8 + 3], v[r * 8 + 4], v[r * 8 + 5], v[r * 8 + 6], v[r * 8 + 7]);
    uint2 o;
    o.x = range_limit(v[r * 8]) | (range_limit(v[r * 8 + 1]) << 8) | (range_limit(v[r * 8 + 2]) << 16) | (range_limit(v[r * 8 + 3]) << 24);
    o.y = range_limit(v[r * 8 + 4]) | (range_limit(v[r * 8 + 5]) << 8) | (range_limit(v[r * 8 + 6]) << 16) | (range_limit(v[r * 8 + 7]) << 24);
    *reinterpret_cast<uint2*>(dst + (long long)r * g.blocks_w[c] * 8) = o;
  }
}

// sample of component c at full-resolution pixel (y, x): the upsampler jdsample.c's jinit_upsampler picks - the triangle
// filters for 2:1 horizontally (alone or with 2:1 vertically) when the component is more than two samples wide and for
// 2:1 vertically alone (libjpeg-turbo's h1v2), plain replication for everything else (ratio 4, narrow components)
__device__ __forceinline__ int jpeg_sample(const uint8_t* __restrict__ pl, const JpegGeom& g, int c, int y, int x) {
  const int stride = g.blocks_w[c] * 8;
  const int hr = g.hr[c], vr = g.vr[c];
  if (hr == 1 && vr == 1) return pl[(long long)y * stride + x];
  const int cw = g.comp_w[c];
  if (hr == 1) {  // h1v2: 3/4 nearer row + 1/4 next nearest; bias 1 (upper) / 2 (lower); edge rows stand in for their neighbours
    const int ch = g.comp_h[c], cy = y >> 1;
    const int cyn = (y & 1) ? min(cy + 1, ch - 1) : max(cy - 1, 0);
    return (3 * pl[(long long)cy * stride + x] + pl[(long long)cyn * stride + x] + ((y & 1) ? 2 : 1)) >> 2;
  }
  if (hr != 2 || cw <= 2) return pl[(long long)(vr == 2 ? y >> 1 : y) * stride + (hr == 4 ? x >> 2 : x >> 1)];
  const int cx = x >> 1;
  if (vr == 1) {  // h2v1: 3/4 nearer + 1/4 further; rounding 1 (even) / 2 (odd); edge columns copied
    const uint8_t* row = pl + (long long)y * stride;
    const int p = row[cx];
    if (x & 1) return cx == cw - 1 ? p : (3 * p + row[cx + 1] + 2) >> 2;
    return cx == 0 ? p : (3 * p + row[cx - 1] + 1) >> 2;
  }
  // h2v2: 9/16 3/16 3/16 1/16; the row above the first / below the last is that row itself
  const int ch = g.comp_h[c], cy = y >> 1;
  const int cyn = (y & 1) ? min(cy + 1, ch - 1) : max(cy - 1, 0);
  const uint8_t* r0 = pl + (long long)cy * stride;
  const uint8_t* r1 = pl + (long long)cyn * stride;
  const int cur = 3 * r0[cx] + r1[cx];
  if (x & 1) {
    if (cx == cw - 1) return (cur * 4 + 7) >> 4;
    return (3 * cur + (3 * r0[cx + 1] + r1[cx + 1]) + 7) >> 4;
  }
  if (cx == 0) return (cur * 4 + 8) >> 4;
  return (3 * cur + (3 * r0[cx - 1] + r1[cx - 1]) + 8) >> 4;
}

struct JpegNorm { float mean[3], stdv[3]; };

// one thread = one pixel: upsample chroma, YCbCr -> RGB (jdcolor.c, SCALEBITS 16); writes uint8 HWC,
// or (kNormalize) K1's output directly: (v/255 - mean)/std as bf16 NHWC with a zero 4th channel
template <bool kNormalize>
__global__ void __launch_bounds__(256)
jpeg_color_kernel(const uint8_t* __restrict__ planes, uint8_t* __restrict__ rgb, const JpegGeom g, int batch,
                  const JpegNorm nm) {
  const long long hw = (long long)g.width * g.height;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= hw * batch) return;
  const int img = (int)(idx / hw);
  const int rem = (int)(idx - (long long)img * hw);
  const int y = rem / g.width, x = rem - y * g.width;
  const uint8_t* base = planes + (long long)img * g.plane_bytes;
  const int yy = jpeg_sample(base + g.plane_off[0], g, 0, y, x);
  int r = yy, gg = yy, b = yy;
  if (g.comps == 3) {
    const int cb = jpeg_sample(base + g.plane_off[1], g, 1, y, x) - 128;
    const int cr = jpeg_sample(base + g.plane_off[2], g, 2, y, x) - 128;
    r = yy + ((91881 * cr + 32768) >> 16);                           // FIX(1.40200)
    b = yy + ((116130 * cb + 32768) >> 16);                          // FIX(1.77200)
    gg = yy + ((-22554 * cb + 32768 + -46802 * cr) >> 16);           // FIX(0.34414), FIX(0.71414)
    r = min(max(r, 0), 255); gg = min(max(gg, 0), 255); b = min(max(b, 0), 255);
  }
  if constexpr (kNormalize) {
    const uint32_t c0 = norm_to_bf16((float)r, nm.mean[0], nm.stdv[0]), c1 = norm_to_bf16((float)gg, nm.mean[1], nm.stdv[1]);
    const uint32_t c2 = norm_to_bf16((float)b, nm.mean[2], nm.stdv[2]);
    reinterpret_cast<uint2*>(rgb)[idx] = make_uint2(c0 | (c1 << 16), c2);
  } else {
    uint8_t* o = rgb + idx * 3;
    o[0] = (uint8_t)r; o[1] = (uint8_t)gg; o[2] = (uint8_t)b;
  }
}

static void make_geom(const ai4e_jpeg_info& in, JpegGeom* g) {
  memset(g, 0, sizeof(*g));
  g->comps = in.components; g->width = in.width; g->height = in.height;
  long long poff = 0;
  for (int c = 0; c < in.components; ++c) {
    g->blocks_w[c] = in.blocks_w[c]; g->blocks_h[c] = in.blocks_h[c];
    g->comp_w[c] = in.comp_w[c]; g->comp_h[c] = in.comp_h[c];
    g->hr[c] = in.hmax / in.h[c]; g->vr[c] = in.vmax / in.v[c];
    g->block_begin[c + 1] = g->block_begin[c] + in.blocks_w[c] * in.blocks_h[c];
    g->coef_off[c] = in.coef_off[c];
    g->plane_off[c] = poff;
    poff += (long long)in.blocks_w[c] * in.blocks_h[c] * 64;
  }
  g->record_int16 = in.record_int16;
  g->plane_bytes = in.plane_bytes;
}

// `norm_mean_std` == nullptr: d_out is uint8 [batch][H][W][3]; else 6 floats (mean[3], std[3]) and d_out
// is what K1 would have produced from those pixels: bf16 [batch][H][W][4], 4th channel zero
int jpeg_reconstruct(const int16_t* d_records, int batch, const ai4e_jpeg_info* info, uint8_t* d_scratch, uint8_t* d_rgb,
                     cudaStream_t stream, const float* norm_mean_std) {
  AI4E_REQUIRE(d_records && info && d_scratch && d_rgb && batch > 0, "jpeg_reconstruct: bad arguments");
  AI4E_REQUIRE((info->components == 1 || info->components == 3) && info->width > 0 && info->height > 0 &&
               info->record_int16 > 0 && info->plane_bytes > 0, "jpeg_reconstruct: geometry not filled in (call ai4e_jpeg_entropy_decode)");
  JpegGeom g;
  make_geom(*info, &g);
  const long long blocks = (long long)batch * g.block_begin[g.comps];
  jpeg_idct_kernel<<<(unsigned)((blocks + 127) / 128), 128, 0, stream>>>(d_records, d_scratch, g, batch);
  AI4E_CUDA_TRY(cudaGetLastError());
  const long long px = (long long)batch * info->width * info->height;
  JpegNorm nm{};
  if (norm_mean_std) {
    for (int i = 0; i < 3; ++i) { nm.mean[i] = norm_mean_std[i]; nm.stdv[i] = norm_mean_std[3 + i]; }
    jpeg_color_kernel<true><<<(unsigned)((px + 255) / 256), 256, 0, stream>>>(d_scratch, d_rgb, g, batch, nm);
  } else {
    jpeg_color_kernel<false><<<(unsigned)((px + 255) / 256), 256, 0, stream>>>(d_scratch, d_rgb, g, batch, nm);
  }
  AI4E_CUDA_TRY(cudaGetLastError());
  return AI4E_OK;
}

}  // namespace ai4e
