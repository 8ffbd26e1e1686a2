// b200ps — baseline JPEG front end for the image input pipeline (input_pipeline.py:68-78: read_file +
// decode_jpeg(channels=3)).  The entropy-coded segment is inherently serial, so it is decoded on the host:
// markers, quantisation and Huffman tables, then every 8x8 block's quantised coefficients in natural (row-major)
// order.  Everything after that — dequantisation, the inverse DCT, chroma upsampling, YCbCr -> RGB, the optional
// resize — is data parallel and runs on the GPU (csrc/kernels/jpeg.cu).
//
// Supported: baseline sequential DCT (SOF0), 8-bit, 1 or 3 components, any sampling factors up to 2x2 per
// component, one interleaved scan, restart intervals.  Progressive / arithmetic / 12-bit files are refused.
#include <cstdint>
#include <cstring>

#include "../common.h"

namespace {

const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                             41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                             30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct Huff {
  bool present = false;
  uint8_t symbols[256];
  int mincode[17], maxcode[18], valptr[17];   // by code length 1..16
  uint8_t look_len[512];                      // 9-bit fast path: code length (0 = longer than 9 bits)
  uint8_t look_sym[512];
  // AC tables: when a code AND its value bits fit the 9-bit window, one lookup yields everything —
  // (value << 8) | (run << 4) | (code length + value bits); 0 = take the general path
  int16_t fast_ac[512];

  void build(const uint8_t counts[16], const uint8_t* syms, int nsyms, bool is_ac) {
    present = true;
    memcpy(symbols, syms, nsyms);
    memset(look_len, 0, sizeof(look_len));
    int code = 0, k = 0;
    for (int len = 1; len <= 16; ++len) {
      valptr[len] = k;
      mincode[len] = code;
      for (int i = 0; i < counts[len - 1]; ++i, ++k, ++code) {
        if (len <= 9) {
          const int first = code << (9 - len), n = 1 << (9 - len);
          for (int j = 0; j < n; ++j) { look_len[first + j] = (uint8_t)len; look_sym[first + j] = syms[k]; }
        }
      }
      maxcode[len] = counts[len - 1] ? code - 1 : -1;
      code <<= 1;
    }
    maxcode[17] = 0x7fffffff;
    for (int i = 0; is_ac && i < 512; ++i) {
      fast_ac[i] = 0;
      const int len = look_len[i];
      if (!len) continue;
      const int rs = look_sym[i], run = rs >> 4, sz = rs & 15;
      if (sz == 0 || len + sz > 9) continue;
      int v = ((i << len) & 511) >> (9 - sz);                 // the sz bits behind the code
      if (v < (1 << (sz - 1))) v += (int)((~0u) << sz) + 1;   // sign extension of the JPEG magnitude code
      if (v >= -128 && v <= 127) fast_ac[i] = (int16_t)(v * 256 + run * 16 + (len + sz));
    }
  }
};

struct BitReader {
  const uint8_t* p;
  const uint8_t* end;
  uint64_t acc = 0;                            // the next bits of the stream, left-aligned
  int nbits = 0;
  bool hit_marker = false;

  void fill() {                                // up to 64 bits ahead; zeros past the data or behind a marker
    while (nbits <= 56) {
      uint64_t byte = 0;
      if (!hit_marker && p < end) {
        byte = *p;
        if (byte == 0xFF) {
          if (p + 1 < end && p[1] == 0x00) p += 2;           // stuffed zero
          else { hit_marker = true; byte = 0; }              // a marker: feed zeros until the caller handles it
        } else {
          ++p;
        }
      }
      acc |= byte << (56 - nbits);
      nbits += 8;
    }
  }
  inline uint32_t peek(int n) { if (nbits < n) fill(); return (uint32_t)(acc >> (64 - n)); }
  inline void skip(int n) { acc <<= n; nbits -= n; }
  inline int get(int n) { if (n == 0) return 0; const uint32_t v = peek(n); skip(n); return (int)v; }
  void reset() { acc = 0; nbits = 0; hit_marker = false; }
};

inline int decode_symbol(BitReader& br, const Huff& h) {
  const uint32_t look = br.peek(9);
  const int l = h.look_len[look];
  if (l) { br.skip(l); return h.look_sym[look]; }
  int code = (int)br.peek(16);
  for (int len = 10; len <= 16; ++len) {
    const int c = code >> (16 - len);
    if (h.maxcode[len] >= 0 && c <= h.maxcode[len] && c >= h.mincode[len]) {
      br.skip(len);
      return h.symbols[h.valptr[len] + c - h.mincode[len]];
    }
  }
  return -1;
}

inline int extend(int v, int s) { return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v; }

inline int rd16(const uint8_t* p) { return (p[0] << 8) | p[1]; }

}  // namespace

using namespace b200ps;

// Header only: image size and component count.
extern "C" int b200ps_jpeg_info(const uint8_t* data, uint64_t len, int32_t* width, int32_t* height, int32_t* ncomp) {
  B200PS_REQUIRE(data && len > 4 && data[0] == 0xFF && data[1] == 0xD8, "jpeg: not a JPEG stream (no SOI)");
  uint64_t pos = 2;
  while (pos + 4 <= len) {
    if (data[pos] != 0xFF) { ++pos; continue; }
    const int m = data[pos + 1];
    if (m == 0xFF) { ++pos; continue; }
    if (m == 0xD8 || (m >= 0xD0 && m <= 0xD7) || m == 0x01) { pos += 2; continue; }
    const int seg = rd16(data + pos + 2);
    if (m == 0xC0 || m == 0xC1 || m == 0xC2) {
      B200PS_REQUIRE(pos + 2 + seg <= len && seg >= 8, "jpeg: truncated frame header");
      *height = rd16(data + pos + 5);
      *width = rd16(data + pos + 7);
      *ncomp = data[pos + 9];
      return 0;
    }
    pos += 2 + seg;
  }
  return set_error(B200PS_ERR_INVALID, "jpeg: no frame header found");
}

// Entropy-decode a whole baseline JPEG.  coef[coef_cap] receives the quantised coefficients, 64 per block in
// natural order, component after component (desc->comp[c].coef_offset, blocks_w x blocks_h blocks each).
extern "C" int b200ps_jpeg_parse(const uint8_t* data, uint64_t len, int16_t* coef, uint64_t coef_cap,
                                 b200ps_jpeg_desc* desc) {
  B200PS_REQUIRE(data && desc && len > 4 && data[0] == 0xFF && data[1] == 0xD8, "jpeg: not a JPEG stream (no SOI)");
  memset(desc, 0, sizeof(*desc));
  uint16_t qt[4][64];
  bool qt_present[4] = {false, false, false, false};
  Huff dc[4], ac[4];
  int comp_id[3] = {0, 0, 0}, comp_tq[3] = {0, 0, 0};
  int restart_interval = 0;
  bool have_frame = false;
  uint64_t pos = 2;
  while (pos + 4 <= len) {
    if (data[pos] != 0xFF) { ++pos; continue; }
    const int m = data[pos + 1];
    if (m == 0xFF) { ++pos; continue; }
    if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) { pos += 2; continue; }
    if (m == 0xD9) break;
    const int seg = rd16(data + pos + 2);
    B200PS_REQUIRE(seg >= 2 && pos + 2 + seg <= len, "jpeg: truncated segment (marker %02x)", m);
    const uint8_t* s = data + pos + 4;
    const int n = seg - 2;
    if (m == 0xDB) {                                            // DQT
      int o = 0;
      while (o < n) {
        const int pq = s[o] >> 4, tq = s[o] & 15;
        B200PS_REQUIRE(tq < 4 && o + 1 + (pq ? 128 : 64) <= n, "jpeg: bad quantisation table");
        for (int i = 0; i < 64; ++i) qt[tq][kZigzag[i]] = pq ? (uint16_t)rd16(s + o + 1 + 2 * i) : s[o + 1 + i];
        qt_present[tq] = true;
        o += 1 + (pq ? 128 : 64);
      }
    } else if (m == 0xC4) {                                     // DHT
      int o = 0;
      while (o + 17 <= n) {
        const int tc = s[o] >> 4, th = s[o] & 15;
        int total = 0;
        for (int i = 0; i < 16; ++i) total += s[o + 1 + i];
        B200PS_REQUIRE(tc < 2 && th < 4 && total <= 256 && o + 17 + total <= n, "jpeg: bad Huffman table");
        (tc ? ac[th] : dc[th]).build(s + o + 1, s + o + 17, total, tc != 0);
        o += 17 + total;
      }
    } else if (m == 0xC0 || m == 0xC1) {                        // SOF0 / SOF1 (Huffman, sequential)
      B200PS_REQUIRE(n >= 6 && s[0] == 8, "jpeg: only 8-bit samples are supported");
      desc->height = rd16(s + 1);
      desc->width = rd16(s + 3);
      desc->ncomp = s[5];
      B200PS_REQUIRE((desc->ncomp == 1 || desc->ncomp == 3) && n >= 6 + 3 * desc->ncomp && desc->width > 0 && desc->height > 0,
                     "jpeg: %d components / empty image not supported", desc->ncomp);
      int hmax = 1, vmax = 1;
      for (int c = 0; c < desc->ncomp; ++c) {
        comp_id[c] = s[6 + 3 * c];
        desc->comp[c].h = s[7 + 3 * c] >> 4;
        desc->comp[c].v = s[7 + 3 * c] & 15;
        comp_tq[c] = s[8 + 3 * c];
        B200PS_REQUIRE(desc->comp[c].h >= 1 && desc->comp[c].h <= 2 && desc->comp[c].v >= 1 && desc->comp[c].v <= 2 && comp_tq[c] < 4,
                       "jpeg: sampling factors beyond 2x2 are not supported");
        hmax = desc->comp[c].h > hmax ? desc->comp[c].h : hmax;
        vmax = desc->comp[c].v > vmax ? desc->comp[c].v : vmax;
      }
      if (desc->ncomp == 1) { desc->comp[0].h = desc->comp[0].v = 1; hmax = vmax = 1; }
      desc->hmax = hmax; desc->vmax = vmax;
      desc->mcus_x = (desc->width + 8 * hmax - 1) / (8 * hmax);
      desc->mcus_y = (desc->height + 8 * vmax - 1) / (8 * vmax);
      int64_t off = 0, poff = 0;
      for (int c = 0; c < desc->ncomp; ++c) {
        auto& cc = desc->comp[c];
        cc.blocks_w = desc->mcus_x * cc.h;
        cc.blocks_h = desc->mcus_y * cc.v;
        cc.coef_offset = off;
        cc.plane_offset = poff;
        off += (int64_t)cc.blocks_w * cc.blocks_h * 64;
        poff += (int64_t)cc.blocks_w * cc.blocks_h * 64;
      }
      desc->total_coefs = off;
      desc->total_blocks = (int32_t)(off / 64);
      have_frame = true;
    } else if (m == 0xC2 || (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC)) {
      return set_error(B200PS_ERR_INVALID, "jpeg: only baseline sequential Huffman files are supported (SOF marker %02x)", m);
    } else if (m == 0xDD) {                                     // DRI
      B200PS_REQUIRE(n >= 2, "jpeg: bad DRI");
      restart_interval = rd16(s);
    } else if (m == 0xDA) {                                     // SOS: the one interleaved scan
      B200PS_REQUIRE(have_frame, "jpeg: scan before frame header");
      const int ns = s[0];
      B200PS_REQUIRE(ns == desc->ncomp && n >= 1 + 2 * ns + 3, "jpeg: only one interleaved scan of all components is supported");
      int td[3], ta[3];
      for (int k = 0; k < ns; ++k) {
        int c = -1;
        for (int q = 0; q < desc->ncomp; ++q) if (comp_id[q] == s[1 + 2 * k]) c = q;
        B200PS_REQUIRE(c == k, "jpeg: scan components out of frame order");
        td[k] = s[2 + 2 * k] >> 4; ta[k] = s[2 + 2 * k] & 15;
        B200PS_REQUIRE(td[k] < 4 && ta[k] < 4 && dc[td[k]].present && ac[ta[k]].present, "jpeg: scan names a missing Huffman table");
      }
      B200PS_REQUIRE(coef && coef_cap >= (uint64_t)desc->total_coefs, "jpeg: coefficient buffer too small (%lld needed)",
                     (long long)desc->total_coefs);
      for (int c = 0; c < desc->ncomp; ++c) {
        B200PS_REQUIRE(qt_present[comp_tq[c]], "jpeg: missing quantisation table");
        for (int i = 0; i < 64; ++i) desc->comp[c].quant[i] = qt[comp_tq[c]][i];
      }
      memset(coef, 0, sizeof(int16_t) * desc->total_coefs);
      BitReader br;
      br.p = data + pos + 2 + seg;
      br.end = data + len;
      int pred[3] = {0, 0, 0};
      int until_restart = restart_interval;
      const int nmcu = desc->mcus_x * desc->mcus_y;
      for (int mcu = 0; mcu < nmcu; ++mcu) {
        if (restart_interval && until_restart == 0) {           // byte-align, consume RSTn, reset predictors
          br.reset();
          while (br.p + 1 < br.end && !(br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7)) ++br.p;
          if (br.p + 1 < br.end) br.p += 2;
          pred[0] = pred[1] = pred[2] = 0;
          until_restart = restart_interval;
        }
        const int my = mcu / desc->mcus_x, mx = mcu % desc->mcus_x;
        for (int c = 0; c < desc->ncomp; ++c) {
          const auto& cc = desc->comp[c];
          for (int v = 0; v < cc.v; ++v)
            for (int h = 0; h < cc.h; ++h) {
              int16_t* blk = coef + cc.coef_offset + ((int64_t)(my * cc.v + v) * cc.blocks_w + mx * cc.h + h) * 64;
              int sbits = decode_symbol(br, dc[td[c]]);
              B200PS_REQUIRE(sbits >= 0 && sbits <= 11, "jpeg: corrupt DC code (MCU %d)", mcu);
              if (sbits) pred[c] += extend(br.get(sbits), sbits);
              blk[0] = (int16_t)pred[c];
              const Huff& hac = ac[ta[c]];
              for (int k = 1; k < 64;) {
                const int f = hac.fast_ac[br.peek(9)];
                if (f) {                                        // run, value and length from one lookup
                  k += (f >> 4) & 15;
                  B200PS_REQUIRE(k < 64, "jpeg: AC run past the block (MCU %d)", mcu);
                  blk[kZigzag[k]] = (int16_t)(f >> 8);
                  br.skip(f & 15);
                  ++k;
                  continue;
                }
                const int rs = decode_symbol(br, hac);
                B200PS_REQUIRE(rs >= 0, "jpeg: corrupt AC code (MCU %d)", mcu);
                const int r = rs >> 4, sz = rs & 15;
                if (sz == 0) {
                  if (r == 15) { k += 16; continue; }
                  break;                                        // end of block
                }
                k += r;
                B200PS_REQUIRE(k < 64, "jpeg: AC run past the block (MCU %d)", mcu);
                blk[kZigzag[k]] = (int16_t)extend(br.get(sz), sz);
                ++k;
              }
            }
        }
        if (restart_interval) --until_restart;
      }
      return 0;                                                 // one scan holds the whole image
    }
    pos += 2 + seg;
  }
  return set_error(B200PS_ERR_INVALID, "jpeg: no scan found");
}
