// Baseline JPEG decoding arithmetic shared by the CUDA kernels (jpeg_sm.cu)
// and the host emulation the CPU tests run (tests/emu/jpeg_emu.cpp): the
// Huffman-coded scan, destuffed, is cut into fixed-size subsequences that are
// decoded independently and synchronise on their neighbours' exit states; then the
// inverse DCT, chroma upsampling and colour conversion of libjpeg-turbo, the
// decoder behind cv::imread (xpano/algorithm/image.cc:83), restated so that
// the pixels are cv::imread's bit for bit: jidctint.c (jpeg_idct_islow),
// jdsample.c (h2v1 / h2v2 fancy upsampling), jdcolor.c (ycc_rgb_convert).
// libjpeg-turbo is a dependency of OpenCV, not vendored by the reference; the
// restatement is pinned against cv2.imdecode in tests/test_jpeg_sm.py.
#pragma once
#include <stdint.h>

#include "hd.h"

#if defined(__CUDACC__)
#define XP_UNROLL _Pragma("unroll")
#else
#define XP_UNROLL
#endif

namespace xp {
namespace jpg {

constexpr int kLook = 10;  // bits resolved by the first-level Huffman table

struct HuffTab {
  uint16_t look[1 << kLook];  // (length << 8) | symbol, 0 = longer code
  int32_t maxcode[18];        // largest code of each length, -1 = none
  int32_t valoff[18];         // index of a length's first symbol - its code
  uint8_t val[256];
};

struct Frame {
  int width, height, ncomp;
  int hs[3], vs[3];          // sampling factors
  int mcus_x, mcus_y, bpm;   // MCU grid, blocks per MCU
  int blk_comp[10], blk_x[10], blk_y[10];  // blocks of an MCU in scan order
  int blk_first[3];          // first block of a component within the MCU
  int dc_tab[3], ac_tab[3];  // Huffman tables (0..1 DC, 2..3 AC)
  int blk_dc[10], blk_ac[10];  // the same per block of the MCU
  int qt[3];
  int restart;               // MCUs per restart interval, 0 = none
  int comp_w[3], comp_h[3];  // samples that exist (downsampled_width/height)
  int plane_w[3], plane_h[3];  // padded to whole MCUs
  int total_blocks;
  int chroma;                // 0: 4:4:4 or grey, 1: h2v1, 2: h2v2
  int fancy;                 // triangle filter (chroma wider than 2 samples)
};

struct SubRec {  // one subsequence of the scan
  uint32_t start, end;  // bit range in the destuffed scan
  uint32_t first_mcu;   // of its restart interval
  uint32_t is_first;    // starts the interval (state known)
};

// zigzag position -> natural (row-major) index
#define XP_JPG_NATURAL                                                   \
  {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,       \
   12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28,      \
   35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,      \
   58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63}

// The scan is decoded from a DESTUFFED copy: the zero after every FF and the
// restart markers are taken out first (what remains of a restart interval
// still starts on a byte), and the bytes are packed big-endian into 32-bit
// words, so that any 32 bits of the stream are two loads and a funnel shift.
// A byte of [begin, end) is dropped when it is the 00 or Dn after an FF, or
// the FF before a Dn.
XP_HD bool destuff_keeps(const uint8_t* d, size_t i, size_t begin,
                         size_t end) {
  const uint8_t v = d[i];
  if (i > begin && d[i - 1] == 0xFF && (v == 0x00 || (v >= 0xD0 && v <= 0xD7)))
    return false;
  if (v == 0xFF && i + 1 < end && d[i + 1] >= 0xD0 && d[i + 1] <= 0xD7)
    return false;
  return true;
}

// Decoder state between two symbols, packed so that one 64-bit load or store
// moves it whole: bit position in the destuffed scan, block of the MCU,
// zigzag index.
struct State {
  uint32_t p;
  int b, z;
};
XP_HD uint64_t pack(const State& s) {
  return ((uint64_t)s.p << 16) | ((uint64_t)s.b << 8) | (uint64_t)s.z;
}
XP_HD State unpack(uint64_t v) {
  State s;
  s.p = (uint32_t)(v >> 16);
  s.b = (int)(v >> 8) & 15;
  s.z = (int)v & 63;
  return s;
}

// the 32 bits of the destuffed scan from bit p on (words are zero past its end)
XP_HD uint32_t peek32(const uint32_t* w, uint32_t p) {
  const uint32_t i = p >> 5, s = p & 31;
  const uint32_t a = w[i], b = w[i + 1];
#if defined(__CUDA_ARCH__)
  return __funnelshift_l(b, a, s);
#else
  return s ? (a << s) | (b >> (32 - s)) : a;
#endif
}

XP_HD void huff_symbol(const HuffTab& t, uint32_t v16, int* len, int* sym) {
  const uint32_t e = t.look[v16 >> (16 - kLook)];
  if (e) {
    *len = (int)(e >> 8);
    *sym = (int)(e & 255);
    return;
  }
  int l = kLook + 1;
  int32_t code = (int32_t)(v16 >> (16 - l));
  while (l <= 16 && code > t.maxcode[l]) {
    ++l;
    code = (int32_t)(v16 >> (l <= 16 ? 16 - l : 0));
  }
  if (l > 16) {  // not a code (only reached while still out of step)
    *len = 1;
    *sym = 0;
    return;
  }
  *len = l;
  *sym = t.val[(code + t.valoff[l]) & 255];
}

// Decode the symbols that START in [s.p, sub_end) and return how many blocks
// they completed. With kWrite (the write pass) the AC coefficients go to
// coef[64 * block + zigzag index] and the DC difference to the previous block
// of the component to dc[block]; `block` is the index of the block in
// progress and blocks from block_limit on are not stored.
template <bool kWrite>
XP_HD int decode_subsequence(const uint32_t* words, uint32_t sub_end,
                             const Frame& f, const HuffTab* tabs, State* s,
                             int16_t* coef, int16_t* dc, int block,
                             int block_limit) {
  uint32_t p = s->p;
  int b = s->b, z = s->z, done = 0;
  while (p < sub_end) {
    const HuffTab& t = tabs[z == 0 ? f.blk_dc[b] : f.blk_ac[b]];
    const uint32_t v = peek32(words, p);
    int len, sym;
    huff_symbol(t, v >> 16, &len, &sym);
    int run = 0, size = sym & 15;
    bool eob = false;
    if (z != 0) {
      run = sym >> 4;
      if (size == 0) {
        if (run == 15)
          run = 16;  // ZRL: sixteen zeros, no value
        else
          eob = true;
      }
    }
    int value = 0;
    if (size) {
      value = (int)((v << len) >> (32 - size));
      if (value < (1 << (size - 1))) value -= (1 << size) - 1;
    }
    p += (uint32_t)(len + size);
    if (eob) {
      z = 64;
    } else {
      z += run;
      if (size || z == 0) {
        if (kWrite && z < 64 && block < block_limit) {
          if (z == 0)
            dc[block] = (int16_t)value;
          else
            coef[(size_t)block * 64 + z] = (int16_t)value;
        }
        ++z;
      }
    }
    if (z >= 64) {
      z = 0;
      b = b + 1 == f.bpm ? 0 : b + 1;
      ++block;
      ++done;
    }
  }
  s->p = p;
  s->b = b;
  s->z = z;
  return done;
}

// ---------------------------------------------------------------------------
// jpeg_idct_islow (jidctint.c), 8x8, 8-bit samples: `in` are the quantised
// coefficients in ZIGZAG order (as the scan stores them; in[0] is ignored, the
// DC value is passed separately), q the quantisation table in natural order,
// out[8][8] samples.
// ---------------------------------------------------------------------------
XP_HD int32_t mul32(int32_t a, int32_t b) {
  return (int32_t)((uint32_t)a * (uint32_t)b);
}
XP_HD int32_t descale(int32_t x, int n) {
  return (int32_t)((uint32_t)x + (1u << (n - 1))) >> n;
}
XP_HD uint8_t idct_limit(int32_t x) {  // range_limit[(x) & RANGE_MASK]
  int32_t v = ((x & 1023) ^ 512) - 512 + 128;
  return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
}

// (unsigned arithmetic: a damaged stream may overflow 32 bits, which wraps
// here as it does on the device instead of being undefined on the host)
XP_HD void idct_1d(const int32_t in[8], int32_t even_shift, int32_t o[8],
                   int shift) {
  typedef uint32_t u32;
  u32 i[8];
  XP_UNROLL
  for (int k = 0; k < 8; ++k) i[k] = (u32)in[k];
  // even part
  u32 z2 = i[2], z3 = i[6];
  u32 z1 = (z2 + z3) * 4433u;
  u32 tmp2 = z1 + z3 * (u32)-15137;
  u32 tmp3 = z1 + z2 * 6270u;
  u32 tmp0 = (i[0] + i[4]) << even_shift;
  u32 tmp1 = (i[0] - i[4]) << even_shift;
  const u32 tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3;
  const u32 tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  // odd part
  tmp0 = i[7], tmp1 = i[5], tmp2 = i[3], tmp3 = i[1];
  z1 = tmp0 + tmp3;
  z2 = tmp1 + tmp2;
  z3 = tmp0 + tmp2;
  u32 z4 = tmp1 + tmp3;
  const u32 z5 = (z3 + z4) * 9633u;
  tmp0 *= 2446u;
  tmp1 *= 16819u;
  tmp2 *= 25172u;
  tmp3 *= 12299u;
  z1 *= (u32)-7373;
  z2 *= (u32)-20995;
  z3 *= (u32)-16069;
  z4 *= (u32)-3196;
  z3 += z5;
  z4 += z5;
  tmp0 += z1 + z3;
  tmp1 += z2 + z4;
  tmp2 += z2 + z3;
  tmp3 += z1 + z4;
  o[0] = descale((int32_t)(tmp10 + tmp3), shift);
  o[7] = descale((int32_t)(tmp10 - tmp3), shift);
  o[1] = descale((int32_t)(tmp11 + tmp2), shift);
  o[6] = descale((int32_t)(tmp11 - tmp2), shift);
  o[2] = descale((int32_t)(tmp12 + tmp1), shift);
  o[5] = descale((int32_t)(tmp12 - tmp1), shift);
  o[3] = descale((int32_t)(tmp13 + tmp0), shift);
  o[4] = descale((int32_t)(tmp13 - tmp0), shift);
}

XP_HD void idct_islow(const int16_t in[64], int dc, const uint16_t q[64],
                      uint8_t out[64]) {
  // natural (row-major) index -> zigzag position
  constexpr uint8_t zz[64] = {
      0,  1,  5,  6,  14, 15, 27, 28, 2,  4,  7,  13, 16, 26, 29, 42,
      3,  8,  12, 17, 25, 30, 41, 43, 9,  11, 18, 24, 31, 40, 44, 53,
      10, 19, 23, 32, 39, 45, 52, 54, 20, 22, 33, 38, 46, 51, 55, 60,
      21, 34, 37, 47, 50, 56, 59, 61, 35, 36, 48, 49, 57, 58, 62, 63};
  int32_t ws[64];
  XP_UNROLL
  for (int c = 0; c < 8; ++c) {  // columns, CONST_BITS - PASS1_BITS
    int32_t i[8], o[8];
    XP_UNROLL
    for (int r = 0; r < 8; ++r)
      i[r] = (int32_t)(8 * r + c == 0 ? dc : in[zz[8 * r + c]]) *
             (int32_t)q[8 * r + c];
    idct_1d(i, 13, o, 11);
    XP_UNROLL
    for (int r = 0; r < 8; ++r) ws[8 * r + c] = o[r];
  }
  XP_UNROLL
  for (int r = 0; r < 8; ++r) {  // rows, CONST_BITS + PASS1_BITS + 3
    int32_t o[8];
    idct_1d(ws + 8 * r, 13, o, 18);
    XP_UNROLL
    for (int c = 0; c < 8; ++c) out[8 * r + c] = idct_limit(o[c]);
  }
}

// ---------------------------------------------------------------------------
// One chroma sample at full resolution (jdsample.c): p = the component plane,
// cw x ch its existing samples, (X, Y) the output pixel.
// ---------------------------------------------------------------------------
XP_HD int upsample_h2v1(const uint8_t* p, int pitch, int cw, int X, int Y) {
  const uint8_t* r = p + (size_t)Y * pitch;
  const int i = X >> 1;
  if (X & 1)
    return i == cw - 1 ? r[i] : (3 * r[i] + r[i + 1] + 2) >> 2;
  return i == 0 ? r[i] : (3 * r[i] + r[i - 1] + 1) >> 2;
}
XP_HD int upsample_h2v2(const uint8_t* p, int pitch, int cw, int ch, int X,
                        int Y) {
  const int i = X >> 1, r = Y >> 1;
  int rn = (Y & 1) ? r + 1 : r - 1;  // the nearer neighbouring row
  rn = rn < 0 ? 0 : rn > ch - 1 ? ch - 1 : rn;
  const uint8_t* a = p + (size_t)r * pitch;
  const uint8_t* b = p + (size_t)rn * pitch;
  const int cur = 3 * a[i] + b[i];
  if (X & 1) {
    if (i == cw - 1) return (cur * 4 + 7) >> 4;
    return (cur * 3 + 3 * a[i + 1] + b[i + 1] + 7) >> 4;
  }
  if (i == 0) return (cur * 4 + 8) >> 4;
  return (cur * 3 + 3 * a[i - 1] + b[i - 1] + 8) >> 4;
}

XP_HD uint8_t clamp8(int v) { return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }
// ycc_rgb_convert (jdcolor.c), SCALEBITS 16
XP_HD void ycc_to_bgr(int y, int cb, int cr, uint8_t bgr[3]) {
  cb -= 128;
  cr -= 128;
  const int r = (91881 * cr + 32768) >> 16;
  const int b = (116130 * cb + 32768) >> 16;
  const int g = (-22554 * cb + 32768 - 46802 * cr) >> 16;
  bgr[0] = clamp8(y + b);
  bgr[1] = clamp8(y + g);
  bgr[2] = clamp8(y + r);
}

XP_HD void pixel_bgr(const Frame& f, const uint8_t* const plane[3],
                     const int pitch[3], int X, int Y, uint8_t bgr[3]) {
  const int y = plane[0][(size_t)Y * pitch[0] + X];
  if (f.ncomp == 1) {
    bgr[0] = bgr[1] = bgr[2] = (uint8_t)y;
    return;
  }
  int cb, cr;
  if (f.chroma && !f.fancy) {  // jdsample.c: replication when width <= 2
    const int r = f.chroma == 2 ? Y >> 1 : Y;
    cb = plane[1][(size_t)r * pitch[1] + (X >> 1)];
    cr = plane[2][(size_t)r * pitch[2] + (X >> 1)];
  } else if (f.chroma == 2) {
    cb = upsample_h2v2(plane[1], pitch[1], f.comp_w[1], f.comp_h[1], X, Y);
    cr = upsample_h2v2(plane[2], pitch[2], f.comp_w[2], f.comp_h[2], X, Y);
  } else if (f.chroma == 1) {
    cb = upsample_h2v1(plane[1], pitch[1], f.comp_w[1], X, Y);
    cr = upsample_h2v1(plane[2], pitch[2], f.comp_w[2], X, Y);
  } else {
    cb = plane[1][(size_t)Y * pitch[1] + X];
    cr = plane[2][(size_t)Y * pitch[2] + X];
  }
  ycc_to_bgr(y, cb, cr, bgr);
}

}  // namespace jpg
}  // namespace xp
