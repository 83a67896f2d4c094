// Baseline JPEG ENCODING arithmetic shared by the CUDA kernels (jpeg_enc.cu)
// and the host emulation the CPU tests run (tests/emu/jpeg_emu.cpp): the
// compressor of libjpeg-turbo, the encoder behind cv::imwrite
// (xpano/pipeline/stitcher_pipeline.cc:94-116), restated so that the STREAM is
// cv::imwrite's byte for byte -- jccolor.c (rgb_ycc_convert), jcsample.c
// (h2v1 / h2v2_downsample and the edge expansion), jccoefct.c (dummy blocks
// of a partial MCU), jfdctint.c (jpeg_fdct_islow), jcdctmgr.c (quantisation),
// jchuff.c (encode_one_block with the Annex K tables), jcmarker.c / jcparam.c
// (headers, quality scaling). libjpeg-turbo is a dependency of OpenCV, not
// vendored by the reference; the restatement is pinned against cv2.imencode
// in tests/test_jpeg_sm.py.
#pragma once
#include <stdint.h>

#include "hd.h"
#include "jpeg_sm.h"

namespace xp {
namespace jpg {

struct EncFrame {
  int width, height;
  int hs, vs;               // luma sampling factors (chroma is 1 x 1)
  int mcus_x, mcus_y, bpm;  // MCU grid, blocks per MCU (hs * vs + 2)
  int comp_w[3], comp_h[3];     // samples that exist
  int blocks_w[3], blocks_h[3];  // width_in_blocks / height_in_blocks
  int total_blocks;
};

struct EncTabs {
  uint16_t code[4][256];  // DC luma, AC luma, DC chroma, AC chroma
  uint8_t len[4][256];
  uint16_t q[2][64];      // luma, chroma; natural order
};

// Where a block of the scan lives: block index -> component, block column and
// row in the component, and whether it is one of jccoefct.c's dummy blocks
// (beyond the component's real blocks in a partial MCU).
struct BlockPos {
  int comp, bx, by, xi, yi;  // xi, yi: position inside the MCU
  bool dummy;
};
XP_HD BlockPos block_pos(const EncFrame& f, int blk) {
  BlockPos p;
  const int m = blk / f.bpm, b = blk - m * f.bpm;
  const int mx = m % f.mcus_x, my = m / f.mcus_x;
  const int ny = f.hs * f.vs;
  if (b < ny) {
    p.comp = 0;
    p.xi = b % f.hs;
    p.yi = b / f.hs;
    p.bx = mx * f.hs + p.xi;
    p.by = my * f.vs + p.yi;
  } else {
    p.comp = b - ny + 1;
    p.xi = p.yi = 0;
    p.bx = mx;
    p.by = my;
  }
  p.dummy = p.bx >= f.blocks_w[p.comp] || p.by >= f.blocks_h[p.comp];
  return p;
}
// the block whose DC a dummy block repeats (jccoefct.c compress_data)
XP_HD int dummy_source(const EncFrame& f, int blk) {
  for (int guard = 0; guard < 4; ++guard) {
    const BlockPos p = block_pos(f, blk);
    if (!p.dummy) return blk;
    if (p.by >= f.blocks_h[p.comp])
      blk -= p.xi + 1;  // a dummy row: the block before the row, in MCU order
    else
      blk -= 1;  // right edge: the block to the left
  }
  return blk;
}

// jccolor.c rgb_ycc_convert, SCALEBITS 16
XP_HD void bgr_to_ycc(int b, int g, int r, int* y, int* cb, int* cr) {
  *y = (19595 * r + 38470 * g + 7471 * b + 32768) >> 16;
  *cb = (-11059 * r - 21709 * g + 32768 * b + (128 << 16) + 32767) >> 16;
  *cr = (32768 * r - 27439 * g - 5329 * b + (128 << 16) + 32767) >> 16;
}

XP_HD int bit_length(int v) {  // of |v| < 2^15
  int n = 0;
  while (v) {
    ++n;
    v >>= 1;
  }
  return n;
}

// jfdctint.c jpeg_fdct_islow, one 8-point pass; `first`: the row pass
XP_HD void fdct_1d(const int32_t d[8], int32_t o[8], bool first) {
  const int32_t tmp0 = d[0] + d[7], tmp7 = d[0] - d[7];
  const int32_t tmp1 = d[1] + d[6], tmp6 = d[1] - d[6];
  const int32_t tmp2 = d[2] + d[5], tmp5 = d[2] - d[5];
  const int32_t tmp3 = d[3] + d[4], tmp4 = d[3] - d[4];
  const int32_t tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3;
  const int32_t tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  const int sh = first ? 11 : 15;  // CONST_BITS -/+ PASS1_BITS
  if (first) {
    o[0] = (int32_t)((uint32_t)(tmp10 + tmp11) << 2);
    o[4] = (int32_t)((uint32_t)(tmp10 - tmp11) << 2);
  } else {
    o[0] = descale(tmp10 + tmp11, 2);
    o[4] = descale(tmp10 - tmp11, 2);
  }
  int32_t z1 = mul32(tmp12 + tmp13, 4433);
  o[2] = descale(z1 + mul32(tmp13, 6270), sh);
  o[6] = descale(z1 + mul32(tmp12, -15137), sh);
  z1 = tmp4 + tmp7;
  int32_t z2 = tmp5 + tmp6, z3 = tmp4 + tmp6, z4 = tmp5 + tmp7;
  const int32_t z5 = mul32(z3 + z4, 9633);
  const int32_t t4 = mul32(tmp4, 2446), t5 = mul32(tmp5, 16819);
  const int32_t t6 = mul32(tmp6, 25172), t7 = mul32(tmp7, 12299);
  z1 = mul32(z1, -7373);
  z2 = mul32(z2, -20995);
  z3 = mul32(z3, -16069) + z5;
  z4 = mul32(z4, -3196) + z5;
  o[7] = descale(t4 + z1 + z3, sh);
  o[5] = descale(t5 + z2 + z4, sh);
  o[3] = descale(t6 + z2 + z3, sh);
  o[1] = descale(t7 + z1 + z4, sh);
}

// samples (0..255, row-major) -> quantised coefficients in ZIGZAG order
XP_HD void fdct_quantise(const uint8_t px[64], const uint16_t q[64],
                         int16_t out[64]) {
  constexpr uint8_t zz[64] = {
      0,  1,  5,  6,  14, 15, 27, 28, 2,  4,  7,  13, 16, 26, 29, 42,
      3,  8,  12, 17, 25, 30, 41, 43, 9,  11, 18, 24, 31, 40, 44, 53,
      10, 19, 23, 32, 39, 45, 52, 54, 20, 22, 33, 38, 46, 51, 55, 60,
      21, 34, 37, 47, 50, 56, 59, 61, 35, 36, 48, 49, 57, 58, 62, 63};
  int32_t ws[64];
  XP_UNROLL
  for (int r = 0; r < 8; ++r) {
    int32_t d[8], o[8];
    XP_UNROLL
    for (int c = 0; c < 8; ++c) d[c] = (int32_t)px[8 * r + c] - 128;
    fdct_1d(d, o, true);
    XP_UNROLL
    for (int c = 0; c < 8; ++c) ws[8 * r + c] = o[c];
  }
  XP_UNROLL
  for (int c = 0; c < 8; ++c) {
    int32_t d[8], o[8];
    XP_UNROLL
    for (int r = 0; r < 8; ++r) d[r] = ws[8 * r + c];
    fdct_1d(d, o, false);
    XP_UNROLL
    for (int r = 0; r < 8; ++r) {
      // jcdctmgr.c: the DCT is scaled by 8; round half away from zero
      const int32_t div = (int32_t)q[8 * r + c] << 3;
      int32_t t = o[r];
      const bool neg = t < 0;
      if (neg) t = -t;
      t += div >> 1;
      t = t >= div ? t / div : 0;
      out[zz[8 * r + c]] = (int16_t)(neg ? -t : t);
    }
  }
}

// jchuff.c encode_one_block: coef in zigzag order, pred = the component's
// previous DC. sink(table, symbol, extra, nbits) receives every Huffman
// symbol (table: dc_tab or ac_tab) and the nbits value bits that follow it.
template <class Sink>
XP_HD void block_symbols(const int16_t coef[64], int pred, int dc_tab,
                         int ac_tab, Sink&& sink) {
  int diff = coef[0] - pred;
  int v = diff < 0 ? -diff : diff;
  int nbits = bit_length(v);
  sink(dc_tab, nbits,
       (uint32_t)(diff < 0 ? diff - 1 : diff) & ((1u << nbits) - 1), nbits);
  int run = 0;
  for (int k = 1; k < 64; ++k) {
    const int c = coef[k];
    if (c == 0) {
      ++run;
      continue;
    }
    while (run > 15) {
      sink(ac_tab, 0xF0, 0u, 0);
      run -= 16;
    }
    v = c < 0 ? -c : c;
    nbits = bit_length(v);
    sink(ac_tab, (run << 4) + nbits,
         (uint32_t)(c < 0 ? c - 1 : c) & ((1u << nbits) - 1), nbits);
    run = 0;
  }
  if (run > 0) sink(ac_tab, 0, 0u, 0);
}

// the same as bits: emit(bits, n) receives every code and value, n <= 16
template <class Emit>
XP_HD void encode_block(const int16_t coef[64], int pred, const EncTabs& t,
                        int dc_tab, int ac_tab, Emit&& emit) {
  block_symbols(coef, pred, dc_tab, ac_tab,
                [&](int tab, int sym, uint32_t extra, int nbits) {
                  emit((uint32_t)t.code[tab][sym], (int)t.len[tab][sym]);
                  if (nbits) emit(extra, nbits);
                });
}

}  // namespace jpg
}  // namespace xp
