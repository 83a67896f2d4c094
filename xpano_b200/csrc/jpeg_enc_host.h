// Host side of the SM JPEG encoder: quality scaling of the Annex K tables
// (jcparam.c jpeg_set_quality / jpeg_add_quant_table, force_baseline),
// derived Huffman codes (jchuff.c jpeg_make_c_derived_tbl), frame geometry
// (jcmaster.c initial_setup / per_scan_setup) and the header bytes as
// jcmarker.c writes them for the settings cv::imwrite uses. Plain C++; also
// compiled into tests/emu/jpeg_emu.cpp.
#pragma once
#include <cstring>
#include <vector>

#include "jpeg_enc.h"
#include "jpeg_std_tables.h"

namespace xp {
namespace jpg {

static const uint8_t kNaturalEnc[64] = XP_JPG_NATURAL;

struct EncHost {
  EncFrame f;
  EncTabs t;
  uint8_t qzz[2][64];  // the tables as DQT carries them (zigzag order)
};

inline void derive_codes(const uint8_t bits[16], const uint8_t* vals, int n,
                         uint16_t code[256], uint8_t len[256]) {
  memset(code, 0, 256 * sizeof(uint16_t));
  memset(len, 0, 256);
  int c = 0, p = 0;
  for (int l = 1; l <= 16; ++l) {
    for (int i = 0; i < bits[l - 1] && p < n; ++i, ++p, ++c) {
      code[vals[p]] = (uint16_t)c;
      len[vals[p]] = (uint8_t)l;
    }
    c <<= 1;
  }
}

// subsampling: 444, 422 or 420
inline bool enc_setup(int width, int height, int quality, int subsampling,
                      EncHost* e) {
  if (width <= 0 || height <= 0 || width > 65500 || height > 65500)
    return false;
  EncFrame& f = e->f;
  memset(&f, 0, sizeof(f));
  f.width = width;
  f.height = height;
  f.hs = subsampling == 444 ? 1 : 2;
  f.vs = subsampling == 420 ? 2 : 1;
  if (subsampling != 444 && subsampling != 422 && subsampling != 420)
    return false;
  f.mcus_x = (width + 8 * f.hs - 1) / (8 * f.hs);
  f.mcus_y = (height + 8 * f.vs - 1) / (8 * f.vs);
  f.bpm = f.hs * f.vs + 2;
  for (int c = 0; c < 3; ++c) {
    const int h = c ? 1 : f.hs, v = c ? 1 : f.vs;
    f.comp_w[c] = (width * h + f.hs - 1) / f.hs;
    f.comp_h[c] = (height * v + f.vs - 1) / f.vs;
    f.blocks_w[c] = (f.comp_w[c] + 7) / 8;
    f.blocks_h[c] = (f.comp_h[c] + 7) / 8;
  }
  const long long total = (long long)f.mcus_x * f.mcus_y * f.bpm;
  if (total > 0x7FFFFFFF / 64) return false;
  f.total_blocks = (int)total;
  // jpeg_quality_scaling + jpeg_add_quant_table(force_baseline = TRUE)
  quality = quality < 1 ? 1 : quality > 100 ? 100 : quality;
  const int scale = quality < 50 ? 5000 / quality : 200 - 2 * quality;
  for (int k = 0; k < 2; ++k) {
    const uint8_t* std_zz = k ? kStdQuantChromaZz : kStdQuantLumaZz;
    for (int z = 0; z < 64; ++z) {
      long v = ((long)std_zz[z] * scale + 50) / 100;
      v = v < 1 ? 1 : v > 255 ? 255 : v;
      e->qzz[k][z] = (uint8_t)v;
      e->t.q[k][kNaturalEnc[z]] = (uint16_t)v;
    }
  }
  derive_codes(kStdDcLumaBits, kStdDcLumaVals, 12, e->t.code[0], e->t.len[0]);
  derive_codes(kStdAcLumaBits, kStdAcLumaVals, 162, e->t.code[1], e->t.len[1]);
  derive_codes(kStdDcChromaBits, kStdDcChromaVals, 12, e->t.code[2],
               e->t.len[2]);
  derive_codes(kStdAcChromaBits, kStdAcChromaVals, 162, e->t.code[3],
               e->t.len[3]);
  return true;
}

// SOI .. SOS as jcmarker.c writes them (JFIF 1.01, two DQT, SOF0, four DHT)
inline void enc_headers(const EncHost& e, std::vector<uint8_t>* out) {
  auto put = [&](std::initializer_list<int> b) {
    for (int v : b) out->push_back((uint8_t)v);
  };
  put({0xFF, 0xD8});
  put({0xFF, 0xE0, 0, 16, 'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0});
  for (int k = 0; k < 2; ++k) {
    put({0xFF, 0xDB, 0, 67, k});
    out->insert(out->end(), e.qzz[k], e.qzz[k] + 64);
  }
  const EncFrame& f = e.f;
  put({0xFF, 0xC0, 0, 17, 8, f.height >> 8, f.height & 255, f.width >> 8,
       f.width & 255, 3, 1, (f.hs << 4) | f.vs, 0, 2, 0x11, 1, 3, 0x11, 1});
  struct {
    int id;
    const uint8_t* bits;
    const uint8_t* vals;
    int n;
  } dht[4] = {{0x00, kStdDcLumaBits, kStdDcLumaVals, 12},
              {0x10, kStdAcLumaBits, kStdAcLumaVals, 162},
              {0x01, kStdDcChromaBits, kStdDcChromaVals, 12},
              {0x11, kStdAcChromaBits, kStdAcChromaVals, 162}};
  for (auto& d : dht) {
    const int len = 2 + 1 + 16 + d.n;
    put({0xFF, 0xC4, len >> 8, len & 255, d.id});
    out->insert(out->end(), d.bits, d.bits + 16);
    out->insert(out->end(), d.vals, d.vals + d.n);
  }
  put({0xFF, 0xDA, 0, 12, 3, 1, 0x00, 2, 0x11, 3, 0x11, 0, 63, 0});
}

}  // namespace jpg
}  // namespace xp
