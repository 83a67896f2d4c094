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

struct EncHost {
  EncFrame f;
  EncTabs t;
  uint8_t qzz[2][64];  // the tables as DQT carries them (zigzag order)
  // the Huffman tables as DHT carries them: DC luma, AC luma, DC chroma, AC
  // chroma (the order of EncTabs)
  uint8_t hbits[4][16];
  uint8_t hvals[4][256];
  int hcount[4];
};

// jchuff.c jpeg_gen_optimal_table: code lengths from the symbol frequencies
// of one table (freq[256] is the reserved pseudo-symbol), limited to 16 bits.
inline void gen_optimal_table(const long freq_in[257], uint8_t bits_out[16],
                              uint8_t* vals, int* nvals) {
  enum { MAX_CLEN = 32 };
  long freq[257];
  int codesize[257], others[257];
  uint8_t bits[MAX_CLEN + 1];
  memset(bits, 0, sizeof(bits));
  memset(codesize, 0, sizeof(codesize));
  for (int i = 0; i < 257; ++i) {
    others[i] = -1;
    freq[i] = freq_in[i];
  }
  freq[256] = 1;  // guarantees that no real symbol gets the all-ones code
  for (;;) {
    int c1 = -1, c2 = -1;
    long v = 1000000000L;
    for (int i = 0; i <= 256; ++i)
      if (freq[i] && freq[i] <= v) v = freq[i], c1 = i;
    v = 1000000000L;
    for (int i = 0; i <= 256; ++i)
      if (freq[i] && freq[i] <= v && i != c1) v = freq[i], c2 = i;
    if (c2 < 0) break;
    freq[c1] += freq[c2];
    freq[c2] = 0;
    codesize[c1]++;
    while (others[c1] >= 0) {
      c1 = others[c1];
      codesize[c1]++;
    }
    others[c1] = c2;
    codesize[c2]++;
    while (others[c2] >= 0) {
      c2 = others[c2];
      codesize[c2]++;
    }
  }
  for (int i = 0; i <= 256; ++i)
    if (codesize[i]) bits[codesize[i] > MAX_CLEN ? MAX_CLEN : codesize[i]]++;
  int i;
  for (i = MAX_CLEN; i > 16; --i)
    while (bits[i] > 0) {
      int j = i - 2;
      while (bits[j] == 0) --j;
      bits[i] -= 2;
      bits[i - 1]++;
      bits[j + 1] += 2;
      bits[j]--;
    }
  while (bits[i] == 0) --i;
  bits[i]--;  // the pseudo-symbol's code
  memcpy(bits_out, bits + 1, 16);
  int p = 0;
  for (int l = 1; l <= MAX_CLEN; ++l)
    for (int j = 0; j <= 255; ++j)
      if (codesize[j] == l) vals[p++] = (uint8_t)j;
  *nvals = p;
}

// install the four tables of e->hbits / hvals / hcount as the codes to use
inline void enc_derive(EncHost* e) {
  for (int k = 0; k < 4; ++k)
    derive_codes(e->hbits[k], e->hvals[k], e->hcount[k], e->t.code[k],
                 e->t.len[k]);
}
// ... from symbol frequencies (optimize_coding): hist[4][257]
inline void enc_optimise(const long hist[4][257], EncHost* e) {
  for (int k = 0; k < 4; ++k)
    gen_optimal_table(hist[k], e->hbits[k], e->hvals[k], &e->hcount[k]);
  enc_derive(e);
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
  const uint8_t* sb[4] = {kStdDcLumaBits, kStdAcLumaBits, kStdDcChromaBits,
                          kStdAcChromaBits};
  const uint8_t* sv[4] = {kStdDcLumaVals, kStdAcLumaVals, kStdDcChromaVals,
                          kStdAcChromaVals};
  const int sn[4] = {12, 162, 12, 162};
  for (int k = 0; k < 4; ++k) {
    memcpy(e->hbits[k], sb[k], 16);
    memcpy(e->hvals[k], sv[k], (size_t)sn[k]);
    e->hcount[k] = sn[k];
  }
  enc_derive(e);
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
  const int ids[4] = {0x00, 0x10, 0x01, 0x11};
  for (int k = 0; k < 4; ++k) {
    const int len = 2 + 1 + 16 + e.hcount[k];
    put({0xFF, 0xC4, len >> 8, len & 255, ids[k]});
    out->insert(out->end(), e.hbits[k], e.hbits[k] + 16);
    out->insert(out->end(), e.hvals[k], e.hvals[k] + e.hcount[k]);
  }
  put({0xFF, 0xDA, 0, 12, 3, 1, 0x00, 2, 0x11, 3, 0x11, 0, 63, 0});
}

}  // namespace jpg
}  // namespace xp
