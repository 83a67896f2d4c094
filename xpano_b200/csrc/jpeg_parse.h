// Host side of the SM JPEG decoder: marker parsing, Huffman table expansion
// and the cut of the entropy-coded scan into subsequences (ITU-T T.81 annex
// B / C / F). Streams it does not cover -- progressive, arithmetic, 12-bit,
// several scans, CMYK, unusual sampling -- are reported as unsupported and
// go to nvJPEG. Plain C++; also compiled into tests/emu/jpeg_emu.cpp.
#pragma once
#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "jpeg_sm.h"

namespace xp {
namespace jpg {

static const uint8_t kNatural[64] = XP_JPG_NATURAL;

struct Parsed {
  Frame f;
  HuffTab huff[4];        // DC 0, DC 1, AC 0, AC 1
  uint16_t quant[3][64];  // per component, natural order
  size_t scan_begin = 0, scan_end = 0;
  std::vector<SubRec> subs;
  std::string why;  // reason when unsupported
};

inline bool build_huff(const uint8_t bits[17], const uint8_t* vals, int nvals,
                       HuffTab* t) {
  memset(t, 0, sizeof(*t));
  int total = 0;
  for (int l = 1; l <= 16; ++l) total += bits[l];
  if (total > 256 || total != nvals) return false;
  memcpy(t->val, vals, (size_t)total);
  int32_t code = 0;
  int p = 0;
  for (int l = 1; l <= 16; ++l) {
    t->valoff[l] = p - code;
    for (int i = 0; i < bits[l]; ++i, ++p, ++code) {
      if (code >= (1 << l)) return false;
      if (l <= kLook) {
        const int lo = code << (kLook - l), n = 1 << (kLook - l);
        for (int j = 0; j < n; ++j)
          t->look[lo + j] = (uint16_t)((l << 8) | vals[p]);
      }
    }
    t->maxcode[l] = bits[l] ? code - 1 : -1;
    code <<= 1;
  }
  t->maxcode[17] = 0xFFFFF;
  return true;
}

inline bool unsupported(Parsed* out, const char* why) {
  out->why = why;
  return false;
}

// sub_bytes: raw bytes per subsequence
inline bool parse(const uint8_t* d, size_t n, int sub_bytes, Parsed* out) {
  Frame& f = out->f;
  memset(&f, 0, sizeof(f));
  if (n < 4 || d[0] != 0xFF || d[1] != 0xD8)
    return unsupported(out, "not a JPEG stream");
  if (n >= 0xFFFF0000u) return unsupported(out, "stream too large");
  uint16_t qt[4][64];
  bool have_qt[4] = {false, false, false, false};
  bool have_huff[4] = {false, false, false, false};
  int comp_id[3] = {0, 0, 0}, comp_q[3] = {0, 0, 0};
  bool have_frame = false;
  int adobe_transform = -1;  // 1 = YCbCr; 0 with three components = RGB
  size_t p = 2;
  for (;;) {
    if (p + 4 > n) return unsupported(out, "truncated header");
    if (d[p] != 0xFF) return unsupported(out, "marker expected");
    const uint8_t m = d[p + 1];
    if (m == 0xFF) {  // fill byte
      ++p;
      continue;
    }
    if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) {
      p += 2;
      continue;
    }
    if (m == 0xD9) return unsupported(out, "no scan");
    const size_t len = ((size_t)d[p + 2] << 8) | d[p + 3];
    if (len < 2 || p + 2 + len > n) return unsupported(out, "bad segment");
    const uint8_t* s = d + p + 4;
    const size_t sl = len - 2;
    if (m == 0xC0 || m == 0xC1) {  // baseline / extended sequential Huffman
      if (have_frame) return unsupported(out, "two frames");
      if (sl < 6 || s[0] != 8) return unsupported(out, "not 8-bit samples");
      f.height = (s[1] << 8) | s[2];
      f.width = (s[3] << 8) | s[4];
      f.ncomp = s[5];
      if (f.width <= 0 || f.height <= 0)
        return unsupported(out, "size missing (DNL)");
      if (f.ncomp != 1 && f.ncomp != 3)
        return unsupported(out, "1 or 3 components");
      if (sl < 6 + 3 * (size_t)f.ncomp) return unsupported(out, "bad SOF");
      for (int c = 0; c < f.ncomp; ++c) {
        comp_id[c] = s[6 + 3 * c];
        f.hs[c] = s[7 + 3 * c] >> 4;
        f.vs[c] = s[7 + 3 * c] & 15;
        comp_q[c] = s[8 + 3 * c];
        if (comp_q[c] > 3) return unsupported(out, "bad quantisation index");
      }
      have_frame = true;
    } else if (m >= 0xC2 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
      return unsupported(out, "not a sequential Huffman frame");
    } else if (m == 0xCC) {
      return unsupported(out, "arithmetic coding");
    } else if (m == 0xC4) {  // DHT
      size_t q = 0;
      while (q < sl) {
        if (q + 17 > sl) return unsupported(out, "bad DHT");
        const int tc = s[q] >> 4, th = s[q] & 15;
        uint8_t bits[17] = {0};
        int total = 0;
        for (int l = 1; l <= 16; ++l) total += (bits[l] = s[q + l]);
        if (q + 17 + (size_t)total > sl) return unsupported(out, "bad DHT");
        if (tc > 1 || th > 1)
          return unsupported(out, "Huffman table index above 1");
        if (!build_huff(bits, s + q + 17, total, &out->huff[tc * 2 + th]))
          return unsupported(out, "bad Huffman table");
        have_huff[tc * 2 + th] = true;
        q += 17 + (size_t)total;
      }
    } else if (m == 0xDB) {  // DQT
      size_t q = 0;
      while (q < sl) {
        const int pq = s[q] >> 4, tq = s[q] & 15;
        if (tq > 3) return unsupported(out, "bad DQT");
        const size_t need = pq ? 128 : 64;
        if (q + 1 + need > sl) return unsupported(out, "bad DQT");
        for (int z = 0; z < 64; ++z)
          qt[tq][kNatural[z]] =
              pq ? (uint16_t)((s[q + 1 + 2 * z] << 8) | s[q + 2 + 2 * z])
                 : s[q + 1 + z];
        have_qt[tq] = true;
        q += 1 + need;
      }
    } else if (m == 0xDD) {  // DRI
      if (sl < 2) return unsupported(out, "bad DRI");
      f.restart = (s[0] << 8) | s[1];
    } else if (m == 0xEE) {  // Adobe: colour transform flag
      if (sl >= 12 && memcmp(s, "Adobe", 5) == 0) adobe_transform = s[11];
    } else if (m == 0xDA) {  // SOS
      if (!have_frame) return unsupported(out, "scan before frame");
      if (sl < 1 || s[0] != f.ncomp || sl < 4 + 2 * (size_t)f.ncomp)
        return unsupported(out, "several scans");
      for (int c = 0; c < f.ncomp; ++c) {
        if (s[1 + 2 * c] != comp_id[c])
          return unsupported(out, "scan component order");
        const int td = s[2 + 2 * c] >> 4, ta = s[2 + 2 * c] & 15;
        if (td > 1 || ta > 1 || !have_huff[td] || !have_huff[2 + ta])
          return unsupported(out, "Huffman table missing");
        f.dc_tab[c] = td;
        f.ac_tab[c] = 2 + ta;
      }
      const uint8_t* t = s + 1 + 2 * f.ncomp;
      if (t[0] != 0 || t[1] != 63 || t[2] != 0)
        return unsupported(out, "spectral selection / approximation");
      out->scan_begin = p + 2 + len;
      break;
    }
    p += 2 + len;
  }
  if (f.ncomp == 3) {
    if (adobe_transform >= 0 && adobe_transform != 1)
      return unsupported(out, "Adobe RGB stream");
    if (adobe_transform < 0 && comp_id[0] == 'R' && comp_id[1] == 'G' &&
        comp_id[2] == 'B')
      return unsupported(out, "RGB component ids");
  }
  for (int c = 0; c < f.ncomp; ++c) {
    if (!have_qt[comp_q[c]]) return unsupported(out, "quantisation table missing");
    memcpy(out->quant[c], qt[comp_q[c]], sizeof(qt[0]));
    f.qt[c] = c;
  }
  // geometry
  if (f.ncomp == 1) {
    f.hs[0] = f.vs[0] = 1;  // a single-component scan is not interleaved
    f.chroma = 0;
  } else {
    if (f.hs[1] != 1 || f.vs[1] != 1 || f.hs[2] != 1 || f.vs[2] != 1)
      return unsupported(out, "chroma sampling factors");
    if (f.hs[0] == 1 && f.vs[0] == 1)
      f.chroma = 0;
    else if (f.hs[0] == 2 && f.vs[0] == 1)
      f.chroma = 1;
    else if (f.hs[0] == 2 && f.vs[0] == 2)
      f.chroma = 2;
    else
      return unsupported(out, "luma sampling factors");
  }
  const int hmax = f.hs[0], vmax = f.vs[0];
  f.mcus_x = (f.width + 8 * hmax - 1) / (8 * hmax);
  f.mcus_y = (f.height + 8 * vmax - 1) / (8 * vmax);
  f.bpm = 0;
  for (int c = 0; c < f.ncomp; ++c) {
    f.blk_first[c] = f.bpm;
    for (int y = 0; y < f.vs[c]; ++y)
      for (int x = 0; x < f.hs[c]; ++x) {
        f.blk_comp[f.bpm] = c;
        f.blk_x[f.bpm] = x;
        f.blk_y[f.bpm] = y;
        ++f.bpm;
      }
    f.comp_w[c] = (f.width * f.hs[c] + hmax - 1) / hmax;
    f.comp_h[c] = (f.height * f.vs[c] + vmax - 1) / vmax;
    f.plane_w[c] = f.mcus_x * f.hs[c] * 8;
    f.plane_h[c] = f.mcus_y * f.vs[c] * 8;
  }
  f.fancy = f.ncomp == 3 && f.comp_w[1] > 2;
  const long long total_mcus = (long long)f.mcus_x * f.mcus_y;
  if (total_mcus * f.bpm > 0x7FFFFFFF / 64)
    return unsupported(out, "image too large");
  f.total_blocks = (int)(total_mcus * f.bpm);

  for (int k = 0; k < f.bpm; ++k) {
    f.blk_dc[k] = f.dc_tab[f.blk_comp[k]];
    f.blk_ac[k] = f.ac_tab[f.blk_comp[k]];
  }

  // The end of the scan and the restart intervals, as byte ranges of the
  // DESTUFFED scan (jpeg_sm.h: destuff_keeps).
  struct Seg {
    size_t begin, end;  // destuffed bytes
  };
  std::vector<Seg> segs;
  size_t end = n;
  if (!f.restart) {
    // One interval. Its end only has to be at or after the scan's last byte
    // (decoding stops with the last block), so the last EOI of the stream
    // serves -- found from the back -- and the scan is not searched at all;
    // its destuffed length is bounded by its raw length (the device knows
    // the exact one).
    const size_t stop =
        n > 65536 + out->scan_begin ? n - 65536 : out->scan_begin;
    for (size_t k = n; k >= stop + 2; --k)
      if (d[k - 2] == 0xFF && d[k - 1] == 0xD9) {
        end = k - 2;
        break;
      }
    segs.push_back({0, end - out->scan_begin});
  } else {
    size_t q = out->scan_begin, kept = 0, seg_begin = 0, last = q;
    while (q < n) {
      const uint8_t* hit =
          static_cast<const uint8_t*>(memchr(d + q, 0xFF, n - q));
      if (!hit) {
        q = n;
        break;
      }
      const size_t at = (size_t)(hit - d);
      if (at + 1 >= n) {
        q = at;
        break;
      }
      const uint8_t m = d[at + 1];
      if (m == 0x00) {  // FF kept, 00 dropped
        kept += at + 1 - last;
        last = q = at + 2;
      } else if (m >= 0xD0 && m <= 0xD7) {  // both dropped
        kept += at - last;
        last = q = at + 2;
        segs.push_back({seg_begin, kept});
        seg_begin = kept;
      } else if (m == 0xFF) {
        return unsupported(out, "fill bytes inside the scan");
      } else {
        q = at;
        break;
      }
    }
    end = q;
    kept += end - last;
    segs.push_back({seg_begin, kept});
  }
  out->scan_end = end;
  if (end <= out->scan_begin) return unsupported(out, "empty scan");
  const long long per = f.restart ? f.restart : total_mcus;
  const long long want = (total_mcus + per - 1) / per;
  if ((long long)segs.size() != want)
    return unsupported(out, "restart markers do not match the interval");
  if ((end - out->scan_begin) >= 0x1FFFFFF0u)
    return unsupported(out, "scan too large");
  out->subs.clear();
  for (size_t k = 0; k < segs.size(); ++k) {
    if (segs[k].end <= segs[k].begin)
      return unsupported(out, "empty restart interval");
    for (size_t b = segs[k].begin; b < segs[k].end; b += (size_t)sub_bytes) {
      SubRec r;
      r.start = (uint32_t)(8 * b);
      r.end = (uint32_t)(8 * std::min(b + (size_t)sub_bytes, segs[k].end));
      r.first_mcu = (uint32_t)(k * per);
      r.is_first = b == segs[k].begin;
      out->subs.push_back(r);
    }
  }
  return true;
}

// Host form of the destuffing kernels (the emulation's): -> big-endian words
// with two zero words behind the last byte; returns the destuffed bytes.
inline size_t destuff(const uint8_t* d, size_t begin, size_t end,
                      std::vector<uint32_t>* words) {
  std::vector<uint8_t> bytes;
  bytes.reserve(end - begin);
  for (size_t i = begin; i < end; ++i)
    if (destuff_keeps(d, i, begin, end)) bytes.push_back(d[i]);
  words->assign(bytes.size() / 4 + 3, 0);
  for (size_t i = 0; i < bytes.size(); ++i)
    (*words)[i / 4] |= (uint32_t)bytes[i] << (24 - 8 * (i % 4));
  return bytes.size();
}

}  // namespace jpg
}  // namespace xp
