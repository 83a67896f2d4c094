// TEST INFRASTRUCTURE: host emulation of the SM JPEG decoder. It compiles the
// SAME headers the CUDA kernels use (xpano_b200/csrc/jpeg_sm.h, jpeg_parse.h)
// and runs the kernels' steps as plain loops -- guess pass, synchronisation
// sweeps, block-index scan, write pass, DC scan, inverse DCT, upsampling and
// colour conversion -- so the CPU suite checks the decoder's arithmetic and
// its synchronisation rule against cv2.imdecode without a GPU.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../xpano_b200/csrc/jpeg_parse.h"

using namespace xp::jpg;

static char g_why[256];

extern "C" const char* emu_jpeg_why() { return g_why; }

// -> 0 ok, 1 unsupported, 2 did not converge. info: {width, height, sweeps,
// subsequences}
extern "C" int emu_jpeg_decode(const uint8_t* d, size_t n, int sub_bytes,
                               uint8_t* bgr, size_t stride, int info[4],
                               int order) {
  Parsed P;
  if (!parse(d, n, sub_bytes, &P)) {
    strncpy(g_why, P.why.c_str(), sizeof(g_why) - 1);
    return 1;
  }
  const Frame& f = P.f;
  info[0] = f.width;
  info[1] = f.height;
  info[3] = (int)P.subs.size();
  if (!bgr) return 0;
  const int ns = (int)P.subs.size();
  std::vector<uint32_t> words;
  const uint32_t total_bits =
      (uint32_t)(8 * destuff(d, P.scan_begin, P.scan_end, &words));
  auto sub_end = [&](int i) {
    return P.subs[i].end < total_bits ? P.subs[i].end : total_bits;
  };
  std::vector<uint64_t> exit_state(ns);
  std::vector<int> done(ns);
  auto start_of = [&](int i, bool guess) {
    State s;
    const SubRec& r = P.subs[i];
    if (r.is_first || guess) {
      s.p = r.start;
      s.b = 0, s.z = 0;
    } else {
      s = unpack(exit_state[i - 1]);
    }
    return s;
  };
  // guess pass
  for (int i = 0; i < ns; ++i) {
    State s = start_of(i, true);
    done[i] = decode_subsequence<false>(words.data(), sub_end(i), f, P.huff, &s, nullptr, nullptr, 0, 0);
    exit_state[i] = pack(s);
  }
  // sweeps until nothing changes; order 0: every subsequence reads the
  // previous sweep's states, 1: ascending in place, 2: descending in place
  int sweeps = 0;
  for (;; ++sweeps) {
    if (sweeps > ns + 2) return 2;
    bool changed = false;
    std::vector<uint64_t> prev;
    if (order == 0) prev = exit_state;
    for (int j = 0; j < ns; ++j) {
      const int i = order == 2 ? ns - 1 - j : j;
      State s;
      if (P.subs[i].is_first || order != 0)
        s = start_of(i, false);
      else
        s = unpack(prev[i - 1]);
      done[i] = decode_subsequence<false>(words.data(), sub_end(i), f, P.huff,
                                          &s, nullptr, nullptr, 0, 0);
      if (pack(s) != exit_state[i]) changed = true;
      exit_state[i] = pack(s);
    }
    if (!changed) break;
  }
  info[2] = sweeps;
  // block index of every subsequence's first symbol
  std::vector<int> base(ns);
  int run = 0;
  for (int i = 0; i < ns; ++i) {
    if (P.subs[i].is_first) run = (int)P.subs[i].first_mcu * f.bpm;
    base[i] = run;
    run += done[i];
  }
  std::vector<int16_t> coef((size_t)f.total_blocks * 64, 0);
  std::vector<int16_t> dcs((size_t)f.total_blocks, 0);
  const long long per = f.restart ? f.restart : (long long)f.mcus_x * f.mcus_y;
  for (int i = 0; i < ns; ++i) {
    State s = start_of(i, false);
    long long limit = ((long long)P.subs[i].first_mcu + per) * f.bpm;
    if (limit > f.total_blocks) limit = f.total_blocks;
    decode_subsequence<true>(words.data(), sub_end(i), f, P.huff, &s,
                             coef.data(), dcs.data(), base[i], (int)limit);
  }
  // DC prediction
  int pred[3] = {0, 0, 0};
  const int mcus = f.mcus_x * f.mcus_y;
  for (int m = 0; m < mcus; ++m) {
    if (f.restart && m % f.restart == 0) pred[0] = pred[1] = pred[2] = 0;
    for (int b = 0; b < f.bpm; ++b) {
      int16_t& dc = dcs[(size_t)m * f.bpm + b];
      pred[f.blk_comp[b]] += dc;
      dc = (int16_t)pred[f.blk_comp[b]];
    }
  }
  // inverse DCT into the component planes
  std::vector<uint8_t> planes[3];
  const uint8_t* pl[3] = {nullptr, nullptr, nullptr};
  int pitch[3] = {0, 0, 0};
  for (int c = 0; c < f.ncomp; ++c) {
    planes[c].assign((size_t)f.plane_w[c] * f.plane_h[c], 0);
    pitch[c] = f.plane_w[c];
  }
  for (int m = 0; m < mcus; ++m) {
    const int mx = m % f.mcus_x, my = m / f.mcus_x;
    for (int b = 0; b < f.bpm; ++b) {
      const int c = f.blk_comp[b];
      uint8_t px[64];
      idct_islow(&coef[((size_t)m * f.bpm + b) * 64],
                 dcs[(size_t)m * f.bpm + b], P.quant[c], px);
      const int bx = mx * f.hs[c] + f.blk_x[b], by = my * f.vs[c] + f.blk_y[b];
      for (int r = 0; r < 8; ++r)
        memcpy(&planes[c][(size_t)(by * 8 + r) * pitch[c] + bx * 8],
               px + 8 * r, 8);
    }
  }
  for (int c = 0; c < f.ncomp; ++c) pl[c] = planes[c].data();
  for (int y = 0; y < f.height; ++y)
    for (int x = 0; x < f.width; ++x)
      pixel_bgr(f, pl, pitch, x, y, bgr + (size_t)y * stride + 3 * (size_t)x);
  return 0;
}

// ---------------------------------------------------------------------------
// The encoder: the same steps as xpano_b200/csrc/jpeg_enc.cu, as loops.
// -> bytes written (0: unsupported or `cap` too small)
// ---------------------------------------------------------------------------
#include "../../xpano_b200/csrc/jpeg_enc_host.h"

extern "C" size_t emu_jpeg_encode(const uint8_t* bgr, int w, int h,
                                  size_t stride, int quality, int subsampling,
                                  int optimize, uint8_t* out, size_t cap) {
  EncHost E;
  if (!enc_setup(w, h, quality, subsampling, &E)) return 0;
  const EncFrame& f = E.f;
  // component planes with libjpeg's edge expansion
  std::vector<uint8_t> plane[3];
  int pw[3], ph[3];
  for (int c = 0; c < 3; ++c) {
    pw[c] = f.blocks_w[c] * 8;
    ph[c] = f.blocks_h[c] * 8;
    plane[c].assign((size_t)pw[c] * ph[c], 0);
  }
  auto ycc = [&](int x, int y, int* Y, int* Cb, int* Cr) {
    x = x < w - 1 ? x : w - 1;
    y = y < h - 1 ? y : h - 1;
    const uint8_t* p = bgr + (size_t)y * stride + 3 * (size_t)x;
    bgr_to_ycc(p[0], p[1], p[2], Y, Cb, Cr);
  };
  for (int y = 0; y < ph[0]; ++y)
    for (int x = 0; x < pw[0]; ++x) {
      int Y, Cb, Cr;
      ycc(x, y, &Y, &Cb, &Cr);
      plane[0][(size_t)y * pw[0] + x] = (uint8_t)Y;
    }
  for (int cy = 0; cy < ph[1]; ++cy)
    for (int cx = 0; cx < pw[1]; ++cx) {
      const int r = cy < f.comp_h[1] - 1 ? cy : f.comp_h[1] - 1;
      int sb = 0, sr = 0;
      for (int dy = 0; dy < f.vs; ++dy)
        for (int dx = 0; dx < f.hs; ++dx) {
          int Y, Cb, Cr;
          ycc(cx * f.hs + dx, r * f.vs + dy, &Y, &Cb, &Cr);
          sb += Cb;
          sr += Cr;
        }
      int vb = sb, vr = sr;
      if (f.hs == 2 && f.vs == 1) {
        vb = (sb + (cx & 1)) >> 1;
        vr = (sr + (cx & 1)) >> 1;
      } else if (f.hs == 2 && f.vs == 2) {
        vb = (sb + 1 + (cx & 1)) >> 2;
        vr = (sr + 1 + (cx & 1)) >> 2;
      }
      plane[1][(size_t)cy * pw[1] + cx] = (uint8_t)vb;
      plane[2][(size_t)cy * pw[2] + cx] = (uint8_t)vr;
    }
  // blocks in scan order
  std::vector<int16_t> coef((size_t)f.total_blocks * 64, 0);
  for (int blk = 0; blk < f.total_blocks; ++blk) {
    const BlockPos p = block_pos(f, blk);
    if (p.dummy) continue;
    uint8_t px[64];
    for (int r = 0; r < 8; ++r)
      memcpy(px + 8 * r,
             &plane[p.comp][(size_t)(p.by * 8 + r) * pw[p.comp] + p.bx * 8], 8);
    fdct_quantise(px, E.t.q[p.comp ? 1 : 0], &coef[(size_t)blk * 64]);
  }
  for (int blk = 0; blk < f.total_blocks; ++blk)
    if (block_pos(f, blk).dummy)
      coef[(size_t)blk * 64] = coef[(size_t)dummy_source(f, blk) * 64];
  // entropy coding
  std::vector<uint8_t> scan;
  uint64_t acc = 0;
  int nacc = 0;
  auto emit = [&](uint32_t bits, int n) {
    acc = (acc << n) | bits;
    nacc += n;
    while (nacc >= 8) {
      const uint8_t b = (uint8_t)(acc >> (nacc - 8));
      scan.push_back(b);
      if (b == 0xFF) scan.push_back(0);
      nacc -= 8;
    }
  };
  int pred[3] = {0, 0, 0};
  if (optimize) {  // symbol statistics first, tables from them
    long hist[4][257];
    memset(hist, 0, sizeof(hist));
    for (int blk = 0; blk < f.total_blocks; ++blk) {
      const int c = block_pos(f, blk).comp;
      const int16_t* b = &coef[(size_t)blk * 64];
      block_symbols(b, pred[c], c ? 2 : 0, c ? 3 : 1,
                    [&](int tab, int sym, uint32_t, int) { hist[tab][sym]++; });
      pred[c] = b[0];
    }
    enc_optimise(hist, &E);
    pred[0] = pred[1] = pred[2] = 0;
  }
  for (int blk = 0; blk < f.total_blocks; ++blk) {
    const int c = block_pos(f, blk).comp;
    const int16_t* b = &coef[(size_t)blk * 64];
    encode_block(b, pred[c], E.t, c ? 2 : 0, c ? 3 : 1, emit);
    pred[c] = b[0];
  }
  if (nacc) emit(0x7F, 7);  // pad the last byte with ones
  std::vector<uint8_t> head;
  enc_headers(E, &head);
  const size_t total = head.size() + scan.size() + 2;
  if (!out || cap < total) return out ? 0 : total;
  memcpy(out, head.data(), head.size());
  memcpy(out + head.size(), scan.data(), scan.size());
  out[total - 2] = 0xFF;
  out[total - 1] = 0xD9;
  return total;
}
