// Per-sample arithmetic of the multiband pyramids, in the reference's exact
// operation order (every multiply and add rounded separately; the library is
// compiled with -fmad=false). Shared by the CUDA kernels (pyramid.cu) and the
// host-side checks in tests/emu.
//
// Spec: external/multiblend/src/pyramid.cpp:584-823 (Shrink / Squeeze),
// :828-1038 (LaplaceCollapse / LaplaceExpand[Shifted] / LaplaceLine2/3),
// :1563-1681 (Out, dither), src/functions.cpp:24-314 (mask pyramid).
#pragma once

#include <cstddef>
#include <cstdint>

#include "hd.h"

namespace xp {
namespace pm {

XP_HD int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// ---------------------------------------------------------------------------
// Shrink: lo[y][j] from hi (hh rows, pitch hp, x/y parity shifts xs, ys).
// ---------------------------------------------------------------------------
struct VForm {
  int kind;           // 0 copy, 1 (A*11+B*4)+C, 2 A*15+B, 3 five-tap, 4 A+B*15
  int r[5];
  float mul;
};

XP_HD VForm shrink_vform(int hh, int ys, int lo_h, int y) {
  VForm f;
  f.mul = 1.0f / 256;
  const int height_odd = (hh & 1) ^ ys;
  const int first_bad = lo_h - (3 - height_odd);
  const int last = hh - 1;
  const int c = 2 * y - 2 - ys;
  f.r[0] = f.r[1] = f.r[2] = f.r[3] = f.r[4] = 0;
  if (y == 0) {  // pyramid.cpp:632-635
    f.kind = 0;
    f.r[0] = 0;
    f.mul = 1.0f / 16;
  } else if (y == 1 && !ys) {  // :639-644
    f.kind = 1;
    f.r[0] = 0;
    f.r[1] = 1;
    f.r[2] = 2;
  } else if (y == 1 && ys) {  // :651-654
    f.kind = 2;
    f.r[0] = 0;
    f.r[1] = 1;
  } else if (y < first_bad || (y == first_bad && height_odd == 0)) {
    f.kind = 3;  // :659-670, :690-701, :710-719
    f.r[0] = c - 2 < 0 ? 0 : c - 2;
    f.r[1] = c - 1 < 0 ? 0 : c - 1;
    f.r[2] = c;
    f.r[3] = c + 1 > last ? last : c + 1;
    f.r[4] = c + 2 > last ? last : c + 2;
  } else if (y == lo_h - 2 && height_odd == 0) {  // :727-731
    f.kind = 4;
    f.r[0] = last - 1;
    f.r[1] = last;
  } else if (y == lo_h - 2) {  // :734-739
    f.kind = 1;
    f.r[0] = last;
    f.r[1] = last - 1;
    f.r[2] = last - 2;
  } else {  // :749-751
    f.kind = 0;
    f.r[0] = last;
    f.mul = 1.0f / 16;
  }
  for (int k = 0; k < 5; ++k)
    if (f.r[k] < 0) f.r[k] = 0;  // degenerate 1-2 row levels
  return f;
}

XP_HD float shrink_vline(const float* hi, int hp, const VForm& f, int col) {
#define XP_HI(k) hi[(size_t)f.r[k] * hp + col]
  switch (f.kind) {
    case 0:
      return XP_HI(0);
    case 1:
      return (XP_HI(0) * 11.0f + XP_HI(1) * 4.0f) + XP_HI(2);
    case 2:
      return XP_HI(0) * 15.0f + XP_HI(1);
    case 3:
      return ((XP_HI(0) + XP_HI(4)) + (XP_HI(1) + XP_HI(3)) * 4.0f) +
             XP_HI(2) * 6.0f;
    default:
      return XP_HI(0) + XP_HI(1) * 15.0f;
  }
#undef XP_HI
}

// column of the (unshifted) hi line that tap k of lo sample j reads
XP_HD int squeeze_col(int j, int k, int hp, int xs) {
  int idx = 2 * j - 4 + k;
  if (idx < 0) idx = 0;
  if (idx > hp - 1) idx = hp - 1;
  idx -= xs;
  if (idx < 0) idx = 0;
  return idx;
}

XP_HD float squeeze_combine(const float t[5], float mul) {
  const float d = t[0] + t[4];
  const float hh = 4.0f * (t[1] + t[3]);
  const float jj = 6.0f * t[2];
  return (d + (hh + jj)) * mul;
}

XP_HD float shrink_sample(const float* hi, int hh, int hp, int xs, int ys,
                          int lo_h, int j, int y) {
  const VForm f = shrink_vform(hh, ys, lo_h, y);
  float t[5];
  for (int k = 0; k < 5; ++k)
    t[k] = shrink_vline(hi, hp, f, squeeze_col(j, k, hp, xs));
  return squeeze_combine(t, f.mul);
}

// ---------------------------------------------------------------------------
// Expand (one horizontal sample of the upsampled lo row)
// ---------------------------------------------------------------------------
XP_HD float lo_at(const float* lo, int lo_pitch, int k) {
  return lo[k < lo_pitch ? k : lo_pitch - 1];
}

XP_HD float expand_h(const float* lo, int lo_pitch, int x, int shifted) {
  const int m = x >> 1;
  if (!shifted) {  // LaplaceExpand :851-896
    if ((x & 1) == 0) {
      const float a = 0.125f * lo_at(lo, lo_pitch, m);
      const float b = 0.75f * lo_at(lo, lo_pitch, m + 1);
      const float c = 0.125f * lo_at(lo, lo_pitch, m + 2);
      return ((x & 3) == 0) ? (a + b) + c : a + (b + c);
    }
    return 0.5f * lo_at(lo, lo_pitch, m + 1) + 0.5f * lo_at(lo, lo_pitch, m + 2);
  }
  // LaplaceExpandShifted :898-946
  if ((x & 1) == 0)
    return 0.5f * lo_at(lo, lo_pitch, m + 1) + 0.5f * lo_at(lo, lo_pitch, m + 2);
  const float a = 0.125f * lo_at(lo, lo_pitch, m + 1);
  const float b = 0.75f * lo_at(lo, lo_pitch, m + 2);
  const float c = 0.125f * lo_at(lo, lo_pitch, m + 3);
  return ((x & 3) == 1) ? a + (b + c) : (a + b) + c;
}

// expand_h for the four samples x = 4i .. 4i+3 at once, from a[k] = lo[2i+k],
// k = 0..4 (no clamping: the caller guarantees 2i+4 < lo_pitch). Same
// association as expand_h for each (x & 3).
XP_HD void expand_h4(const float a[5], int shifted, float e[4]) {
  if (!shifted) {
    e[0] = (0.125f * a[0] + 0.75f * a[1]) + 0.125f * a[2];
    e[1] = 0.5f * a[1] + 0.5f * a[2];
    e[2] = 0.125f * a[1] + (0.75f * a[2] + 0.125f * a[3]);
    e[3] = 0.5f * a[2] + 0.5f * a[3];
  } else {
    e[0] = 0.5f * a[1] + 0.5f * a[2];
    e[1] = 0.125f * a[1] + (0.75f * a[2] + 0.125f * a[3]);
    e[2] = 0.5f * a[2] + 0.5f * a[3];
    e[3] = (0.125f * a[2] + 0.75f * a[3]) + 0.125f * a[4];
  }
}

// vertical part of expand_sample: t = y + ys
XP_HD float expand_v(int t_even, float e0, float e1, float e2) {
  if (t_even) return (e0 + e2) * 0.125f + e1 * 0.75f;  // LaplaceLine3
  return (e1 + e2) * 0.5f;                             // LaplaceLine2
}

// vertical form of Shrink on already loaded rows v[k] = hi[f.r[k]][col]
XP_HD float shrink_vcombine(int kind, float v0, float v1, float v2, float v3,
                            float v4) {
  switch (kind) {
    case 0:
      return v0;
    case 1:
      return (v0 * 11.0f + v1 * 4.0f) + v2;
    case 2:
      return v0 * 15.0f + v1;
    case 3:
      return ((v0 + v4) + (v1 + v3) * 4.0f) + v2 * 6.0f;
    default:
      return v0 + v1 * 15.0f;
  }
}

// expand(lo) at hi sample (x, y); hi level has shifts (xs, ys)
XP_HD float expand_sample(const float* lo, int lo_pitch, int xs, int ys, int x,
                          int y) {
  const int t = y + ys;
  const int j = (t >> 1) + 1;
  const float* r1 = lo + (size_t)j * lo_pitch;
  const float* r2 = r1 + lo_pitch;
  const float e1 = expand_h(r1, lo_pitch, x, xs);
  const float e2 = expand_h(r2, lo_pitch, x, xs);
  if ((t & 1) == 0) {  // LaplaceLine3 :960-975
    const float e0 = expand_h(r1 - lo_pitch, lo_pitch, x, xs);
    return (e0 + e2) * 0.125f + e1 * 0.75f;
  }
  return (e1 + e2) * 0.5f;  // LaplaceLine2 :948-958
}

// ---------------------------------------------------------------------------
// Mask pyramid (dense restatement of Squish / ShrinkMasks)
// ---------------------------------------------------------------------------
XP_HD float mask_h_combine(float a, float b, float c, float d, float e) {
  // inside a run the reference copies the value instead of filtering it
  if (a == b && b == c && c == d && d == e) return c;
  return (b + d) * 0.25f + (a + e + 6 * c) * 0.0625f;
}

XP_HD float mask_v_combine(float v0, float v1, float v2, float v3, float v4) {
  return (v0 + v4) * 0.0625f + (v1 + v3) * 0.25f + v2 * 0.375f;
}

// ---------------------------------------------------------------------------
// Out: dither + clamp (pyramid.cpp:1621-1634, 2054-2066)
// ---------------------------------------------------------------------------
XP_HD float dither_value(int y, int x) {
  const float k[16] = {-0.125f,  0.375f,   0.0f,     0.4999f,
                       0.125f,   -0.375f,  0.25f,    -0.25f,
                       -0.0625f, 0.4375f,  -0.1875f, 0.3125f,
                       0.1875f,  -0.3125f, 0.0625f,  -0.4375f};
  return k[(y & 3) * 4 + (x & 3)];
}

}  // namespace pm
}  // namespace xp
