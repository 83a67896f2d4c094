// The four raster-scan recurrences of the blend path, as problems for the
// wavefront driver (wavefront.h). Each step() is the reference's per-pixel
// update restated on (own-row predecessor, three values of the previous row).
//
//   InpaintFwd / InpaintBwd : external/multiblend/src/image.cpp:550-812
//   SeamBwd / SeamFwd       : external/multiblend/src/multiblend.cpp:254-823
//
// Values are u64. Seams: dist << 32 | image id, top bit = "that image does
// not cover this pixel" (maskval, multiblend.cpp:74-76). Inpaint: dist << 32 |
// index of the source pixel whose colour the reference would have copied here
// (the reference copies the whole pixel in place from the winning neighbour;
// carrying the winner's source index and gathering once at the end is the
// same function of the input).
#pragma once

#include <cstdint>

#include "wavefront.h"

namespace xp {

typedef uint64_t u64;
typedef uint32_t u32;

constexpr u64 kTopBit = 0x8000000000000000ull;
constexpr u64 kLow63 = 0x7fffffffffffffffull;
constexpr u64 kDtMax = 0x9000000000000000ull;
XP_HD u64 d32(u32 d) { return (u64)d << 32; }

template <class C>
struct CoverTraits;
template <>
struct CoverTraits<u32> {
  static constexpr u32 rs_bit = 0x80000000u;
  static constexpr int max_images = 31;
};
template <>
struct CoverTraits<u64> {
  static constexpr u64 rs_bit = 0x8000000000000000ull;
  static constexpr int max_images = 63;
};

XP_HD int popc_c(u32 v) {
#if defined(__CUDA_ARCH__)
  return __popc(v);
#else
  return __builtin_popcount(v);
#endif
}
XP_HD int popc_c(u64 v) {
#if defined(__CUDA_ARCH__)
  return __popcll(v);
#else
  return __builtin_popcountll(v);
#endif
}
XP_HD int ctz_c(u32 v) {
#if defined(__CUDA_ARCH__)
  return __ffs((int)v) - 1;
#else
  return __builtin_ctz(v);
#endif
}
XP_HD int ctz_c(u64 v) {
#if defined(__CUDA_ARCH__)
  return __ffsll((long long)v) - 1;
#else
  return __builtin_ctzll(v);
#endif
}
XP_HD int clz_top(u32 v) {  // index of the highest set bit
#if defined(__CUDA_ARCH__)
  return 31 - __clz((int)v);
#else
  return 31 - __builtin_clz(v);
#endif
}
XP_HD int clz_top(u64 v) {
#if defined(__CUDA_ARCH__)
  return 63 - __clzll((long long)v);
#else
  return 63 - __builtin_clzll(v);
#endif
}

// maskval (multiblend.cpp:74-76): re-evaluate the cover flag of v's label
template <class C>
XP_HD u64 maskval(u64 v, C bits) {
  const u32 lab = (u32)v;
  const bool covered = (bits >> (lab & (sizeof(C) * 8 - 1))) & 1;
  return (v & kLow63) | (covered ? 0 : kTopBit);
}

struct NoRow {};

// ---------------------------------------------------------------------------
// Seams, backward pass: rows bottom -> top, x right -> left.
// Processing coordinates: q = W-1-x, rho = H-1-y.
// ---------------------------------------------------------------------------
template <class C>
struct SeamBwd {
  typedef C In;
  struct Row {
    bool first_run;  // still inside the first run of the bottom row
  };
  int n_bands, W, H;
  const C* cover;  // H x W, bit i = image i opaque, top bit = run start
  u64* out;        // H x W backward values

  XP_HD wf::Band band(int i) const {
    return wf::Band{0, i * wf::kLanes, W, H, i - 1};
  }
  XP_HD Row row_init(const wf::Band&, int) const { return Row{true}; }
  XP_HD size_t at(int rho, int q) const {
    return (size_t)(H - 1 - rho) * W + (W - 1 - q);
  }
  XP_HD In load_in(const wf::Band&, int rho, int q) const {
    return cover[at(rho, q)];
  }
  XP_HD u64 load_val(const wf::Band&, int rho, int q) const {
    return out[at(rho, q)];
  }
  XP_HD void store(const wf::Band&, int rho, int q, u64 v) const {
    out[at(rho, q)] = v;
  }
  // last = (x+1, y); w0 = (x+1, y+1); w1 = (x, y+1); w2 = (x-1, y+1)
  XP_HD u64 step(const wf::Band&, Row& row, int rho, int q, In cov, u64 last,
                 u64 w0, u64 w1, u64 w2) const {
    const C bits = cov & ~CoverTraits<C>::rs_bit;
    const bool was_first = row.first_run;
    if (cov & CoverTraits<C>::rs_bit) row.first_run = false;
    if (popc_c(bits) == 1) return (u64)ctz_c(bits);  // :327-331
    if (rho == 0) {                                  // bottom row :333-346
      if (was_first) return kDtMax;
      return maskval(last, bits) + d32(3);
    }
    u64 d = ~0ull, t;
    if (q > 0) {
      d = maskval(last + d32(3), bits);
      t = maskval(w0 + d32(4), bits);
      if (t < d) d = t;
    }
    if (q < W - 1) {
      t = maskval(w2 + d32(4), bits);
      if (t < d) d = t;
    }
    t = maskval(w1 + d32(3), bits);
    if (t < d) d = t;
    return d;
  }
};

// ---------------------------------------------------------------------------
// Seams, forward pass: rows top -> bottom, x left -> right. q = x, rho = y.
// ---------------------------------------------------------------------------
template <class C>
struct SeamFwdIn {
  u64 bw;
  C cov;
};

template <class C>
struct SeamFwd {
  typedef SeamFwdIn<C> In;
  typedef NoRow Row;
  int n_bands, W, H, reverse;
  const C* cover;
  const u64* bwd;
  u64* out;         // H x W forward values (also the band hand-over rows)
  uint8_t* labels;  // H x W owner image

  XP_HD wf::Band band(int i) const {
    return wf::Band{0, i * wf::kLanes, W, H, i - 1};
  }
  XP_HD Row row_init(const wf::Band&, int) const { return Row{}; }
  XP_HD In load_in(const wf::Band&, int rho, int q) const {
    const size_t p = (size_t)rho * W + q;
    return In{bwd[p], cover[p]};
  }
  XP_HD u64 load_val(const wf::Band&, int rho, int q) const {
    return out[(size_t)rho * W + q];
  }
  XP_HD void store(const wf::Band&, int rho, int q, u64 v) const {
    const size_t p = (size_t)rho * W + q;
    out[p] = v;
    labels[p] = (uint8_t)(v & 0xffu);
  }
  // last = (x-1, y); w0 = (x-1, y-1); w1 = (x, y-1); w2 = (x+1, y-1)
  XP_HD u64 step(const wf::Band&, Row&, int y, int x, In in, u64 last, u64 w0,
                 u64 w1, u64 w2) const {
    const C bits = in.cov & ~CoverTraits<C>::rs_bit;
    const bool rs = (in.cov & CoverTraits<C>::rs_bit) != 0;
    const int cnt = popc_c(bits);
    if (cnt == 1) return (u64)ctz_c(bits);  // :607-616
    u64 best = in.bw, t;                    // ReadSeamDT, unmasked
    if (y > 0) {
      if (x > 0) {
        t = maskval(w0 + d32(4), bits);
        if (t < best) best = t;
      }
      t = maskval(w1 + d32(3), bits);
      if (t < best) best = t;
      if (x < W - 1) {
        t = maskval(w2 + d32(4), bits);
        if (t < best) best = t;
      }
    }
    if (x > 0) {
      // re-masked only on the top row, at x == 1 and at a run start
      // (:629-630, :704-705); raw inside a run (:750)
      t = last + d32(3);
      if (y == 0 || x == 1 || rs) t = maskval(t, bits);
      if (t < best) best = t;
    }
    if ((best & kTopBit) && cnt != 0)  // arbitrary seam :732-742
      best = kTopBit | (u64)(reverse ? clz_top(bits) : ctz_c(bits));
    return best;
  }
};

// ---------------------------------------------------------------------------
// Inpaint. One job per image; the bbox (left, top, w, h) comes from trim.
// ---------------------------------------------------------------------------
struct InpaintJob {
  const void* data;  // interleaved 4-sample pixels of the untrimmed image
  int tiff_w;        // pixels per row of `data`
  int left, top, w, h;
  u64* fwd;          // h x w forward values
  u64* bwd;          // h x w final values (low 32 bits = source index)
  int first_band;    // index of this job's band 0 in the band table
};

// band table entry: job and band-within-job, ordered band-major so that a
// band's predecessor always has a smaller global index
struct InpaintBandRef {
  int job, local, dep;
};

template <int BPP>
XP_HD bool px_opaque(const void* data, size_t idx) {
  if (BPP == 8)
    return (static_cast<const u32*>(data)[idx] >> 24) == 0xffu;
  return (static_cast<const u64*>(data)[idx] >> 48) == 0xffffull;
}

// forward pass: q = x, rho = y (image.cpp:550-694)
template <int BPP>
struct InpaintFwd {
  typedef u32 In;  // 1 = opaque
  typedef NoRow Row;
  int n_bands;
  const InpaintJob* jobs;
  const InpaintBandRef* bands;

  XP_HD wf::Band band(int i) const {
    const InpaintBandRef r = bands[i];
    const InpaintJob& j = jobs[r.job];
    return wf::Band{r.job, r.local * wf::kLanes, j.w, j.h, r.dep};
  }
  XP_HD Row row_init(const wf::Band&, int) const { return Row{}; }
  XP_HD u32 self(const InpaintJob& j, int y, int x) const {
    return (u32)((size_t)(j.top + y) * j.tiff_w + j.left + x);
  }
  XP_HD In load_in(const wf::Band& B, int y, int x) const {
    const InpaintJob& j = jobs[B.job];
    return px_opaque<BPP>(j.data, self(j, y, x)) ? 1u : 0u;
  }
  XP_HD u64 load_val(const wf::Band& B, int y, int x) const {
    return jobs[B.job].fwd[(size_t)y * B.W + x];
  }
  XP_HD void store(const wf::Band& B, int y, int x, u64 v) const {
    jobs[B.job].fwd[(size_t)y * B.W + x] = v;
  }
  // last = (x-1, y); w0 = (x-1, y-1); w1 = (x, y-1); w2 = (x+1, y-1)
  XP_HD u64 step(const wf::Band& B, Row&, int y, int x, In opaque, u64 last,
                 u64 w0, u64 w1, u64 w2) const {
    const u32 me = self(jobs[B.job], y, x);
    if (opaque) return (u64)me;  // dist 0 (:659-662)
    u32 d, src;
    if (y == 0) {  // :573-583
      if (x == 0) return d32(0x80000000u) | me;
      d = (u32)(last >> 32) + 3;
      src = (u32)last;
    } else if (x == 0) {  // :585-603, tie goes to up-right
      const u32 b = (u32)(w1 >> 32) + 3;
      const u32 c = B.W > 1 ? (u32)(w2 >> 32) + 4 : 0xffffffffu;
      if (b < c) {
        d = b;
        src = (u32)w1;
      } else {
        d = c;
        src = (u32)w2;
      }
    } else {  // :609-641
      d = (u32)(last >> 32) + 3;
      src = (u32)last;
      const u32 a = (u32)(w0 >> 32) + 4;
      const u32 b = (u32)(w1 >> 32) + 3;
      const u32 c = x < B.W - 1 ? (u32)(w2 >> 32) + 4 : 0xffffffffu;
      if (a < d) {
        d = a;
        src = (u32)w0;
      }
      if (b < d) {
        d = b;
        src = (u32)w1;
      }
      if (c < d) {
        d = c;
        src = (u32)w2;
      }
    }
    return d32(d) | src;
  }
};

// backward pass: q = w-1-x, rho = h-1-y (image.cpp:711-812)
struct InpaintBwd {
  typedef u64 In;  // forward value of this pixel (dist 0 <=> opaque)
  typedef NoRow Row;
  int n_bands;
  const InpaintJob* jobs;
  const InpaintBandRef* bands;

  XP_HD wf::Band band(int i) const {
    const InpaintBandRef r = bands[i];
    const InpaintJob& j = jobs[r.job];
    return wf::Band{r.job, r.local * wf::kLanes, j.w, j.h, r.dep};
  }
  XP_HD Row row_init(const wf::Band&, int) const { return Row{}; }
  XP_HD size_t at(const wf::Band& B, int rho, int q) const {
    return (size_t)(B.H - 1 - rho) * B.W + (B.W - 1 - q);
  }
  XP_HD In load_in(const wf::Band& B, int rho, int q) const {
    return jobs[B.job].fwd[at(B, rho, q)];
  }
  XP_HD u64 load_val(const wf::Band& B, int rho, int q) const {
    return jobs[B.job].bwd[at(B, rho, q)];
  }
  XP_HD void store(const wf::Band& B, int rho, int q, u64 v) const {
    jobs[B.job].bwd[at(B, rho, q)] = v;
  }
  // last = (x+1, y); w0 = (x+1, y+1); w1 = (x, y+1); w2 = (x-1, y+1)
  XP_HD u64 step(const wf::Band& B, Row&, int rho, int q, In fwd, u64 last,
                 u64 w0, u64 w1, u64 w2) const {
    u32 best = (u32)(fwd >> 32);
    u32 src = (u32)fwd;
    if (best == 0) return fwd;  // opaque: cur = 0, d = 3
    // candidate from the right neighbour: its final distance + 3; at the row
    // start the reference seeds d with 0x80000000 (or nothing on the last line)
    const bool has_right = q > 0;
    const u32 d = has_right ? (u32)(last >> 32) + 3 : 0x80000000u;
    if (rho == 0) {  // last line :711-747
      if (has_right && d < best) {
        best = d;
        src = (u32)last;
      }
      return d32(best) | src;
    }
    const u32 a = q < B.W - 1 ? (u32)(w2 >> 32) + 4 : 0x80000000u;  // x-1
    const u32 b = (u32)(w1 >> 32) + 3;
    const u32 c = q > 0 ? (u32)(w0 >> 32) + 4 : 0x80000000u;        // x+1
    if (a < best) {
      best = a;
      src = (u32)w2;
    }
    if (b < best) {
      best = b;
      src = (u32)w1;
    }
    if (c < best) {
      best = c;
      src = (u32)w0;
    }
    if (d < best) {
      best = d;
      src = (u32)last;
    }
    return d32(best) | src;
  }
};

}  // namespace xp
