// The four raster-scan recurrences of the blend path, as problems for the
// wavefront driver (wavefront.h). Each step() is the reference's per-pixel
// update restated on (own-row predecessor, three values of the previous row).
//
//   InpaintFwd / InpaintBwd : external/multiblend/src/image.cpp:550-812
//   SeamBwd / SeamFwd       : external/multiblend/src/multiblend.cpp:254-823
//
// Seam values: the reference's u64 is dist << 32 | image id with the top bit
// meaning "that image does not cover this pixel" (maskval,
// multiblend.cpp:74-76). They are carried here packed in 32 bits,
// flag << 31 | dist23 << 8 | id, by an order-preserving map (seam_pack /
// seam_unpack below): real distances stay below 2^22 for any canvas with
// W + H < kSeamMaxExtent, and the DT_MAX family 0x10000000 + k maps to
// 2^22 + k. Comparisons, +3/+4 and re-masking commute with the map, so the
// unpacked result is bit-identical to the reference's.
// Inpaint values are u64: dist << 32 | index of the source pixel whose colour
// the reference would have copied here (the reference copies the whole pixel
// in place from the winning neighbour; carrying the winner's source index and
// gathering once at the end is the same function of the input).
//
// Every step() is branch-free (selects only): one warp per band runs it as a
// serial dependent chain, so instruction count and divergence are the cost.
#pragma once

#include <cstdint>

#include "wavefront.h"

namespace xp {

typedef uint64_t u64;
typedef uint32_t u32;

// A value another block polls for (wavefront.h: the band below reads the last
// row of this band with ld.relaxed.gpu) is published with a store of the same
// strength: a weak store paired with a strong load is a data race under the
// PTX memory model, even though whole-word stores reach L2 intact today.
XP_HD void publish(u32* p, u32 v) {
#if defined(__CUDA_ARCH__)
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v)
               : "memory");
#else
  *p = v;
#endif
}
XP_HD void publish(u64* p, u64 v) {
#if defined(__CUDA_ARCH__)
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v)
               : "memory");
#else
  *p = v;
#endif
}

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

// ---------------------------------------------------------------------------
// packed seam values
// ---------------------------------------------------------------------------
constexpr u32 kTop32 = 0x80000000u;
constexpr u32 kInf32 = 0xffffffffu;
constexpr u32 kDtMax32 = 0xC0000000u;  // flag | (1 << 22) << 8 | label 0
constexpr u32 kD3 = 3u << 8, kD4 = 4u << 8;
constexpr int kSeamMaxExtent = 1000000;  // W + H bound for the 23-bit field

XP_HD u64 seam_unpack(u32 v) {
  const u32 d23 = (v >> 8) & 0x7fffffu;
  const u32 dist =
      d23 >= 0x400000u ? d23 - 0x400000u + 0x10000000u : d23;
  return ((u64)(v >> 31) << 63) | ((u64)dist << 32) | (u64)(v & 0xffu);
}

// maskval (multiblend.cpp:74-76): re-evaluate the cover flag of v's label
template <class C>
XP_HD u32 mask32(u32 v, C bits) {
  const u32 covered = (u32)(bits >> (v & (sizeof(C) * 8 - 1))) & 1u;
  return (v & 0x7fffffffu) | ((covered ^ 1u) << 31);
}
XP_HD u32 min32(u32 a, u32 b) { return a < b ? a : b; }
// min(a + b, c): one instruction (VIADDMNMX) on the device
XP_HD u32 add_min32(u32 a, u32 b, u32 c) {
#if defined(__CUDA_ARCH__)
  return __viaddmin_u32(a, b, c);
#else
  return min32(a + b, c);
#endif
}

// Two encodings of the per-pixel cover word.
//  kCoverBits  : bit i = image i opaque, top bit = run start (u32: <= 31
//                images, u64: <= 63).
//  kCoverDigest: u32, <= 24 images; what a step would otherwise recompute
//                for every pixel is stored ready-made:
//                [31:8] cover bits, reversed (image i at bit 31 - i, so that
//                `word << label` brings the label's bit to the top: maskval
//                is one funnel shift and one LOP3), [4:0] image an arbitrary
//                seam picks (lowest, or highest with `reverse`), [5] run
//                start, [6] exactly one image covers, [7] some image covers.
enum { kCoverBits = 0, kCoverDigest = 1 };
constexpr int kDigestMaxImages = 24;

template <class C, int F>
struct Cov {
  static XP_HD u32 mask(u32 v, C c) { return mask32(v, bits(c)); }
  static XP_HD u32 mask_if(u32 v, C c, bool on) { return on ? mask(v, c) : v; }
  static XP_HD C bits(C c) { return c & ~CoverTraits<C>::rs_bit; }
  static XP_HD bool rs(C c) { return (c & CoverTraits<C>::rs_bit) != 0; }
  static XP_HD bool seed(C c) { return popc_c(bits(c)) == 1; }
  static XP_HD bool any(C c) { return bits(c) != 0; }
  static XP_HD int low(C c) { return ctz_c((C)(c | CoverTraits<C>::rs_bit)); }
  static XP_HD int pick(C c, int reverse) {
    return reverse ? clz_top((C)(bits(c) | 1)) : low(c);
  }
};
template <>
struct Cov<u32, kCoverDigest> {
  // maskval: the label's (reversed) cover bit shifted to the top, inverted,
  // replaces v's flag. Labels of live values are < 24; bits 7..0 of the word
  // can only reach the top for labels >= 24.
  static XP_HD u32 mask(u32 v, u32 c) {
#if defined(__CUDA_ARCH__)
    // two instructions on the step's dependent chain: funnel shift with a
    // wrapped amount, then (v & ~m) | (~top & m) as one LOP3 (LUT 0x72)
    u32 top, r;
    asm("shf.l.wrap.b32 %0, %1, %2, %3;" : "=r"(top) : "r"(0u), "r"(c), "r"(v));
    asm("lop3.b32 %0, %1, %2, %3, 0x72;"
        : "=r"(r)
        : "r"(v), "r"(top), "r"(0x80000000u));
    return r;
#else
    const u32 top = c << (v & 31u);
    return (v & 0x7fffffffu) | (~top & kTop32);
#endif
  }
  // maskval when `on`, v otherwise: the LOP3's select mask carries the choice
  static XP_HD u32 mask_if(u32 v, u32 c, bool on) {
#if defined(__CUDA_ARCH__)
    u32 top, r;
    asm("shf.l.wrap.b32 %0, %1, %2, %3;" : "=r"(top) : "r"(0u), "r"(c), "r"(v));
    asm("lop3.b32 %0, %1, %2, %3, 0x72;"
        : "=r"(r)
        : "r"(v), "r"(top), "r"(on ? 0x80000000u : 0u));
    return r;
#else
    return on ? mask(v, c) : v;
#endif
  }
  static XP_HD bool rs(u32 c) { return (c & (1u << 5)) != 0; }
  static XP_HD bool seed(u32 c) { return (c & (1u << 6)) != 0; }
  static XP_HD bool any(u32 c) { return (c & (1u << 7)) != 0; }
  static XP_HD int low(u32 c) { return (int)(c & 31u); }
  static XP_HD int pick(u32 c, int) { return (int)(c & 31u); }
};
XP_HD u32 brev32(u32 v) {
#if defined(__CUDA_ARCH__)
  return __brev(v);
#else
  v = ((v >> 1) & 0x55555555u) | ((v & 0x55555555u) << 1);
  v = ((v >> 2) & 0x33333333u) | ((v & 0x33333333u) << 2);
  v = ((v >> 4) & 0x0f0f0f0fu) | ((v & 0x0f0f0f0fu) << 4);
  v = ((v >> 8) & 0x00ff00ffu) | ((v & 0x00ff00ffu) << 8);
  return (v >> 16) | (v << 16);
#endif
}
XP_HD u32 make_digest(u32 bits, bool run_start, int reverse) {
  const int cnt = popc_c(bits);
  const int pick = bits ? (reverse ? clz_top(bits) : ctz_c(bits)) : 0;
  return brev32(bits) | (u32)pick | (run_start ? 1u << 5 : 0u) |
         (cnt == 1 ? 1u << 6 : 0u) | (cnt != 0 ? 1u << 7 : 0u);
}

struct NoRow {};

// ---------------------------------------------------------------------------
// Seams, backward pass: rows bottom -> top, x right -> left.
// Processing coordinates: q = W-1-x, rho = H-1-y.
// ---------------------------------------------------------------------------
template <class C, int F = kCoverBits>
struct SeamBwd {
  typedef Cov<C, F> Cv;
  typedef u32 Val;
  typedef C In0;
  typedef wf::NoInput In1;
  struct Row {
    bool first_run;  // still inside the first run of the bottom row
  };
  int n_bands, W, H;
  const C* cover;  // H x W, bit i = image i opaque, top bit = run start
  u32* out;        // H x W packed backward values

  XP_HD wf::Band band(int i) const {
    return wf::Band{0, i * wf::kLanes, W, H, i - 1};
  }
  XP_HD Row row_init(const wf::Band&, int) const { return Row{true}; }
  XP_HD size_t at(int rho, int q) const {
    return (size_t)(H - 1 - rho) * W + (W - 1 - q);
  }
  XP_HD const C* in0_ptr(const wf::Band&, int rho, int q) const {
    return cover + at(rho, q);
  }
  XP_HD const void* in1_ptr(const wf::Band&, int, int) const { return 0; }
  XP_HD u32 load_val(const wf::Band&, int rho, int q) const {
    return out[at(rho, q)];
  }
  XP_HD const u32* val_ptr(const wf::Band&, int rho, int q) const {
    return out + at(rho, q);
  }
  XP_HD u32 sentinel() const { return kInf32; }
  XP_HD void store(const wf::Band&, int rho, int q, u32 v) const {
    publish(out + at(rho, q), v);
  }
  // chunks made of seeds only skip the recurrence; not on the bottom band,
  // whose first row carries the first_run state
  XP_HD bool allow_trivial(const wf::Band& B) const { return B.rho0 != 0; }
  XP_HD bool trivial(C cov, In1) const { return Cv::seed(cov); }
  XP_HD u32 trivial_value(const wf::Band&, const Row&, int, int, C cov,
                          In1) const {
    return (u32)Cv::low(cov);
  }
  // Tiles no image covers: maskval flags every candidate, so the step is a
  // plain 3-4 chamfer minimum. The flag of the carried state is recomputed
  // by every consumer (all of them go through maskval), so it may be set on
  // entry and the chain through `last` is the add-min alone.
  static constexpr bool kHasUncovered = true;
  XP_HD bool uncovered(C cov, In1) const { return !Cv::any(cov); }
  XP_HD u32 uncovered_up(u32 up) const { return up | kTop32; }
  XP_HD u32 uncovered_last(u32 last) const { return last | kTop32; }
  template <bool kEdgeY>
  XP_HD u32 step_uncovered(Row& row, int rho, C cov, In1, u32 last, u32 w0,
                           u32 w1, u32 w2) const {
    u32 pre = add_min32(w2, kD4, add_min32(w1, kD3, w0 + kD4));
    bool bottom_first = false;
    if (kEdgeY) {
      const bool was_first = row.first_run;
      if (Cv::rs(cov)) row.first_run = false;
      if (rho == 0) pre = kInf32, bottom_first = was_first;
    }
    u32 d = add_min32(last, kD3, pre);
    if (kEdgeY && bottom_first) d = kDtMax32;
    return d;
  }
  // last = (x+1, y); w0 = (x+1, y+1); w1 = (x, y+1); w2 = (x-1, y+1).
  // kEdgeX = false: the caller guarantees 0 < q < W-1;
  // kEdgeY = false: the caller guarantees rho > 0.
  template <bool kEdgeX, bool kEdgeY>
  XP_HD u32 step(const wf::Band&, Row& row, bool active, int rho, int q, C cov,
                 In1, u32 last, u32 w0, u32 w1, u32 w2) const {
    // w2 arrives by shuffle from the lane above and depends on that lane's
    // value two steps back: over the whole warp, `last` -> shuffle -> c2 ->
    // pre -> `last` is the loop-carried cycle, so everything that does not
    // need w2 is folded into `early` first. The chain through `last` itself
    // is maskval (2 instructions) and one add-min; the add commutes with
    // maskval (it cannot carry into the label or the flag).
    u32 c0 = Cv::mask(w0 + kD4, cov);
    const u32 c1 = Cv::mask(w1 + kD3, cov);
    u32 c2 = Cv::mask(w2 + kD4, cov);
    bool left = true;  // the (x+1, y) neighbour exists
    if (kEdgeX) {
      if (q <= 0) left = false, c0 = kInf32;
      if (q >= W - 1) c2 = kInf32;
    }
    u32 early = min32(c0, c1);
    bool bottom_first = false;
    if (kEdgeY) {
      const bool was_first = row.first_run;
      if (active && Cv::rs(cov)) row.first_run = false;
      // bottom row :333-346: the left neighbour alone, or DT_MAX
      if (rho == 0) early = kInf32, c2 = kInf32, bottom_first = was_first;
    }
    // a seed's value is below anything a neighbour can offer :327-331
    if (Cv::seed(cov)) early = (u32)Cv::low(cov);
    const u32 pre = min32(early, c2);
    u32 d = add_min32(Cv::mask(last, cov), kD3, pre);
    if (kEdgeX && !left) d = pre;
    if (kEdgeY && bottom_first && !Cv::seed(cov)) d = kDtMax32;
    return d;
  }
};

// ---------------------------------------------------------------------------
// Seams, forward pass: rows top -> bottom, x left -> right. q = x, rho = y.
// ---------------------------------------------------------------------------
template <class C, int F = kCoverBits>
struct SeamFwd {
  typedef Cov<C, F> Cv;
  typedef u32 Val;
  typedef u32 In0;  // packed backward value of this pixel
  typedef C In1;    // cover word
  typedef NoRow Row;
  int n_bands, W, H, reverse;
  const C* cover;
  const u32* bwd;
  u32* out;         // H x W packed forward values (also the hand-over rows)
  uint8_t* labels;  // H x W owner image

  XP_HD wf::Band band(int i) const {
    return wf::Band{0, i * wf::kLanes, W, H, i - 1};
  }
  XP_HD Row row_init(const wf::Band&, int) const { return Row{}; }
  XP_HD const u32* in0_ptr(const wf::Band&, int rho, int q) const {
    return bwd + (size_t)rho * W + q;
  }
  XP_HD const C* in1_ptr(const wf::Band&, int rho, int q) const {
    return cover + (size_t)rho * W + q;
  }
  XP_HD u32 load_val(const wf::Band&, int rho, int q) const {
    return out[(size_t)rho * W + q];
  }
  XP_HD const u32* val_ptr(const wf::Band&, int rho, int q) const {
    return out + ((size_t)rho * W + q);
  }
  XP_HD u32 sentinel() const { return kInf32; }
  XP_HD void store(const wf::Band&, int rho, int q, u32 v) const {
    const size_t p = (size_t)rho * W + q;
    publish(out + p, v);
    labels[p] = (uint8_t)(v & 0xffu);
  }
  XP_HD bool allow_trivial(const wf::Band&) const { return true; }
  XP_HD bool trivial(u32, C cov) const { return Cv::seed(cov); }
  XP_HD u32 trivial_value(const wf::Band&, const Row&, int, int, u32,
                          C cov) const {
    return (u32)Cv::low(cov);
  }
  // Tiles no image covers (see SeamBwd): the row above is always re-masked,
  // i.e. flagged; the left neighbour enters raw inside a run (:750), so its
  // flag is kept as it is.
  static constexpr bool kHasUncovered = true;
  XP_HD bool uncovered(u32, C cov) const { return !Cv::any(cov); }
  XP_HD u32 uncovered_up(u32 up) const { return up | kTop32; }
  XP_HD u32 uncovered_last(u32 last) const { return last; }
  template <bool kEdgeY>
  XP_HD u32 step_uncovered(Row&, int y, u32 bw, C cov, u32 last, u32 w0,
                           u32 w1, u32 w2) const {
    u32 pre = add_min32(w2, kD4, add_min32(w1, kD3, w0 + kD4));
    pre = min32(pre, bw);
    bool remask = Cv::rs(cov);
    if (kEdgeY) {
      if (y == 0) pre = bw, remask = true;
    }
    return add_min32(last | (remask ? kTop32 : 0u), kD3, pre);
  }
  // last = (x-1, y); w0 = (x-1, y-1); w1 = (x, y-1); w2 = (x+1, y-1).
  // kEdgeX = false: the caller guarantees 1 < x < W-1;
  // kEdgeY = false: the caller guarantees y > 0.
  template <bool kEdgeX, bool kEdgeY>
  XP_HD u32 step(const wf::Band&, Row&, bool, int y, int x, u32 bw, C cov,
                 u32 last, u32 w0, u32 w1, u32 w2) const {
    u32 c0 = Cv::mask(w0 + kD4, cov);
    u32 c1 = Cv::mask(w1 + kD3, cov);
    u32 c2 = Cv::mask(w2 + kD4, cov);
    // the left neighbour is re-masked only on the top row, at x == 1 and at
    // a run start (:629-630, :704-705); raw inside a run (:750)
    const bool remask =
        Cv::rs(cov) || (kEdgeY && y == 0) || (kEdgeX && x == 1);
    bool left = true;
    if (kEdgeX) {
      if (x <= 0) left = false, c0 = kInf32;
      if (x >= W - 1) c2 = kInf32;
    }
    if (kEdgeY) {
      if (y == 0) c0 = kInf32, c1 = kInf32, c2 = kInf32;
    }
    // ReadSeamDT value enters unmasked. Everything that does not need w2 is
    // folded into `early` (see SeamBwd::step):
    //  * arbitrary seam :732-742 -- the minimum is flagged iff every
    //    candidate is; flag|pick (distance 0) is below every other flagged
    //    value (those are at distance >= 3) and above every unflagged one, so
    //    taking the minimum with it makes the result come out right whatever
    //    w2 and the left neighbour are;
    //  * a seed's value :607-616 is below anything a neighbour can offer.
    u32 early = min32(min32(bw, c0), c1);
    if (Cv::any(cov))
      early = min32(early, kTop32 | (u32)Cv::pick(cov, reverse));
    if (Cv::seed(cov)) early = (u32)Cv::low(cov);
    const u32 pre = min32(early, c2);
    u32 best = add_min32(Cv::mask_if(last, cov, remask), kD3, pre);
    if (kEdgeX && !left) best = pre;
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
struct PixelOf {
  typedef u32 type;
};
template <>
struct PixelOf<16> {
  typedef u64 type;
};
XP_HD bool pixel_opaque(u32 px) { return (px >> 24) == 0xffu; }
XP_HD bool pixel_opaque(u64 px) { return (px >> 48) == 0xffffull; }

template <int BPP>
XP_HD bool px_opaque(const void* data, size_t idx) {
  if (BPP == 8)
    return (static_cast<const u32*>(data)[idx] >> 24) == 0xffu;
  return (static_cast<const u64*>(data)[idx] >> 48) == 0xffffull;
}

XP_HD u32 hi32(u64 v) { return (u32)(v >> 32); }
XP_HD u32 lo32(u64 v) { return (u32)v; }

// forward pass: q = x, rho = y (image.cpp:550-694)
template <int BPP>
struct InpaintFwd {
  typedef u64 Val;
  typedef typename PixelOf<BPP>::type In0;  // the pixel (alpha on top)
  typedef wf::NoInput In1;
  struct Row {
    u32 base;  // index of (x = 0, this row) in the untrimmed image
  };
  int n_bands;
  const InpaintJob* jobs;
  const InpaintBandRef* bands;

  XP_HD wf::Band band(int i) const {
    const InpaintBandRef r = bands[i];
    const InpaintJob& j = jobs[r.job];
    return wf::Band{r.job, r.local * wf::kLanes, j.w, j.h, r.dep};
  }
  XP_HD Row row_init(const wf::Band& B, int y) const {
    const InpaintJob& j = jobs[B.job];
    return Row{(u32)((size_t)(j.top + y) * j.tiff_w + j.left)};
  }
  XP_HD const In0* in0_ptr(const wf::Band& B, int y, int x) const {
    const InpaintJob& j = jobs[B.job];
    return static_cast<const In0*>(j.data) +
           ((size_t)(j.top + y) * j.tiff_w + j.left + x);
  }
  XP_HD const void* in1_ptr(const wf::Band&, int, int) const { return 0; }
  XP_HD u64 load_val(const wf::Band& B, int y, int x) const {
    return jobs[B.job].fwd[(size_t)y * B.W + x];
  }
  XP_HD const u64* val_ptr(const wf::Band& B, int y, int x) const {
    return jobs[B.job].fwd + ((size_t)y * B.W + x);
  }
  XP_HD u64 sentinel() const { return ~0ull; }
  XP_HD void store(const wf::Band& B, int y, int x, u64 v) const {
    publish(jobs[B.job].fwd + ((size_t)y * B.W + x), v);
  }
  // no uncovered-tile variant (wavefront.h); never called
  static constexpr bool kHasUncovered = false;
  XP_HD bool uncovered(In0, In1) const { return false; }
  XP_HD u64 uncovered_up(u64 v) const { return v; }
  XP_HD u64 uncovered_last(u64 v) const { return v; }
  template <bool kEdgeY>
  XP_HD u64 step_uncovered(Row&, int, In0, In1, u64 last, u64, u64,
                           u64) const {
    return last;
  }
  XP_HD bool allow_trivial(const wf::Band&) const { return true; }
  XP_HD bool trivial(In0 px, In1) const { return pixel_opaque(px); }
  // `row` is the calling lane's; the wavefront asks for other rows of the
  // band too, which sit at a fixed stride from it
  XP_HD u64 trivial_value(const wf::Band& B, const Row& row, int dy, int x,
                          In0, In1) const {
    return (u64)(row.base + (u32)(dy * jobs[B.job].tiff_w) + (u32)x);
  }
  // last = (x-1, y); w0 = (x-1, y-1); w1 = (x, y-1); w2 = (x+1, y-1).
  // kEdgeX = false: the caller guarantees 0 < x < W-1;
  // kEdgeY = false: the caller guarantees y > 0.
  template <bool kEdgeX, bool kEdgeY>
  XP_HD u64 step(const wf::Band& B, Row& row, bool, int y, int x, In0 px, In1,
                 u64 last, u64 w0, u64 w1, u64 w2) const {
    const u32 me = row.base + (u32)x;
    // own row :573-583 / :609-641; the very first pixel seeds 0x80000000
    u32 d = hi32(last) + 3, s = lo32(last);
    u32 a = hi32(w0) + 4, b = hi32(w1) + 3, c = hi32(w2) + 4;
    if (kEdgeX) {
      if (x == 0) d = (kEdgeY && y == 0) ? 0x80000000u : kInf32, s = me;
      if (x == 0) a = kInf32;
      if (x >= B.W - 1) c = kInf32;
    }
    if (kEdgeY) {
      if (y == 0) a = kInf32, b = kInf32, c = kInf32;
    }
    if (a < d) d = a, s = lo32(w0);
    if (b < d) d = b, s = lo32(w1);
    // in column 0 a tie between up and up-right goes to up-right (:585-603)
    if (c < d || (kEdgeX && x == 0 && c == d)) d = c, s = lo32(w2);
    if (pixel_opaque(px)) d = 0, s = me;  // :659-662
    return d32(d) | s;
  }
};

// backward pass: q = w-1-x, rho = h-1-y (image.cpp:711-812)
struct InpaintBwd {
  typedef u64 Val;
  typedef u64 In0;  // forward value of this pixel (dist 0 <=> opaque)
  typedef wf::NoInput In1;
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
  XP_HD const u64* in0_ptr(const wf::Band& B, int rho, int q) const {
    return jobs[B.job].fwd + at(B, rho, q);
  }
  XP_HD const void* in1_ptr(const wf::Band&, int, int) const { return 0; }
  XP_HD u64 load_val(const wf::Band& B, int rho, int q) const {
    return jobs[B.job].bwd[at(B, rho, q)];
  }
  XP_HD const u64* val_ptr(const wf::Band& B, int rho, int q) const {
    return jobs[B.job].bwd + at(B, rho, q);
  }
  XP_HD u64 sentinel() const { return ~0ull; }
  XP_HD void store(const wf::Band& B, int rho, int q, u64 v) const {
    publish(jobs[B.job].bwd + at(B, rho, q), v);
  }
  // no uncovered-tile variant (wavefront.h); never called
  static constexpr bool kHasUncovered = false;
  XP_HD bool uncovered(In0, In1) const { return false; }
  XP_HD u64 uncovered_up(u64 v) const { return v; }
  XP_HD u64 uncovered_last(u64 v) const { return v; }
  template <bool kEdgeY>
  XP_HD u64 step_uncovered(Row&, int, In0, In1, u64 last, u64, u64,
                           u64) const {
    return last;
  }
  XP_HD bool allow_trivial(const wf::Band&) const { return true; }
  XP_HD bool trivial(u64 fwd, In1) const { return (fwd >> 32) == 0; }
  XP_HD u64 trivial_value(const wf::Band&, const Row&, int, int, u64 fwd,
                          In1) const {
    return fwd;
  }
  // last = (x+1, y); w0 = (x+1, y+1); w1 = (x, y+1); w2 = (x-1, y+1).
  // kEdgeX = false: the caller guarantees 0 < q < W-1;
  // kEdgeY = false: the caller guarantees rho > 0.
  template <bool kEdgeX, bool kEdgeY>
  XP_HD u64 step(const wf::Band& B, Row&, bool, int rho, int q, u64 fwd, In1,
                 u64 last, u64 w0, u64 w1, u64 w2) const {
    u32 best = hi32(fwd), src = lo32(fwd);
    const bool opaque = best == 0;  // cur = 0, d = 3
    // at the row start the reference seeds the running distance with
    // 0x80000000 (nothing on the last line)
    u32 a = hi32(w2) + 4;  // x-1
    u32 b = hi32(w1) + 3;
    u32 c = hi32(w0) + 4;  // x+1
    u32 d = hi32(last) + 3;
    const bool bottom = kEdgeY && rho == 0;  // last line :711-747: right
                                             // neighbour only
    if (kEdgeX) {
      if (q >= B.W - 1) a = 0x80000000u;
      if (q <= 0) c = 0x80000000u, d = bottom ? kInf32 : 0x80000000u;
    }
    if (kEdgeY) {
      if (bottom) a = kInf32, b = kInf32, c = kInf32;
    }
    if (a < best) best = a, src = lo32(w2);
    if (b < best) best = b, src = lo32(w1);
    if (c < best) best = c, src = lo32(w0);
    if (d < best) best = d, src = lo32(last);
    const u64 v = d32(best) | src;
    return opaque ? fwd : v;
  }
};

}  // namespace xp
