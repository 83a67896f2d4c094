// Streaming form of the per-image pyramid work (SURVEY.md 8(d) byte model):
//
//   * level 0 of an image pyramid is never materialised as floats: the
//     inpainted pixels stay in the interleaved source buffer (the inpaint
//     result is gathered into it in place, seams.cu) and are converted on the
//     fly;
//   * the Laplacian is never materialised either: expand(lo) - hi is formed in
//     registers and multiplied straight into the canvas pyramid, and only
//     where the image's mask is non-zero.
//
// Same per-sample arithmetic, in the same order, as the staged kernels in
// pyramid.cu (pyramid_math.h), which remain as the keep_debug path.
// Replaces external/multiblend/src/pyramid.cpp Copy :233-408 + Shrink
// :584-823 (first level), LaplaceCollapse :828-1038 + functions.cpp
// CompositeLine :319-411 (all levels below the top).
#include "blend_impl.h"
#include "common.h"
#include "pyramid_dev.h"
#include "shrink_tiled.h"
#include "pyramid_math.h"

namespace xp {

// ---------------------------------------------------------------------------
// virtual level 0: the trimmed, inpainted source pixels as floats
// ---------------------------------------------------------------------------
template <int BPP>
struct Level0 {
  typedef typename PixelOf<BPP>::type Px;
  const Px* data;
  int tiff_w, left, top, w, h;
  float mul;
  // raw pixel at level coordinates; columns past w replicate the last pixel
  // (pyramid.cpp:268-330)
  __device__ __forceinline__ Px px(int row, int col) const {
    const int c = col < w ? col : w - 1;
    return __ldg(data + ((size_t)(top + row) * tiff_w + left + c));
  }
  static __device__ __forceinline__ float chan(Px v, int c) {
    if (BPP == 8) return (float)((u32)(v >> (8 * c)) & 0xffu);
    return (float)((u32)((u64)v >> (16 * c)) & 0xffffu);
  }
  __device__ __forceinline__ float at(int c, int row, int col) const {
    if (row >= h) return 0.0f;  // rows the staged path leaves zeroed
    return chan(px(row, col), c) * mul;
  }
};

template <int BPP>
static Level0<BPP> make_level0(const ImgDev& im, float mul) {
  return Level0<BPP>{static_cast<const typename PixelOf<BPP>::type*>(im.data),
                     im.tiff_w, im.left, im.top, im.w, im.h, mul};
}

// ---------------------------------------------------------------------------
// shrink level 0 -> 1 straight from the source pixels, 3 channels together
// ---------------------------------------------------------------------------
template <int BPP>
struct PixelSrc : Level0<BPP> {
  typedef typename Level0<BPP>::Px Raw;
  static constexpr int kChannels = 3;
  // 8-bit pixels: B, G, R in the low three bytes of a 32-bit word
  static constexpr bool kPackedBytes = BPP == 8;
  __device__ __forceinline__ Raw raw(int, int row, int col) const {
    return row < this->h ? this->px(row, col) : (Raw)0;  // zeroed rows
  }
  __device__ __forceinline__ float value(Raw v, int c) const {
    return this->chan(v, c) * this->mul;
  }
};

// level 0 -> 1 of n images straight from their source pixels, in launches of
// up to kShrinkBatch images
template <int BPP>
static void shrink0_batch_t(const ImgDev* im, float mul, const PyrGeom* g,
                            float* const* pyr, int n, void* d_tmaps,
                            Mailbox* mbox, cudaStream_t s) {
  for (int i0 = 0; i0 < n; i0 += kShrinkBatch) {
    const int m = std::min(kShrinkBatch, n - i0);
    ShrinkBatch<PixelSrc<BPP>> batch;
    batch.tma = 0;
    batch.tmaps = static_cast<const TensorMapBytes*>(d_tmaps) + i0;
    TensorMapBytes host_maps[kShrinkBatch];
    memset(host_maps, 0, sizeof(host_maps));
    int gx = 0, gy = 0;
    for (int k = 0; k < m; ++k) {
      const PyrGeom& G = g[i0 + k];
      const LevelGeom& H = G.lv[0];
      const LevelGeom& L = G.lv[1];
      PixelSrc<BPP> src;
      static_cast<Level0<BPP>&>(src) = make_level0<BPP>(im[i0 + k], mul);
      batch.job[k] = ShrinkJob<PixelSrc<BPP>>{
          src, H.h, H.pitch, H.xs, H.ys, pyr[i0 + k] + L.offset, L.h, L.pitch,
          G.channel_floats};
      gx = std::max(gx, div_up(L.pitch, kShrinkTileW));
      gy = std::max(gy, div_up(L.h, kShrinkTileH));
      // 8-bit pixels: interior tiles are staged by TMA (one bulk tensor copy
      // of 2*64+4 x 2*8+3 pixels per block) where the driver gives a map
      if (BPP == 8 && mul == 1.0f && d_tmaps && mbox &&
          getenv("XP_NO_TMA") == nullptr) {
        const ImgDev& d = im[i0 + k];
        if (make_tensor_map_2d(host_maps[k].b, 4, d.data,
                               (unsigned long long)d.tiff_w,
                               (unsigned long long)d.tiff_h,
                               (unsigned long long)d.tiff_w * 4,
                               2 * kShrinkTileW + 8, 2 * kShrinkTileH + 3))
          batch.tma |= 1u << k;
      }
    }
    if (batch.tma)
      mbox->to_device(const_cast<TensorMapBytes*>(batch.tmaps), host_maps,
                      sizeof(TensorMapBytes) * m, s);
    for (int k = m; k < kShrinkBatch; ++k) batch.job[k] = batch.job[0];
    shrink_batch_kernel<<<dim3(gx, gy, m), 256, 0, s>>>(batch, 1);
    XP_LAUNCHED();
  }
}

void launch_shrink0_batch(const ImgDev* im, int bpp, float mul,
                          const PyrGeom* g, float* const* pyr, int n,
                          void* d_tmaps, Mailbox* mbox, cudaStream_t s) {
  if (bpp == 8)
    shrink0_batch_t<8>(im, mul, g, pyr, n, d_tmaps, mbox, s);
  else
    shrink0_batch_t<16>(im, mul, g, pyr, n, d_tmaps, mbox, s);
}

void launch_shrink0(const ImgDev& im, int bpp, float mul, const PyrGeom& g,
                    float* pyr, cudaStream_t s) {
  const LevelGeom& H = g.lv[0];
  const LevelGeom& L = g.lv[1];
  dim3 gr(div_up(L.pitch, kShrinkTileW), div_up(L.h, kShrinkTileH), 1);
  if (bpp == 8) {
    PixelSrc<8> src;
    static_cast<Level0<8>&>(src) = make_level0<8>(im, mul);
    shrink_tiled_kernel<<<gr, 256, 0, s>>>(src, H.h, H.pitch, H.xs, H.ys,
                                           pyr + L.offset, L.h, L.pitch,
                                           g.channel_floats);
  } else {
    PixelSrc<16> src;
    static_cast<Level0<16>&>(src) = make_level0<16>(im, mul);
    shrink_tiled_kernel<<<gr, 256, 0, s>>>(src, H.h, H.pitch, H.xs, H.ys,
                                           pyr + L.offset, L.h, L.pitch,
                                           g.channel_floats);
  }
  XP_LAUNCHED();
}

// ---------------------------------------------------------------------------
// fused Laplace + composite for one level below the top:
//   out(X, Y) += (expand(lo) - hi)(clamp(X - x_off), clamp(Y - y_off)) * mask
// over the mask rectangle, 3 channels together, four consecutive image
// samples per thread. `Hi` is either a float level or the virtual level 0.
// ---------------------------------------------------------------------------
struct FloatLevel {
  const float* p;  // channel 0 of the level
  int pitch;
  size_t cs;  // channel stride
  __device__ __forceinline__ float at(int c, int row, int col) const {
    return p[c * cs + (size_t)row * pitch + col];
  }
  // samples (row, 4g .. 4g+3) of all three channels
  __device__ __forceinline__ void load4(int row, int g, float h[3][4]) const {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float4 v = *reinterpret_cast<const float4*>(
          p + c * cs + (size_t)row * pitch + 4 * g);
      h[c][0] = v.x, h[c][1] = v.y, h[c][2] = v.z, h[c][3] = v.w;
    }
  }
};

template <int BPP>
struct PixelLevel : Level0<BPP> {
  __device__ __forceinline__ void load4(int row, int g, float h[3][4]) const {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const typename Level0<BPP>::Px v = this->px(row, 4 * g + k);
#pragma unroll
      for (int c = 0; c < 3; ++c) h[c][k] = this->chan(v, c) * this->mul;
    }
  }
};

__device__ __forceinline__ int floor_div4(int a) {
  return a >= 0 ? a >> 2 : -((-a + 3) >> 2);
}

// kFirstTouch (level 0): the label map gives every canvas pixel to exactly
// one image with weight 1, so the level is written once instead of zeroed
// and accumulated into (0 + x, as the reference computes it).
//
// A thread owns four samples of two consecutive canvas rows whose image rows
// t = iy + ys (even) and t + 1 share their lo rows, so every horizontally
// expanded lo row is formed once for both (as in laplace_kernel); rows that
// are clamped to the image, lie outside the mask rectangle or touch the
// image's right edge take the one-row path.
template <class Hi, bool kFirstTouch>
struct LapComp {
  Hi hi;
  int in_w, in_h, xs, ys;
  const float* __restrict__ lo;
  int lo_pitch;
  size_t in_cs;
  float* __restrict__ out;
  int out_pitch;
  size_t out_cs;
  int x_off, y_off;
  MaskLevel m;
  const float* __restrict__ mask;
  const uint8_t* __restrict__ labels;
  int image;

  // weights of samples X0 .. X0+3 of canvas row Y; false if all are zero
  __device__ __forceinline__ bool weights(int X0, int Y, float mv[4]) const {
    bool any = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int X = X0 + k;
      float v = 0.0f;
      if (X >= m.x0 && X < m.x0 + m.w) {
        if (labels)
          v = labels[(size_t)Y * m.cw + X] == image ? 1.0f : 0.0f;
        else
          v = mask[(size_t)(Y - m.y0) * m.w + (X - m.x0)];
      }
      mv[k] = v;
      any = any || v != 0.0f;
    }
    return any;
  }
  __device__ __forceinline__ void put(int c, int X0, int Y, const float mv[4],
                                      const float lap[4]) const {
    float* row = out + c * out_cs + (size_t)Y * out_pitch + X0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (mv[k] != 0.0f)
        row[k] = (kFirstTouch ? 0.0f : row[k]) + lap[k] * mv[k];
  }
  // one canvas row (any position: clamped addressing, functions.cpp:319-411)
  __device__ __forceinline__ void row(int g, int X0, int Y) const {
    float mv[4];
    if (!weights(X0, Y, mv)) return;  // functions.cpp:347-351 (skip)
    const int iy = pm::clampi(Y - y_off, 0, in_h - 1);
    if (g >= 0 && 4 * g + 3 <= in_w - 1) {
      float h[3][4];
      hi.load4(iy, g, h);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float e[4], lap[4];
        expand4(lo + c * in_cs, lo_pitch, xs, ys, g, iy, e);
#pragma unroll
        for (int k = 0; k < 4; ++k) lap[k] = e[k] - h[c][k];
        put(c, X0, Y, mv, lap);
      }
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float lap[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int ix = pm::clampi(4 * g + k, 0, in_w - 1);
          lap[k] = mv[k] != 0.0f
                       ? pm::expand_sample(lo + c * in_cs, lo_pitch, xs, ys,
                                           ix, iy) -
                             hi.at(c, iy, ix)
                       : 0.0f;
        }
        put(c, X0, Y, mv, lap);
      }
    }
  }
  // canvas rows Ya, Ya + 1 with image rows iya (iya + ys even), iya + 1, both
  // inside the image and the rectangle, group g inside the image
  __device__ __forceinline__ void pair(int g, int X0, int Ya, int iya) const {
    float ma[4], mb[4];
    const bool any_a = weights(X0, Ya, ma), any_b = weights(X0, Ya + 1, mb);
    if (!any_a && !any_b) return;
    float ha[3][4], hb[3][4];
    hi.load4(iya, g, ha);
    hi.load4(iya + 1, g, hb);
    const int j = ((iya + ys) >> 1) + 1;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float* r1 = lo + c * in_cs + (size_t)j * lo_pitch;
      float e0[4], e1[4], e2[4], la[4], lb[4];
      expand_row4(r1 - lo_pitch, g, xs, e0);
      expand_row4(r1, g, xs, e1);
      expand_row4(r1 + lo_pitch, g, xs, e2);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        la[k] = pm::expand_v(1, e0[k], e1[k], e2[k]) - ha[c][k];
        lb[k] = pm::expand_v(0, e0[k], e1[k], e2[k]) - hb[c][k];
      }
      if (any_a) put(c, X0, Ya, ma, la);
      if (any_b) put(c, X0, Ya + 1, mb, lb);
    }
  }
};

// kPairs: two rows per thread (level 0, tens of thousands of blocks); the
// upper levels are a few hundred blocks each and keep one row per thread --
// halving their grids costs more in parallelism than the shared rows save.
template <class Hi, bool kFirstTouch, bool kPairs>
__global__ void __launch_bounds__(256)
    lapcomp_kernel(const LapComp<Hi, kFirstTouch> k, int y_first) {
  const int g =
      floor_div4(k.m.x0 - k.x_off) + blockIdx.x * blockDim.x + threadIdx.x;
  const int y_end = k.m.y0 + k.m.h;
  if (!kPairs) {
    const int Y = k.m.y0 + blockIdx.y * blockDim.y + threadIdx.y;
    const int X0 = 4 * g + k.x_off;
    if (Y < y_end && X0 < k.m.x0 + k.m.w) k.row(g, X0, Y);
    return;
  }
  // row pairs start at y_first (<= m.y0), chosen so that image row
  // y_first - y_off + ys is even
  const int Ya = y_first + 2 * (blockIdx.y * blockDim.y + threadIdx.y);
  if (Ya >= y_end) return;
  const int X0 = 4 * g + k.x_off;
  if (X0 >= k.m.x0 + k.m.w) return;
  const int iya = Ya - k.y_off;
  if (Ya >= k.m.y0 && Ya + 1 < y_end && iya >= 0 && iya + 1 <= k.in_h - 1 &&
      g >= 0 && 4 * g + 3 <= k.in_w - 1 && 2 * g + 4 < k.lo_pitch) {
    k.pair(g, X0, Ya, iya);
  } else {
    if (Ya >= k.m.y0) k.row(g, X0, Ya);
    if (Ya + 1 < y_end) k.row(g, X0, Ya + 1);
  }
}

void launch_lapcomp(const ImgDev& im, int bpp, float mul, const PyrGeom& in_g,
                    const float* in_pyr, const PyrGeom& out_g, float* out_pyr,
                    int level, const uint8_t* labels, int image,
                    const MaskLevel& m, const float* mask, cudaStream_t s) {
  if (m.w <= 0 || m.h <= 0) return;
  const LevelGeom& I = in_g.lv[level];
  const LevelGeom& Lo = in_g.lv[level + 1];
  const LevelGeom& O = out_g.lv[level];
  const int x_off = (I.x - O.x) >> level;
  const int y_off = (I.y - O.y) >> level;
  const int g0 = (m.x0 - x_off) >= 0 ? (m.x0 - x_off) >> 2
                                     : -((-(m.x0 - x_off) + 3) >> 2);
  const int last = m.x0 + m.w - 1 - x_off;
  const int g1 = last >= 0 ? last >> 2 : -((-last + 3) >> 2);
  // row pairs: the first one starts on an image row of even t = iy + ys
  const int y_first = m.y0 - ((m.y0 - y_off + I.ys) & 1);
  const int pairs = (m.y0 + m.h - y_first + 1) / 2;
  dim3 b(32, 8);
  dim3 gr(div_up(g1 - g0 + 1, 32), div_up(level == 0 ? pairs : m.h, 8));
  const float* lo = in_pyr + Lo.offset;
  float* out = out_pyr + O.offset;
  const float* mk = level == 0 ? nullptr : mask + m.offset;
  const uint8_t* lb = level == 0 ? labels : nullptr;
#define XP_LAPCOMP(HI, FIRST)                                                 \
  {                                                                           \
    const LapComp<decltype(HI), FIRST> k{                                     \
        HI,       I.w,     I.h,    I.xs, I.ys, lo, Lo.pitch,                  \
        in_g.channel_floats, out,  O.pitch, out_g.channel_floats, x_off,      \
        y_off,    m,       mk,     lb,   image};                              \
    lapcomp_kernel<decltype(HI), FIRST, FIRST><<<gr, b, 0, s>>>(k, y_first);  \
  }
  if (level > 0) {
    FloatLevel hi{in_pyr + I.offset, I.pitch, in_g.channel_floats};
    XP_LAPCOMP(hi, false);
  } else if (bpp == 8) {
    PixelLevel<8> hi;
    static_cast<Level0<8>&>(hi) = make_level0<8>(im, mul);
    XP_LAPCOMP(hi, true);
  } else {
    PixelLevel<16> hi;
    static_cast<Level0<16>&>(hi) = make_level0<16>(im, mul);
    XP_LAPCOMP(hi, true);
  }
#undef XP_LAPCOMP
  XP_LAUNCHED();
}

}  // namespace xp
