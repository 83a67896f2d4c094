// Pyramid / blend kernels: level-0 copy, 5-tap reduce, expand-subtract
// (Laplace and Collapse), mask pyramid, masked composite, offset correction,
// dither/clamp/pack.
//
// Replaces external/multiblend/src/pyramid.cpp (Copy :233-408, Shrink :584-823,
// LaplaceCollapse :828-1038, Add :1084-1110, Out :1563-1794),
// src/functions.cpp (ShrinkMasks :172-314, CompositeLine :319-411) and the
// blend loop / offset correction of src/multiblend.cpp:1057-1131, 1232-1267,
// plus xpano/algorithm/blenders.cc:64-77,99-112,151-154 (pack / unpack).
// All arithmetic lives in pyramid_math.h; the whole library is built with
// -fmad=false so results are bit-comparable with the SSE reference.
#include "blend_impl.h"
#include "common.h"
#include "pyramid_dev.h"
#include "shrink_tiled.h"
#include "pyramid_math.h"

namespace xp {

PyrGeom make_pyramid(int width, int height, int levels, int x, int y) {
  PyrGeom g;
  XP_REQUIRE(levels <= kMaxLevels, "too many pyramid levels (%d)", levels);
  int b = 0, req = 2;
  size_t off = 0;
  for (int n = 0; n < levels; ++n) {
    LevelGeom& L = g.lv[n];
    L.xs = ((x - b) & (req - 1)) != 0;
    L.ys = ((y - b) & (req - 1)) != 0;
    L.w = width;
    L.h = height;
    L.x = x;
    L.y = y;
    L.pitch = (width + L.xs + 7) & ~7;
    L.rows = (height + L.ys + 3) & ~3;
    L.offset = off;
    off += (size_t)L.pitch * L.rows;
    x -= L.xs << n;
    y -= L.ys << n;
    x -= req;
    y -= req;
    b -= req;
    req <<= 1;
    width = (width + L.xs + 6) >> 1;
    height = (height + L.ys + 6) >> 1;
  }
  g.n = levels;
  g.channel_floats = off;
  return g;
}

static dim3 grid3(int w, int h, int z, dim3 b) {
  return dim3(div_up(w, (int)b.x), div_up(h, (int)b.y), z);
}

// ---------------------------------------------------------------------------
// level 0 copy (+ inpaint gather + de-interleave)
// ---------------------------------------------------------------------------
template <int BPP>
__global__ void copy_level0_kernel(ImgDev im, const u64* __restrict__ src_index,
                                   float mul, int pitch, size_t chan_stride,
                                   float* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= pitch || y >= im.h) return;
  const int xx = x < im.w ? x : im.w - 1;  // pad: replicate the last pixel
  size_t idx;
  if (src_index)
    idx = (u32)src_index[(size_t)y * im.w + xx];
  else
    idx = (size_t)(im.top + y) * im.tiff_w + im.left + xx;
  float c0, c1, c2;
  if (BPP == 8) {
    const u32 v = static_cast<const u32*>(im.data)[idx];
    c0 = (float)(v & 0xff);
    c1 = (float)((v >> 8) & 0xff);
    c2 = (float)((v >> 16) & 0xff);
  } else {
    const u64 v = static_cast<const u64*>(im.data)[idx];
    c0 = (float)(u32)(v & 0xffff);
    c1 = (float)(u32)((v >> 16) & 0xffff);
    c2 = (float)(u32)((v >> 32) & 0xffff);
  }
  if (mul != 1.0f) {
    c0 *= mul;
    c1 *= mul;
    c2 *= mul;
  }
  const size_t o = (size_t)y * pitch + x;
  out[o] = c0;
  out[o + chan_stride] = c1;
  out[o + 2 * chan_stride] = c2;
}

void launch_copy_level0(const ImgDev& im, int bpp, const u64* src_index,
                        float mul, const PyrGeom& g, float* pyr,
                        cudaStream_t s) {
  const LevelGeom& L = g.lv[0];
  dim3 b(32, 8);
  dim3 gr = grid3(L.pitch, L.h, 1, b);
  if (bpp == 8)
    copy_level0_kernel<8><<<gr, b, 0, s>>>(im, src_index, mul, L.pitch,
                                           g.channel_floats, pyr + L.offset);
  else
    copy_level0_kernel<16><<<gr, b, 0, s>>>(im, src_index, mul, L.pitch,
                                            g.channel_floats, pyr + L.offset);
  XP_LAUNCHED();
}

// ---------------------------------------------------------------------------
// shrink: four lo samples per thread. Away from the left/right borders the
// eleven hi columns they need are read as aligned float4 (three, or four when
// the level is x-shifted) from each of the form's rows; groups that touch a
// border clamp per sample (pm::shrink_sample).
// ---------------------------------------------------------------------------
__device__ __forceinline__ float4 ldg4(const float* p) {
  return __ldg(reinterpret_cast<const float4*>(p));
}

// level -> level + 1 of n pyramids (geometry g[i], base pointer pyr[i]), in
// launches of up to kShrinkBatch pyramids
void launch_shrink_batch(const PyrGeom* g, float* const* pyr, int n, int level,
                         int channels, cudaStream_t s) {
  for (int i0 = 0; i0 < n; i0 += kShrinkBatch) {
    const int m = std::min(kShrinkBatch, n - i0);
    ShrinkBatch<PlaneSrc> batch;
    batch.tma = 0;
    batch.tmaps = nullptr;
    int gx = 0, gy = 0;
    for (int k = 0; k < m; ++k) {
      const PyrGeom& G = g[i0 + k];
      const LevelGeom& H = G.lv[level];
      const LevelGeom& L = G.lv[level + 1];
      batch.job[k] = ShrinkJob<PlaneSrc>{
          PlaneSrc{pyr[i0 + k] + H.offset, H.pitch, G.channel_floats},
          H.h, H.pitch, H.xs, H.ys, pyr[i0 + k] + L.offset, L.h, L.pitch,
          G.channel_floats};
      gx = std::max(gx, div_up(L.pitch, kShrinkTileW));
      gy = std::max(gy, div_up(L.h, kShrinkTileH));
    }
    for (int k = m; k < kShrinkBatch; ++k) batch.job[k] = batch.job[0];
    shrink_batch_kernel<<<dim3(gx, gy, m * channels), 256, 0, s>>>(batch,
                                                                  channels);
    XP_LAUNCHED();
  }
}

void launch_shrink(const PyrGeom& g, float* pyr, int level, int channels,
                   cudaStream_t s) {
  const LevelGeom& H = g.lv[level];
  const LevelGeom& L = g.lv[level + 1];
  PlaneSrc src{pyr + H.offset, H.pitch, g.channel_floats};
  dim3 gr(div_up(L.pitch, kShrinkTileW), div_up(L.h, kShrinkTileH), channels);
  shrink_tiled_kernel<<<gr, 256, 0, s>>>(src, H.h, H.pitch, H.xs, H.ys,
                                         pyr + L.offset, L.h, L.pitch,
                                         g.channel_floats);
  XP_LAUNCHED();
}

// ---------------------------------------------------------------------------
// laplace / collapse: hi <- expand(lo) - hi, four hi samples per thread
// ---------------------------------------------------------------------------
// kSum: also accumulate, in double, the new hi values over the pixels of the
// xor mask (offset correction, multiblend.cpp:1232-1258) -- used on the last
// collapse step so level 0 is not read again for the sum. One partial per
// block, combined in a fixed order by xor_sum_final_kernel: deterministic.
// A thread owns four samples of two consecutive hi rows that share their lo
// rows: t = y + ys even (lo rows p, p+1, p+2) and t + 1 (lo rows p+1, p+2),
// so every horizontally expanded lo row is formed once for both.
template <bool kSum>
__global__ void __launch_bounds__(256)
    laplace_kernel(float* __restrict__ hi, int hh, int hp, int xs, int ys,
                   const float* __restrict__ lo, int lo_pitch,
                   size_t chan_stride, const uint8_t* __restrict__ xor_mask,
                   int W, double* __restrict__ partials) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int p = blockIdx.y * blockDim.y + threadIdx.y;  // row pair
  const int ya = 2 * p - ys, yb = ya + 1;               // ya + ys is even
  double acc = 0.0;
  if (4 * i < hp && ya < hh) {
    const size_t co = (size_t)blockIdx.z * chan_stride;
    const bool has_a = ya >= 0, has_b = yb < hh;
    float ea[4], eb[4];
    if (2 * i + 4 < lo_pitch) {
      const float* r0 = lo + co + (size_t)p * lo_pitch;
      float e0[4], e1[4], e2[4];
      expand_row4(r0, i, xs, e0);
      expand_row4(r0 + lo_pitch, i, xs, e1);
      expand_row4(r0 + 2 * (size_t)lo_pitch, i, xs, e2);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        ea[k] = pm::expand_v(1, e0[k], e1[k], e2[k]);
        eb[k] = pm::expand_v(0, e0[k], e1[k], e2[k]);
      }
    } else {
      if (has_a) expand4(lo + co, lo_pitch, xs, ys, i, ya, ea);
      if (has_b) expand4(lo + co, lo_pitch, xs, ys, i, yb, eb);
    }
    float4* da = reinterpret_cast<float4*>(hi + co +
                                           (size_t)(has_a ? ya : 0) * hp + 4 * i);
    float4* db = reinterpret_cast<float4*>(hi + co +
                                           (size_t)(has_b ? yb : 0) * hp + 4 * i);
    float4 ha = make_float4(0, 0, 0, 0), hb = ha;
    if (has_a) ha = *da;
    if (has_b) hb = *db;
    const float4 ra =
        make_float4(ea[0] - ha.x, ea[1] - ha.y, ea[2] - ha.z, ea[3] - ha.w);
    const float4 rb =
        make_float4(eb[0] - hb.x, eb[1] - hb.y, eb[2] - hb.z, eb[3] - hb.w);
    if (has_a) *da = ra;
    if (has_b) *db = rb;
    if (kSum) {
      const int x = 4 * i;
      if (has_a) {
        const uint8_t* m = xor_mask + (size_t)ya * W;
        if (x < W && m[x]) acc += ra.x;
        if (x + 1 < W && m[x + 1]) acc += ra.y;
        if (x + 2 < W && m[x + 2]) acc += ra.z;
        if (x + 3 < W && m[x + 3]) acc += ra.w;
      }
      if (has_b) {
        const uint8_t* m = xor_mask + (size_t)yb * W;
        if (x < W && m[x]) acc += rb.x;
        if (x + 1 < W && m[x + 1]) acc += rb.y;
        if (x + 2 < W && m[x + 2]) acc += rb.z;
        if (x + 3 < W && m[x + 3]) acc += rb.w;
      }
    }
  }
  if (kSum) {
    // one partial per warp (a warp is one row pair of the block): no block
    // barrier
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(~0u, acc, o);
    if (threadIdx.x == 0)
      partials[(((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x +
                blockIdx.x) * blockDim.y + threadIdx.y] = acc;
  }
}

// row pairs of a level with h rows and parity shift ys: pair p holds rows
// 2p - ys and 2p - ys + 1
static int laplace_pairs(int h, int ys) { return (h - 1 + ys) / 2 + 1; }

void launch_laplace(const PyrGeom& g, float* pyr, int level, int channels,
                    cudaStream_t s) {
  const LevelGeom& H = g.lv[level];
  const LevelGeom& L = g.lv[level + 1];
  dim3 b(32, 8);
  laplace_kernel<false><<<grid3(H.pitch / 4, laplace_pairs(H.h, H.ys),
                                channels, b), b, 0, s>>>(
      pyr + H.offset, H.h, H.pitch, H.xs, H.ys, pyr + L.offset, L.pitch,
      g.channel_floats, nullptr, 0, nullptr);
  XP_LAUNCHED();
}

// ---------------------------------------------------------------------------
// mask pyramid
// ---------------------------------------------------------------------------
struct LabelSrc {
  const uint8_t* labels;
  int W, H, image;
  __device__ __forceinline__ float at(int r, int c) const {
    return labels[(size_t)r * W + c] == image ? 1.0f : 0.0f;
  }
  __device__ __forceinline__ int width() const { return W; }
  __device__ __forceinline__ int height() const { return H; }
};

struct RectSrc {
  const float* data;
  int x0, y0, w, h, cw, ch;
  __device__ __forceinline__ float at(int r, int c) const {
    const int rx = c - x0, ry = r - y0;
    if ((unsigned)rx >= (unsigned)w || (unsigned)ry >= (unsigned)h) return 0.f;
    return data[(size_t)ry * w + rx];
  }
  __device__ __forceinline__ int width() const { return cw; }
  __device__ __forceinline__ int height() const { return ch; }
};

// A block produces 32 x 8 mask samples: the source window (19 rows x 67
// columns after clamping) is staged in shared memory once, squished
// horizontally per row, then combined vertically. The first line of a level
// is the squished first source line verbatim (functions.cpp:195-214).
constexpr int kMaskTileRows = 2 * 8 + 4, kMaskTileCols = 2 * 32 + 4;

// Uniformity of a label window, tested on the aligned 32-bit words that cover
// its rows (up to three bytes more on either side of a row: a conservative
// test -- a window that is uniform together with those bytes is uniform).
// -> 1: every label is `image`, 0: none is, -1: mixed (or undecided).
// Nearly every tile is uniform; this test replaces the byte-wise staging of
// 19 x 67 labels for them.
template <class Src>
__device__ __forceinline__ int mask_window_uniform(const Src&, int, int, int,
                                                   int, int) {
  return -1;
}
template <>
__device__ __forceinline__ int mask_window_uniform<LabelSrc>(
    const LabelSrc& src, int rmin, int nrows, int cmin, int ncols, int tid) {
  constexpr int kWords = (kMaskTileCols + 3) / 4 + 1;  // words per row
  const unsigned pat = (unsigned)src.image * 0x01010101u;
  const size_t total = (size_t)src.W * src.H;
  bool all_one = true, all_zero = true;
  for (int i = tid; i < nrows * kWords; i += 256) {
    const int r = i / kWords, k = i - r * kWords;
    const size_t first = (size_t)(rmin + r) * src.W + cmin;
    const size_t w0 = first >> 2, w1 = (first + ncols - 1) >> 2;
    const size_t w = w0 + k;
    if (w <= w1 && 4 * w + 3 < total) {
      const unsigned v =
          __ldg(reinterpret_cast<const unsigned*>(src.labels) + w);
      all_one = all_one && v == pat;
      all_zero = all_zero && __vcmpeq4(v, pat) == 0u;
    } else if (w <= w1) {
      all_one = all_zero = false;  // last bytes of the map: slow path
    }
  }
  all_one = __syncthreads_and(all_one);
  all_zero = __syncthreads_and(all_zero);
  return all_one ? 1 : (all_zero ? 0 : -1);
}

template <class Src>
__device__ __forceinline__ void mask_shrink_tile(const Src& src,
                                                 const MaskLevel& out,
                                                 float* dst) {
  __shared__ float tile[kMaskTileRows][kMaskTileCols + 1];
  __shared__ float hs[kMaskTileRows][32];
  const int tid = threadIdx.y * 32 + threadIdx.x;
  const int X0 = out.x0 + blockIdx.x * 32, Y0 = out.y0 + blockIdx.y * 8;
  const int wmax = src.width() - 1, hmax = src.height() - 1;
  const int rmin = pm::clampi(2 * Y0 - 4, 0, hmax);
  const int rmax = pm::clampi(2 * (Y0 + 7), 0, hmax);
  const int cmin = pm::clampi(2 * X0 - 4, 0, wmax);
  const int cmax = pm::clampi(2 * (X0 + 31), 0, wmax);
  const int nrows = rmax - rmin + 1, ncols = cmax - cmin + 1;
  {
    const int u = mask_window_uniform(src, rmin, nrows, cmin, ncols, tid);
    if (u >= 0) {
      const int x = blockIdx.x * 32 + threadIdx.x;
      const int y = blockIdx.y * 8 + threadIdx.y;
      if (x < out.w && y < out.h)
        dst[(size_t)y * out.w + x] = u ? 1.0f : 0.0f;
      return;
    }
  }
  bool all_one = true, all_zero = true;
  for (int i = tid; i < nrows * kMaskTileCols; i += 256) {
    const int r = i / kMaskTileCols, c = i - r * kMaskTileCols;
    if (c < ncols) {
      const float v = src.at(rmin + r, cmin + c);
      tile[r][c] = v;
      all_one = all_one && v == 1.0f;
      all_zero = all_zero && v == 0.0f;
    }
  }
  // A window of equal weights (inside or outside the owner region: nearly
  // every tile) gives that weight for every sample: the horizontal rule
  // copies it, the vertical taps sum to one exactly for 0 and 1.
  all_one = __syncthreads_and(all_one);
  all_zero = __syncthreads_and(all_zero);
  if (all_one || all_zero) {
    const int x = blockIdx.x * 32 + threadIdx.x;
    const int y = blockIdx.y * 8 + threadIdx.y;
    if (x < out.w && y < out.h)
      dst[(size_t)y * out.w + x] = all_one ? 1.0f : 0.0f;
    return;
  }
  for (int i = tid; i < nrows * 32; i += 256) {
    const int r = i >> 5, x = i & 31;
    const int X = X0 + x;
    float t[5];
#pragma unroll
    for (int k = 0; k < 5; ++k)
      t[k] = tile[r][pm::clampi(2 * X - 4 + k, 0, wmax) - cmin];
    hs[r][x] = pm::mask_h_combine(t[0], t[1], t[2], t[3], t[4]);
  }
  __syncthreads();
  const int x = blockIdx.x * 32 + threadIdx.x;
  const int y = blockIdx.y * 8 + threadIdx.y;
  if (x >= out.w || y >= out.h) return;
  const int Y = out.y0 + y;
  float v;
  if (Y == 0) {
    v = hs[0 - rmin][threadIdx.x];
  } else {
    float t[5];
#pragma unroll
    for (int k = 0; k < 5; ++k)
      t[k] = hs[pm::clampi(2 * Y - 4 + k, 0, hmax) - rmin][threadIdx.x];
    v = pm::mask_v_combine(t[0], t[1], t[2], t[3], t[4]);
  }
  dst[(size_t)y * out.w + x] = v;
}

template <class Src>
__global__ void __launch_bounds__(256)
    mask_shrink_kernel(Src src, MaskLevel out, float* dst) {
  mask_shrink_tile(src, out, dst);
}

// one level of every image's mask pyramid in one launch (blockIdx.z = image)
template <class Src>
struct MaskJob {
  Src src;
  MaskLevel out;
  float* dst;
};

template <class Src>
__global__ void __launch_bounds__(256)
    mask_shrink_batch_kernel(const MaskJob<Src>* __restrict__ jobs) {
  const MaskJob<Src> j = jobs[blockIdx.z];
  if ((int)blockIdx.x * 32 >= j.out.w || (int)blockIdx.y * 8 >= j.out.h) return;
  mask_shrink_tile(j.src, j.out, j.dst);
}

size_t mask_batch_bytes(int n) {
  return (size_t)n * std::max(sizeof(MaskJob<LabelSrc>), sizeof(MaskJob<RectSrc>));
}

// levels[i] / masks[i]: mask levels and buffer of image i. `d_jobs` is device
// scratch of mask_batch_bytes(n) bytes per level, filled through the mailbox.
void launch_mask_pyramids(const uint8_t* labels, int CW, int CH, int n,
                          int n_levels,
                          const std::vector<std::vector<MaskLevel>>& levels,
                          const std::vector<float*>& masks, void* d_jobs,
                          Mailbox& mbox, cudaStream_t s) {
  for (int l = 1; l < n_levels; ++l) {
    int gx = 0, gy = 0;
    for (int i = 0; i < n; ++i) {
      gx = std::max(gx, div_up(levels[i][l].w, 32));
      gy = std::max(gy, div_up(levels[i][l].h, 8));
    }
    if (gx == 0 || gy == 0) continue;
    unsigned char* slot =
        static_cast<unsigned char*>(d_jobs) + (size_t)(l - 1) * mask_batch_bytes(n);
    dim3 b(32, 8), g(gx, gy, n);
    if (l == 1) {
      std::vector<MaskJob<LabelSrc>> jobs(n);
      for (int i = 0; i < n; ++i)
        jobs[i] = MaskJob<LabelSrc>{LabelSrc{labels, CW, CH, i}, levels[i][1],
                                    masks[i] + levels[i][1].offset};
      mbox.to_device(slot, jobs.data(), sizeof(jobs[0]) * n, s);
      mask_shrink_batch_kernel<<<g, b, 0, s>>>(
          reinterpret_cast<const MaskJob<LabelSrc>*>(slot));
    } else {
      std::vector<MaskJob<RectSrc>> jobs(n);
      for (int i = 0; i < n; ++i) {
        const MaskLevel& in = levels[i][l - 1];
        jobs[i] = MaskJob<RectSrc>{
            RectSrc{masks[i] + in.offset, in.x0, in.y0, in.w, in.h, in.cw,
                    in.ch},
            levels[i][l], masks[i] + levels[i][l].offset};
      }
      mbox.to_device(slot, jobs.data(), sizeof(jobs[0]) * n, s);
      mask_shrink_batch_kernel<<<g, b, 0, s>>>(
          reinterpret_cast<const MaskJob<RectSrc>*>(slot));
    }
    XP_LAUNCHED();
  }
}

void launch_mask_level1(const uint8_t* labels, int CW, int CH, int image,
                        const MaskLevel& out, float* mask, cudaStream_t s) {
  if (out.w <= 0 || out.h <= 0) return;
  dim3 b(32, 8);
  LabelSrc src{labels, CW, CH, image};
  mask_shrink_kernel<<<grid3(out.w, out.h, 1, b), b, 0, s>>>(
      src, out, mask + out.offset);
  XP_LAUNCHED();
}

void launch_mask_shrink(const MaskLevel& in, const MaskLevel& out, float* mask,
                        cudaStream_t s) {
  if (out.w <= 0 || out.h <= 0) return;
  dim3 b(32, 8);
  RectSrc src{mask + in.offset, in.x0, in.y0, in.w, in.h, in.cw, in.ch};
  mask_shrink_kernel<<<grid3(out.w, out.h, 1, b), b, 0, s>>>(
      src, out, mask + out.offset);
  XP_LAUNCHED();
}

// ---------------------------------------------------------------------------
// composite: out (+)= in * mask over the mask's rectangle
// ---------------------------------------------------------------------------
__global__ void composite_kernel(const float* __restrict__ in, int in_w,
                                 int in_h, int in_pitch, size_t in_cs,
                                 float* __restrict__ out, int out_pitch,
                                 size_t out_cs, int x_off, int y_off,
                                 MaskLevel m, const float* __restrict__ mask,
                                 const uint8_t* __restrict__ labels,
                                 int image) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= m.w || y >= m.h) return;
  const int X = m.x0 + x, Y = m.y0 + y;
  float mv;
  if (labels)
    mv = labels[(size_t)Y * m.cw + X] == image ? 1.0f : 0.0f;
  else
    mv = mask[(size_t)y * m.w + x];
  if (mv == 0.0f) return;  // functions.cpp:347-351 (skip)
  const int ix = pm::clampi(X - x_off, 0, in_w - 1);
  const int iy = pm::clampi(Y - y_off, 0, in_h - 1);
  const size_t io = (size_t)iy * in_pitch + ix;
  const size_t oo = (size_t)Y * out_pitch + X;
#pragma unroll
  for (int c = 0; c < 3; ++c)
    out[oo + c * out_cs] = out[oo + c * out_cs] + in[io + c * in_cs] * mv;
}

void launch_composite(const PyrGeom& in_g, const float* in_pyr,
                      const PyrGeom& out_g, float* out_pyr, int level,
                      const uint8_t* labels, int image, const MaskLevel& m,
                      const float* mask, cudaStream_t s) {
  if (m.w <= 0 || m.h <= 0) return;
  const LevelGeom& I = in_g.lv[level];
  const LevelGeom& O = out_g.lv[level];
  const int x_off = (I.x - O.x) >> level;
  const int y_off = (I.y - O.y) >> level;
  dim3 b(32, 8);
  composite_kernel<<<grid3(m.w, m.h, 1, b), b, 0, s>>>(
      in_pyr + I.offset, I.w, I.h, I.pitch, in_g.channel_floats,
      out_pyr + O.offset, O.pitch, out_g.channel_floats, x_off, y_off, m,
      level == 0 ? nullptr : mask + m.offset, level == 0 ? labels : nullptr,
      image);
  XP_LAUNCHED();
}

// pad columns [w, pitch) with the last pixel (functions.cpp:403-410)
__global__ void pad_kernel(float* lvl, int w, int h, int pitch,
                           size_t chan_stride) {
  const int pad = pitch - w;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= pad * h) return;
  const int y = i / pad, x = w + i % pad;
  float* row = lvl + (size_t)blockIdx.z * chan_stride + (size_t)y * pitch;
  row[x] = row[w - 1];
}

void launch_pad_level(const PyrGeom& g, float* pyr, int level, int channels,
                      cudaStream_t s) {
  const LevelGeom& L = g.lv[level];
  const int pad = L.pitch - L.w;
  if (pad <= 0) return;
  pad_kernel<<<dim3(div_up(pad * L.h, 256), 1, channels), 256, 0, s>>>(
      pyr + L.offset, L.w, L.h, L.pitch, g.channel_floats);
  XP_LAUNCHED();
}

// ---------------------------------------------------------------------------
// offset correction: double sum of level 0 over the xor mask
// (multiblend.cpp:1232-1258). Fixed-shape two-stage tree: deterministic.
// ---------------------------------------------------------------------------
constexpr int kSumBlocks = 1024;

__global__ void xor_sum_kernel(const float* __restrict__ lvl0, int pitch,
                               size_t chan_stride,
                               const uint8_t* __restrict__ xor_mask, int W,
                               int H, double* __restrict__ partials) {
  __shared__ double sh[3][8];
  const size_t n = (size_t)W * H;
  const size_t per = (n + gridDim.x - 1) / gridDim.x;
  const size_t beg = per * blockIdx.x;
  const size_t end = beg + per < n ? beg + per : n;
  double acc[3] = {0, 0, 0};
  for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
    if (xor_mask[i]) {
      const size_t o = (i / W) * (size_t)pitch + (i % W);
      acc[0] += lvl0[o];
      acc[1] += lvl0[o + chan_stride];
      acc[2] += lvl0[o + 2 * chan_stride];
    }
  }
  for (int c = 0; c < 3; ++c) {
    double v = acc[c];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(~0u, v, o);
    if ((threadIdx.x & 31) == 0) sh[c][threadIdx.x >> 5] = v;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    double v = 0;
    for (int k = 0; k < 8; ++k) v += sh[threadIdx.x][k];
    partials[threadIdx.x * gridDim.x + blockIdx.x] = v;
  }
}

__global__ void __launch_bounds__(1024)
    xor_sum_final_kernel(const double* __restrict__ partials, int n,
                         double* sums) {
  // one block per channel; fixed assignment and a fixed tree: deterministic
  __shared__ double sh[1024];
  const int c = blockIdx.x, t = threadIdx.x;
  double v = 0;
  for (int i = t; i < n; i += 1024) v += partials[(size_t)c * n + i];
  sh[t] = v;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (t < o) sh[t] += sh[t + o];
    __syncthreads();
  }
  if (t == 0) sums[c] = sh[0];
}

size_t collapse0_partials(const PyrGeom& g) {
  const LevelGeom& H = g.lv[0];
  return (size_t)3 * div_up(H.pitch / 4, 32) *
         div_up(laplace_pairs(H.h, H.ys), 8) * 8;
}

// last collapse step (level 1 -> 0 of the canvas pyramid) + the xor sums
void launch_collapse0_sum(const PyrGeom& g, float* pyr,
                          const uint8_t* xor_mask, int W, double* d_partials,
                          double* d_sums, cudaStream_t s) {
  const LevelGeom& H = g.lv[0];
  const LevelGeom& L = g.lv[1];
  dim3 b(32, 8);
  dim3 gr = grid3(H.pitch / 4, laplace_pairs(H.h, H.ys), 3, b);
  laplace_kernel<true><<<gr, b, 0, s>>>(pyr + H.offset, H.h, H.pitch, H.xs,
                                        H.ys, pyr + L.offset, L.pitch,
                                        g.channel_floats, xor_mask, W,
                                        d_partials);
  XP_LAUNCHED();
  xor_sum_final_kernel<<<3, 1024, 0, s>>>(d_partials, (int)(gr.x * gr.y * 8),
                                        d_sums);
  XP_LAUNCHED();
}

void launch_xor_sum(const PyrGeom& g, const float* pyr,
                    const uint8_t* xor_mask, int W, int H, double* d_partials,
                    double* d_sums, cudaStream_t s) {
  const LevelGeom& L = g.lv[0];
  xor_sum_kernel<<<kSumBlocks, 256, 0, s>>>(pyr + L.offset, L.pitch,
                                            g.channel_floats, xor_mask, W, H,
                                            d_partials);
  XP_LAUNCHED();
  xor_sum_final_kernel<<<3, 1024, 0, s>>>(d_partials, kSumBlocks, d_sums);
  XP_LAUNCHED();
}

// ---------------------------------------------------------------------------
// Out: offset add + dither + clamp + round-half-even, planar and packed BGR
// with pixels outside the mask zeroed (pyramid.cpp:1563-1794,
// blenders.cc:64-77,151-154)
// ---------------------------------------------------------------------------
// Four consecutive pixels of the flat H*W output per thread: the plane and
// BGR stores are aligned vectors whatever W is; the level-0 rows (pitch != W)
// are read as scalars.
template <typename T>
struct OutVec;
template <>
struct OutVec<uint8_t> {
  static __device__ __forceinline__ void plane4(uint8_t* p, const uint8_t q[4]) {
    *reinterpret_cast<uchar4*>(p) = make_uchar4(q[0], q[1], q[2], q[3]);
  }
  static __device__ __forceinline__ void bgr12(uint8_t* p, const uint8_t v[12]) {
    uint32_t w[3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
      w[k] = v[4 * k] | (v[4 * k + 1] << 8) | (v[4 * k + 2] << 16) |
             ((uint32_t)v[4 * k + 3] << 24);
    uint32_t* d = reinterpret_cast<uint32_t*>(p);
    d[0] = w[0], d[1] = w[1], d[2] = w[2];
  }
};
template <>
struct OutVec<uint16_t> {
  static __device__ __forceinline__ void plane4(uint16_t* p,
                                                const uint16_t q[4]) {
    *reinterpret_cast<ushort4*>(p) = make_ushort4(q[0], q[1], q[2], q[3]);
  }
  static __device__ __forceinline__ void bgr12(uint16_t* p,
                                               const uint16_t v[12]) {
    uint2* d = reinterpret_cast<uint2*>(p);
#pragma unroll
    for (int k = 0; k < 3; ++k)
      d[k] = make_uint2(v[4 * k] | ((uint32_t)v[4 * k + 1] << 16),
                        v[4 * k + 2] | ((uint32_t)v[4 * k + 3] << 16));
  }
};

// pm::dither_value as a constant table (pyramid.cpp:1621-1634)
__constant__ float kDitherTable[16] = {
    -0.125f, 0.375f,  0.0f,     0.4999f, 0.125f,  -0.375f,  0.25f,   -0.25f,
    -0.0625f, 0.4375f, -0.1875f, 0.3125f, 0.1875f, -0.3125f, 0.0625f, -0.4375f};

// offset correction (multiblend.cpp:1259-1267): mean of the inputs minus mean
// of the collapsed level 0 over the exclusively covered pixels, per channel;
// out[3] != 0 tells the out kernel to apply it
__global__ void offset_kernel(const CanvasStats* __restrict__ stats,
                              const double* __restrict__ sums, int apply_add,
                              int in_bpp, int out_bpp, float* out) {
  const int c = threadIdx.x;
  const unsigned long long total_px = stats->totals[3];
  const bool on = apply_add && total_px != 0;
  if (c < 3) {
    float add = 0.0f;
    if (on) {
      float avg = __ull2float_rn(stats->totals[c]) / __ull2float_rn(total_px);
      if (out_bpp != in_bpp) avg = out_bpp == 8 ? avg / 256 : avg * 256;
      const float output_avg =
          __double2float_rn(sums[c]) / __ull2float_rn(total_px);
      add = avg - output_avg;
    }
    out[c] = add;
  } else if (c == 3) {
    out[3] = on ? 1.0f : 0.0f;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
    out_kernel(const float* __restrict__ lvl0, int pitch, size_t chan_stride,
               int W, int H, int dither, const float* __restrict__ offset,
               const uint8_t* __restrict__ full_mask, T* __restrict__ planes,
               T* __restrict__ bgr) {
  const float maxv = sizeof(T) == 1 ? 255.0f : 65535.0f;
  // offset correction, the same for every pixel: add[3] + "apply" flag
  const float add0 = __ldg(offset + 0), add1 = __ldg(offset + 1),
              add2 = __ldg(offset + 2);
  const bool adding = __ldg(offset + 3) != 0.0f;
  const float s_add[3] = {add0, add1, add2};
  // the canvas has fewer than 2^32 pixels (checked by the blender)
  const unsigned n = (unsigned)W * (unsigned)H;
  const unsigned p0 = 4u * (blockIdx.x * blockDim.x + threadIdx.x);
  if (p0 >= n) return;
  int y = (int)(p0 / (unsigned)W), x = (int)(p0 - (unsigned)y * (unsigned)W);
  const unsigned cs = (unsigned)chan_stride;
  T q[3][4];
  bool masked[4];
  int cnt = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (p0 + k < n) {
      masked[k] = full_mask[p0 + k] != 0;
      const float dv = dither ? kDitherTable[(y & 3) * 4 + (x & 3)] : 0.0f;
      const unsigned o = (unsigned)y * (unsigned)pitch + (unsigned)x;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float v = lvl0[o + c * cs];
        if (adding) v = s_add[c] + v;
        if (dither) v = v + dv;
        v = v > 0.0f ? v : 0.0f;
        v = v < maxv ? v : maxv;
        q[c][k] = (T)__float2int_rn(v);
      }
      ++cnt;
      if (++x == W) x = 0, ++y;
    } else {
      masked[k] = false;
      q[0][k] = q[1][k] = q[2][k] = 0;
    }
  }
  if (cnt == 4) {
    if ((n & 3) == 0) {  // plane bases c*n keep the vector alignment
#pragma unroll
      for (int c = 0; c < 3; ++c)
        OutVec<T>::plane4(planes + ((size_t)c * n + p0), q[c]);
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int k = 0; k < 4; ++k) planes[(size_t)c * n + p0 + k] = q[c][k];
    }
    T v[12];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) v[3 * k + c] = masked[k] ? q[c][k] : (T)0;
    OutVec<T>::bgr12(bgr + 3 * (size_t)p0, v);
  } else {
    for (int k = 0; k < cnt; ++k)
      for (int c = 0; c < 3; ++c) {
        planes[(size_t)c * n + p0 + k] = q[c][k];
        bgr[3 * ((size_t)p0 + k) + c] = masked[k] ? q[c][k] : (T)0;
      }
  }
}

void launch_out(const PyrGeom& g, const float* pyr, int W, int H, int out_bpp,
                int dither, int apply_add, const CanvasStats* d_stats,
                const double* d_sums, int in_bpp, const uint8_t* full_mask,
                float* d_offset, void* planes, void* bgr, cudaStream_t s) {
  const LevelGeom& L = g.lv[0];
  const size_t groups = div_up((size_t)W * H, (size_t)4);
  const unsigned blocks = (unsigned)div_up(groups, (size_t)256);
  offset_kernel<<<1, 32, 0, s>>>(d_stats, d_sums, apply_add, in_bpp, out_bpp,
                                 d_offset);
  XP_LAUNCHED();
  if (out_bpp == 8)
    out_kernel<uint8_t><<<blocks, 256, 0, s>>>(
        pyr + L.offset, L.pitch, g.channel_floats, W, H, dither, d_offset,
        full_mask, (uint8_t*)planes, (uint8_t*)bgr);
  else
    out_kernel<uint16_t><<<blocks, 256, 0, s>>>(
        pyr + L.offset, L.pitch, g.channel_floats, W, H, dither, d_offset,
        full_mask, (uint16_t*)planes, (uint16_t*)bgr);
  XP_LAUNCHED();
}

// ---------------------------------------------------------------------------
// debug / packing helpers
// ---------------------------------------------------------------------------
template <int BPP>
__global__ void extract_chan_kernel(ImgDev im, const u64* src_index,
                                    int channel, void* out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= im.w || y >= im.h) return;
  size_t idx = src_index ? (size_t)(u32)src_index[(size_t)y * im.w + x]
                         : (size_t)(im.top + y) * im.tiff_w + im.left + x;
  if (BPP == 8)
    static_cast<uint8_t*>(out)[(size_t)y * im.w + x] =
        (uint8_t)(static_cast<const u32*>(im.data)[idx] >> (8 * channel));
  else
    static_cast<uint16_t*>(out)[(size_t)y * im.w + x] =
        (uint16_t)(static_cast<const u64*>(im.data)[idx] >> (16 * channel));
}

void launch_extract_chan(const ImgDev& im, int bpp, const u64* src_index,
                         int channel, void* out, cudaStream_t s) {
  dim3 b(32, 8);
  dim3 g = grid3(im.w, im.h, 1, b);
  if (bpp == 8)
    extract_chan_kernel<8><<<g, b, 0, s>>>(im, src_index, channel, out);
  else
    extract_chan_kernel<16><<<g, b, 0, s>>>(im, src_index, channel, out);
  XP_LAUNCHED();
}

template <int BPP>
__global__ void extract_opaque_kernel(ImgDev im, uint8_t* out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= im.w || y >= im.h) return;
  out[(size_t)y * im.w + x] =
      px_opaque<BPP>(im.data, (size_t)(im.top + y) * im.tiff_w + im.left + x);
}

void launch_extract_opaque(const ImgDev& im, int bpp, uint8_t* out,
                           cudaStream_t s) {
  dim3 b(32, 8);
  dim3 g = grid3(im.w, im.h, 1, b);
  if (bpp == 8)
    extract_opaque_kernel<8><<<g, b, 0, s>>>(im, out);
  else
    extract_opaque_kernel<16><<<g, b, 0, s>>>(im, out);
  XP_LAUNCHED();
}

__global__ void labels_to_i32_kernel(const uint8_t* labels, size_t n,
                                     int32_t* out) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) out[i] = labels[i];
}

void launch_labels_to_i32(const uint8_t* labels, size_t n, int32_t* out,
                          cudaStream_t s) {
  labels_to_i32_kernel<<<(unsigned)div_up(n, (size_t)256), 256, 0, s>>>(
      labels, n, out);
  XP_LAUNCHED();
}

// BGR(+mask) or already-interleaved samples -> 4-sample pixels
// (blenders.cc:99-112: alpha = mask != 0 ? max : 0)
template <typename T>
__global__ void pack_bgra_kernel(const T* __restrict__ src, size_t src_pitch,
                                 const uint8_t* __restrict__ mask,
                                 size_t mask_pitch, int w, int h, int spp_in,
                                 T* __restrict__ out, int out_pitch) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= w || y >= h) return;
  const T amax = (T)(sizeof(T) == 1 ? 255 : 65535);
  const T* s = src + (size_t)y * src_pitch + (size_t)x * spp_in;
  T a = amax;
  if (spp_in == 4) a = s[3];
  if (mask) a = mask[(size_t)y * mask_pitch + x] != 0 ? amax : (T)0;
  T* d = out + ((size_t)y * out_pitch + x) * 4;
  d[0] = s[0];
  d[1] = s[1];
  d[2] = s[2];
  d[3] = a;
}

void launch_pack_bgra(const void* bgr, size_t bgr_pitch, const uint8_t* mask,
                      size_t mask_pitch, int w, int h, int depth, int spp_in,
                      void* out, int out_pitch, cudaStream_t s) {
  dim3 b(32, 8);
  dim3 g = grid3(w, h, 1, b);
  if (depth == 8)
    pack_bgra_kernel<uint8_t><<<g, b, 0, s>>>((const uint8_t*)bgr, bgr_pitch,
                                              mask, mask_pitch, w, h, spp_in,
                                              (uint8_t*)out, out_pitch);
  else
    pack_bgra_kernel<uint16_t><<<g, b, 0, s>>>(
        (const uint16_t*)bgr, bgr_pitch / 2, mask, mask_pitch, w, h, spp_in,
        (uint16_t*)out, out_pitch);
  XP_LAUNCHED();
}

// debug: level 0 after the offset add (what the oracle keeps as "collapsed")
__global__ void debug_collapsed_kernel(const float* __restrict__ lvl0,
                                       size_t n, size_t chan_stride,
                                       int apply_add,
                                       const CanvasStats* __restrict__ stats,
                                       const double* __restrict__ sums,
                                       int in_bpp, int out_bpp,
                                       float* __restrict__ out) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = blockIdx.y;
  float v = lvl0[i + c * chan_stride];
  const unsigned long long total_px = stats->totals[3];
  if (apply_add && total_px != 0) {
    float avg = __ull2float_rn(stats->totals[c]) / __ull2float_rn(total_px);
    if (out_bpp != in_bpp) avg = out_bpp == 8 ? avg / 256 : avg * 256;
    const float output_avg =
        __double2float_rn(sums[c]) / __ull2float_rn(total_px);
    v = (avg - output_avg) + v;
  }
  out[i + c * n] = v;
}

void launch_debug_collapsed(const PyrGeom& g, const float* pyr, int apply_add,
                            const CanvasStats* d_stats, const double* d_sums,
                            int in_bpp, int out_bpp, float* out,
                            cudaStream_t s) {
  const LevelGeom& L = g.lv[0];
  const size_t n = (size_t)L.pitch * L.h;
  debug_collapsed_kernel<<<dim3((unsigned)div_up(n, (size_t)256), 3), 256, 0,
                           s>>>(pyr + L.offset, n, g.channel_floats, apply_add,
                                d_stats, d_sums, in_bpp, out_bpp, out);
  XP_LAUNCHED();
}

}  // namespace xp
