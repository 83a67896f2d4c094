// xpano's second blender, blenders::MultiBandOpenCV (xpano/algorithm/
// blenders.h:41-50, blenders.cc:160-175; BlendingMethod::kOpenCV,
// algorithm.cc:135-136): cv::detail::MultiBandBlender(try_gpu = false,
// num_bands = 5, weight_type = CV_32F) fed with CV_16SC3 images, the result
// converted to CV_8U. SURVEY.md 8(f4).
//
// Same arithmetic as OpenCV's CPU code (restated in oracle/mbcv_oracle.py,
// which is pinned against cv2): 16-bit integer Laplacian pyramids
// ([1 4 6 4 1] pyrDown with (v + 128) >> 8, pyrUp with (v + 32) >> 6,
// saturating subtract / add), float weight pyramids, the truncating
// short(src * weight) accumulation and short(sum / (weight + eps))
// normalisation. Everything lives on the device between feed and fetch;
// Laplacian levels are never stored per image: pyrUp, subtract, weighting and
// accumulation into the panorama's pyramid are one kernel per level.
#include <algorithm>
#include <cmath>
#include <vector>

#include "common.h"

namespace xp {

namespace {

__device__ __forceinline__ int r101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
  return i;
}

__device__ __forceinline__ short sat16(int v) {
  return (short)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v));
}

// Level 0 of one feed: the image with a BORDER_REFLECT frame as 16-bit
// samples, and the weight map mask / 255 with a zero frame.
__global__ void mbcv_border_kernel(const uint8_t* __restrict__ img,
                                   size_t istride,
                                   const uint8_t* __restrict__ mask,
                                   size_t mstride, int w, int h, int left,
                                   int top, int W, int H,
                                   short* __restrict__ dst,
                                   float* __restrict__ wdst) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= W || y >= H) return;
  int sx = x - left, sy = y - top;
  const bool inside = sx >= 0 && sx < w && sy >= 0 && sy < h;
  // BORDER_REFLECT: fedcba|abcdefgh|hgfedcb
  while (sx < 0 || sx >= w) sx = sx < 0 ? -sx - 1 : 2 * w - 1 - sx;
  while (sy < 0 || sy >= h) sy = sy < 0 ? -sy - 1 : 2 * h - 1 - sy;
  const uint8_t* p = img + (size_t)sy * istride + 3 * (size_t)sx;
  short* d = dst + ((size_t)y * W + x) * 3;
  d[0] = p[0];
  d[1] = p[1];
  d[2] = p[2];
  // convertTo(CV_32F, 1. / 255.): float(v * double(1 / 255.))
  wdst[(size_t)y * W + x] =
      inside ? (float)((double)mask[(size_t)sy * mstride + sx] * (1. / 255.))
             : 0.f;
}

// cv::pyrDown on CV_16SC3 (one thread per destination sample)
__global__ void mbcv_down16_kernel(const short* __restrict__ src, int sw,
                                   int sh, short* __restrict__ dst, int dw,
                                   int dh) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // sample in the row
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= dw * 3 || y >= dh) return;
  const int x = i / 3, c = i - 3 * x;
  int cx[5], v = 0;
#pragma unroll
  for (int k = 0; k < 5; ++k) cx[k] = r101(2 * x + k - 2, sw) * 3 + c;
  const int wgt[5] = {1, 4, 6, 4, 1};
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const short* r = src + (size_t)r101(2 * y + k - 2, sh) * sw * 3;
    const int row = r[cx[2]] * 6 + (r[cx[1]] + r[cx[3]]) * 4 + r[cx[0]] +
                    r[cx[4]];
    v += row * wgt[k];
  }
  dst[(size_t)y * dw * 3 + i] = (short)((v + 128) >> 8);
}

// cv::pyrDown on CV_32F, the scalar order of pyramids.cpp
__global__ void mbcv_down32_kernel(const float* __restrict__ src, int sw,
                                   int sh, float* __restrict__ dst, int dw,
                                   int dh) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= dw || y >= dh) return;
  int cx[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) cx[k] = r101(2 * x + k - 2, sw);
  float row[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const float* r = src + (size_t)r101(2 * y + k - 2, sh) * sw;
    row[k] = __fadd_rn(
        __fadd_rn(__fadd_rn(__fmul_rn(r[cx[2]], 6.f),
                            __fmul_rn(__fadd_rn(r[cx[1]], r[cx[3]]), 4.f)),
                  r[cx[0]]),
        r[cx[4]]);
  }
  const float v = __fadd_rn(
      __fadd_rn(__fadd_rn(__fmul_rn(row[2], 6.f),
                          __fmul_rn(__fadd_rn(row[1], row[3]), 4.f)),
                row[0]),
      row[4]);
  dst[(size_t)y * dw + x] = __fmul_rn(v, 1.f / 256.f);
}

// cv::pyrUp to exactly twice the size, one sample: left / top neighbours
// reflect (101), right / bottom ones replicate
__device__ __forceinline__ int up16(const short* __restrict__ lo, int lw,
                                    int lh, int X, int Y, int c) {
  const int x = X >> 1, y = Y >> 1;
  const int xl = r101(x - 1, lw), xr = min(x + 1, lw - 1);
  const int yu = r101(y - 1, lh), yd = min(y + 1, lh - 1);
  auto hrow = [&](int yy) {
    const short* r = lo + (size_t)yy * lw * 3 + c;
    return (X & 1) ? (r[3 * x] + r[3 * xr]) * 4
                   : r[3 * xl] + r[3 * x] * 6 + r[3 * xr];
  };
  const int v = (Y & 1) ? (hrow(y) + hrow(yd)) * 4
                        : hrow(y) * 6 + hrow(yu) + hrow(yd);
  return (v + 32) >> 6;
}

// One level of MultiBandBlender::feed: the Laplacian sample
// sat(hi - pyrUp(lo)) (or hi itself at the top level), weighted and added
// into the panorama's pyramid; the weight is added to the band's weight map.
template <bool kTop>
__global__ void mbcv_accumulate_kernel(const short* __restrict__ hi, int w,
                                       int h, const short* __restrict__ lo,
                                       int lw, int lh,
                                       const float* __restrict__ wgt,
                                       short* __restrict__ dst,
                                       float* __restrict__ dwgt, int dW,
                                       int x0, int y0) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= w || y >= h) return;
  const float wv = wgt[(size_t)y * w + x];
  const short* s = hi + ((size_t)y * w + x) * 3;
  short* d = dst + ((size_t)(y + y0) * dW + (x + x0)) * 3;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    int lap = s[c];
    if (!kTop) lap = sat16(lap - up16(lo, lw, lh, x, y, c));
    // dst += static_cast<short>(src * weight)
    d[c] = (short)(d[c] + (short)__float2int_rz(__fmul_rn((float)lap, wv)));
  }
  float* dw = dwgt + (size_t)(y + y0) * dW + (x + x0);
  *dw = __fadd_rn(*dw, wv);
}

// normalizeUsingWeightMap: short(v / (weight + WEIGHT_EPS))
__global__ void mbcv_normalize_kernel(short* __restrict__ lap,
                                      const float* __restrict__ wgt, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float den = __fadd_rn(wgt[i], 1e-5f);
  short* p = lap + 3 * i;
#pragma unroll
  for (int c = 0; c < 3; ++c)
    p[c] = (short)__float2int_rz(__fdiv_rn((float)p[c], den));
}

// restoreImageFromLaplacePyr, one step: hi = sat(pyrUp(lo) + hi)
__global__ void mbcv_restore_kernel(short* __restrict__ hi, int w, int h,
                                    const short* __restrict__ lo, int lw,
                                    int lh) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= w * 3 || y >= h) return;
  const int x = i / 3, c = i - 3 * x;
  short* p = hi + (size_t)y * w * 3 + i;
  *p = sat16(up16(lo, lw, lh, x, y, c) + *p);
}

// MultiBandBlender::blend tail + MultiBandOpenCV::blend: mask = weight >
// WEIGHT_EPS, pixels outside it zero, convertTo(CV_8U)
__global__ void mbcv_out_kernel(const short* __restrict__ lap,
                                const float* __restrict__ wgt, int W, int w,
                                int h, uint8_t* __restrict__ bgr,
                                uint8_t* __restrict__ mask) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= w || y >= h) return;
  const bool on = wgt[(size_t)y * W + x] > 1e-5f;
  const short* s = lap + ((size_t)y * W + x) * 3;
  uint8_t* d = bgr + ((size_t)y * w + x) * 3;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int v = on ? s[c] : 0;
    d[c] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
  }
  mask[(size_t)y * w + x] = on ? 255 : 0;
}

}  // namespace

class MultiBand {
 public:
  int actual_num_bands = 5, device = 0;
  int nb = 0;
  int fx = 0, fy = 0, fw = 0, fh = 0;  // dst_roi_final_
  int rx = 0, ry = 0, rw = 0, rh = 0;  // dst_roi_ (padded)
  bool prepared = false, blended = false;
  std::vector<DevBuf> lap, wts;        // the panorama's pyramid and weights
  std::vector<int2> size;              // per level (w, h)
  DevBuf out_bgr, out_mask;
  cudaStream_t stream = nullptr;

  ~MultiBand() {
    if (stream) {
      int cur = 0;
      cudaGetDevice(&cur);
      if (cur != device) cudaSetDevice(device);
      cudaStreamSynchronize(stream);
      lap.clear();
      wts.clear();
      out_bgr.release();
      out_mask.release();
      cudaStreamDestroy(stream);
      if (cur != device) cudaSetDevice(cur);
    }
  }

  void prepare(int x, int y, int w, int h) {
    XP_REQUIRE(w > 0 && h > 0, "prepare: empty ROI");
    select_device(device);
    if (!stream) stream = create_stream(false);
    XP_CUDA(cudaStreamSynchronize(stream));
    fx = x, fy = y, fw = w, fh = h;
    const double max_len = (double)std::max(w, h);
    nb = std::min(actual_num_bands,
                  (int)std::ceil(std::log(max_len) / std::log(2.0)));
    const int m = 1 << nb;
    w += (m - w % m) % m;
    h += (m - h % m) % m;
    rx = x, ry = y, rw = w, rh = h;
    lap.clear();
    wts.clear();
    size.clear();
    lap.resize(nb + 1);
    wts.resize(nb + 1);
    int lw = w, lh = h;
    for (int i = 0; i <= nb; ++i) {
      size.push_back(make_int2(lw, lh));
      lap[i].reserve(sizeof(short) * 3 * (size_t)lw * lh);
      wts[i].reserve(sizeof(float) * (size_t)lw * lh);
      XP_CUDA(cudaMemsetAsync(lap[i].p, 0, sizeof(short) * 3 * (size_t)lw * lh,
                              stream));
      XP_CUDA(cudaMemsetAsync(wts[i].p, 0, sizeof(float) * (size_t)lw * lh,
                              stream));
      lw = (lw + 1) / 2;
      lh = (lh + 1) / 2;
    }
    prepared = true;
    blended = false;
  }

  void feed(const void* bgr, size_t stride, const uint8_t* mask,
            size_t mstride, int w, int h, int tlx, int tly, int mem) {
    XP_REQUIRE(prepared && !blended, "feed before prepare() or after blend()");
    XP_REQUIRE(bgr && mask && w > 0 && h > 0, "feed: empty image");
    select_device(device);
    if (stride == 0) stride = (size_t)w * 3;
    if (mstride == 0) mstride = (size_t)w;
    XP_REQUIRE(stride >= (size_t)w * 3 && mstride >= (size_t)w,
               "feed: stride smaller than a row");
    // the sub-rectangle of the panorama this image is kept in
    // (MultiBandBlender::feed): a gap of 3 * 2^bands, corners on multiples
    // of 2^bands so that every level is exactly half the previous one
    const int gap = 3 * (1 << nb), m = 1 << nb;
    int tnx = std::max(rx, tlx - gap), tny = std::max(ry, tly - gap);
    int bnx = std::min(rx + rw, tlx + w + gap);
    int bny = std::min(ry + rh, tly + h + gap);
    XP_REQUIRE(bnx > tnx && bny > tny,
               "feed: the image lies outside the prepared ROI");
    tnx = rx + (((tnx - rx) >> nb) << nb);
    tny = ry + (((tny - ry) >> nb) << nb);
    int width = bnx - tnx, height = bny - tny;
    width += (m - width % m) % m;
    height += (m - height % m) % m;
    bnx = tnx + width;
    bny = tny + height;
    const int dy = std::max(bny - (ry + rh), 0);
    const int dx = std::max(bnx - (rx + rw), 0);
    tnx -= dx, bnx -= dx, tny -= dy, bny -= dy;
    const int top = tly - tny, left = tlx - tnx;
    XP_REQUIRE(top >= 0 && left >= 0 && bny - tly - h >= 0 &&
                   bnx - tlx - w >= 0,
               "feed: the image does not fit the prepared ROI");

    cudaStream_t s = stream;
    const uint8_t* dimg = static_cast<const uint8_t*>(bgr);
    const uint8_t* dmask = mask;
    DevBuf simg, smask;
    if (mem == XP_MEM_HOST) {
      simg.reserve((size_t)w * 3 * h);
      smask.reserve((size_t)w * h);
      copy_in_2d(simg.p, (size_t)w * 3, bgr, stride, (size_t)w * 3, (size_t)h,
                 mem, s);
      copy_in_2d(smask.p, (size_t)w, mask, mstride, (size_t)w, (size_t)h, mem,
                 s);
      dimg = simg.as<uint8_t>();
      dmask = smask.as<uint8_t>();
      stride = (size_t)w * 3;
      mstride = (size_t)w;
    }
    // Gaussian pyramids of the framed image (16-bit) and of its weight map
    std::vector<DevBuf> g(nb + 1), gw(nb + 1);
    std::vector<int2> gs(nb + 1);
    int lw = width, lh = height;
    for (int i = 0; i <= nb; ++i) {
      gs[i] = make_int2(lw, lh);
      g[i].reserve(sizeof(short) * 3 * (size_t)lw * lh);
      gw[i].reserve(sizeof(float) * (size_t)lw * lh);
      lw = (lw + 1) / 2;
      lh = (lh + 1) / 2;
    }
    const dim3 b(32, 8);
    mbcv_border_kernel<<<dim3(div_up(width, 32), div_up(height, 8)), b, 0,
                         s>>>(dimg, stride, dmask, mstride, w, h, left, top,
                              width, height, g[0].as<short>(),
                              gw[0].as<float>());
    XP_LAUNCHED();
    for (int i = 0; i < nb; ++i) {
      mbcv_down16_kernel<<<dim3(div_up(gs[i + 1].x * 3, 32),
                                div_up(gs[i + 1].y, 8)),
                           b, 0, s>>>(g[i].as<short>(), gs[i].x, gs[i].y,
                                      g[i + 1].as<short>(), gs[i + 1].x,
                                      gs[i + 1].y);
      XP_LAUNCHED();
      mbcv_down32_kernel<<<dim3(div_up(gs[i + 1].x, 32),
                                div_up(gs[i + 1].y, 8)),
                           b, 0, s>>>(gw[i].as<float>(), gs[i].x, gs[i].y,
                                      gw[i + 1].as<float>(), gs[i + 1].x,
                                      gs[i + 1].y);
      XP_LAUNCHED();
    }
    int x0 = tnx - rx, y0 = tny - ry;
    for (int i = 0; i <= nb; ++i) {
      const dim3 grd(div_up(gs[i].x, 32), div_up(gs[i].y, 8));
      if (i < nb)
        mbcv_accumulate_kernel<false><<<grd, b, 0, s>>>(
            g[i].as<short>(), gs[i].x, gs[i].y, g[i + 1].as<short>(),
            gs[i + 1].x, gs[i + 1].y, gw[i].as<float>(), lap[i].as<short>(),
            wts[i].as<float>(), size[i].x, x0, y0);
      else
        mbcv_accumulate_kernel<true><<<grd, b, 0, s>>>(
            g[i].as<short>(), gs[i].x, gs[i].y, nullptr, 0, 0,
            gw[i].as<float>(), lap[i].as<short>(), wts[i].as<float>(),
            size[i].x, x0, y0);
      XP_LAUNCHED();
      x0 /= 2;
      y0 /= 2;
    }
    // the per-feed pyramids return to the pool when this function leaves
    XP_CUDA(cudaStreamSynchronize(s));
  }

  void blend(int out_wh[2]) {
    XP_REQUIRE(prepared, "blend before prepare()");
    XP_REQUIRE(!blended, "blend() already ran on this blender");
    select_device(device);
    cudaStream_t s = stream;
    for (int i = 0; i <= nb; ++i) {
      const size_t n = (size_t)size[i].x * size[i].y;
      mbcv_normalize_kernel<<<(unsigned)div_up(n, (size_t)256), 256, 0, s>>>(
          lap[i].as<short>(), wts[i].as<float>(), n);
      XP_LAUNCHED();
    }
    const dim3 b(32, 8);
    for (int i = nb; i > 0; --i) {
      mbcv_restore_kernel<<<dim3(div_up(size[i - 1].x * 3, 32),
                                 div_up(size[i - 1].y, 8)),
                            b, 0, s>>>(lap[i - 1].as<short>(), size[i - 1].x,
                                       size[i - 1].y, lap[i].as<short>(),
                                       size[i].x, size[i].y);
      XP_LAUNCHED();
    }
    out_bgr.reserve((size_t)fw * fh * 3);
    out_mask.reserve((size_t)fw * fh);
    mbcv_out_kernel<<<dim3(div_up(fw, 32), div_up(fh, 8)), b, 0, s>>>(
        lap[0].as<short>(), wts[0].as<float>(), size[0].x, fw, fh,
        out_bgr.as<uint8_t>(), out_mask.as<uint8_t>());
    XP_LAUNCHED();
    XP_CUDA(cudaStreamSynchronize(s));
    blended = true;
    if (out_wh) out_wh[0] = fw, out_wh[1] = fh;
  }

  void fetch(void* dst_bgr, size_t dst_stride, uint8_t* dst_mask,
             size_t mask_stride, int mem) {
    XP_REQUIRE(blended, "fetch before blend()");
    select_device(device);
    if (dst_stride == 0) dst_stride = (size_t)fw * 3;
    if (mask_stride == 0) mask_stride = (size_t)fw;
    if (dst_bgr)
      copy_out_2d(dst_bgr, dst_stride, out_bgr.p, (size_t)fw * 3,
                  (size_t)fw * 3, (size_t)fh, mem, stream);
    if (dst_mask)
      copy_out_2d(dst_mask, mask_stride, out_mask.p, (size_t)fw, (size_t)fw,
                  (size_t)fh, mem, stream);
    XP_CUDA(cudaStreamSynchronize(stream));
  }
};

}  // namespace xp

extern "C" {

xp_status xp_multiband_create(int num_bands, int device, xp_multiband** out) {
  return xp::guard([&] {
    XP_REQUIRE(out, "multiband_create: null argument");
    *out = nullptr;
    XP_REQUIRE(num_bands >= 0 && num_bands <= 16,
               "multiband_create: num_bands out of range");
    auto* b = new xp::MultiBand();
    b->actual_num_bands = num_bands;
    b->device = device;
    *out = reinterpret_cast<xp_multiband*>(b);
  });
}

void xp_multiband_destroy(xp_multiband* b) {
  delete reinterpret_cast<xp::MultiBand*>(b);
}

xp_status xp_multiband_prepare(xp_multiband* b, int x, int y, int w, int h) {
  return xp::guard([&] {
    XP_REQUIRE(b, "multiband_prepare: null handle");
    reinterpret_cast<xp::MultiBand*>(b)->prepare(x, y, w, h);
  });
}

xp_status xp_multiband_feed(xp_multiband* b, const void* bgr, size_t stride,
                            const uint8_t* mask, size_t mask_stride, int w,
                            int h, int tl_x, int tl_y, int mem) {
  return xp::guard([&] {
    XP_REQUIRE(b, "multiband_feed: null handle");
    reinterpret_cast<xp::MultiBand*>(b)->feed(bgr, stride, mask, mask_stride,
                                              w, h, tl_x, tl_y, mem);
  });
}

xp_status xp_multiband_blend(xp_multiband* b, int out_wh[2], int* num_bands) {
  return xp::guard([&] {
    XP_REQUIRE(b, "multiband_blend: null handle");
    auto* m = reinterpret_cast<xp::MultiBand*>(b);
    m->blend(out_wh);
    if (num_bands) *num_bands = m->nb;
  });
}

xp_status xp_multiband_fetch(xp_multiband* b, void* dst_bgr, size_t dst_stride,
                             uint8_t* dst_mask, size_t mask_stride, int mem) {
  return xp::guard([&] {
    XP_REQUIRE(b, "multiband_fetch: null handle");
    reinterpret_cast<xp::MultiBand*>(b)->fetch(dst_bgr, dst_stride, dst_mask,
                                               mask_stride, mem);
  });
}

}  // extern "C"
