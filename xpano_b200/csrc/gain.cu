// Exposure step of Stitcher::EstimateSeams (xpano/algorithm/stitcher.cc:
// 273-278): cv::detail::BlocksGainCompensator::feed / apply as xpano
// constructs it (stitcher.cc:200-201: 32x32 blocks, one feed, two gain
// filtering iterations). SURVEY.md 8(f1).
//
// The per-pixel work -- for every pair of overlapping blocks the number of
// pixels both masks cover and the two sums of |bgr| over them -- runs on the
// device, one CTA per block pair. What is left is a sparse symmetric positive
// definite system with one unknown per block (about a thousand for eight
// images at the seam scale) and a 3x3 smoothing of the block gains; both stay
// on the host in double / float as in the reference (cv::solve on doubles,
// sepFilter2D on floats). The result, one small float gain map per image, is
// what xp_composer_set_images / xp_warp_aux take.
#include <algorithm>
#include <cmath>
#include <vector>

#include "common.h"

namespace xp {

struct GainImage {
  int cx, cy, w, h;        // corner and size of the warped seam-scale image
  int nx, ny, bw, bh;      // block grid (BlocksCompensator::feed)
  int first;               // index of its first block
  const uint8_t* bgr;      // device, interleaved 8-bit BGR
  size_t stride;
  const uint8_t* mask;     // device, 8-bit (255 = inside)
  size_t mstride;
};

struct GainPair {
  int a, b;                // images
  int i, j;                // blocks, i <= j
  int x0, y0, w, h;        // intersection rectangle, panorama coordinates
};

struct GainPairDev {
  const uint8_t* pa;       // first pixel of the rectangle in image a
  const uint8_t* pb;
  const uint8_t* ma;
  const uint8_t* mb;
  unsigned sa, sb, sma, smb;  // strides in bytes
  int w, h;
};

struct GainStat {
  double s1, s2;
  long long count;
};

// One CTA per block pair: intersect = (mask_a == 255) & (mask_b == 255),
// count, sum |a|, sum |b| with |p| = sqrt(b^2 + g^2 + r^2) in double
// (cv::norm(Vec3b), GainCompensator::singleFeed).
__global__ void __launch_bounds__(128)
    gain_pairs_kernel(const GainPairDev* __restrict__ pairs,
                      GainStat* __restrict__ out) {
  const GainPairDev p = pairs[blockIdx.x];
  double s1 = 0, s2 = 0;
  int cnt = 0;
  const int n = p.w * p.h;
  for (int k = threadIdx.x; k < n; k += blockDim.x) {
    const int y = k / p.w, x = k - y * p.w;
    if (p.ma[(size_t)y * p.sma + x] != 255 ||
        p.mb[(size_t)y * p.smb + x] != 255)
      continue;
    const uint8_t* a = p.pa + (size_t)y * p.sa + 3 * x;
    const uint8_t* b = p.pb + (size_t)y * p.sb + 3 * x;
    ++cnt;
    s1 += sqrt((double)(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]));
    s2 += sqrt((double)(b[0] * b[0] + b[1] * b[1] + b[2] * b[2]));
  }
  for (int o = 16; o > 0; o >>= 1) {
    s1 += __shfl_down_sync(~0u, s1, o);
    s2 += __shfl_down_sync(~0u, s2, o);
    cnt += __shfl_down_sync(~0u, cnt, o);
  }
  __shared__ double sh1[4], sh2[4];
  __shared__ int shc[4];
  if ((threadIdx.x & 31) == 0) {
    sh1[threadIdx.x >> 5] = s1;
    sh2[threadIdx.x >> 5] = s2;
    shc[threadIdx.x >> 5] = cnt;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    GainStat r;
    r.s1 = (sh1[0] + sh1[1]) + (sh1[2] + sh1[3]);
    r.s2 = (sh2[0] + sh2[1]) + (sh2[2] + sh2[3]);
    r.count = shc[0] + shc[1] + shc[2] + shc[3];
    out[blockIdx.x] = r;
  }
}

// BlocksCompensator::apply on an interleaved BGR image in place: the gain map
// resized to the image size (INTER_LINEAR, float), multiply, saturate.
// Coefficients as in warp.cu's fused kernel (resizeGeneric_ HResizeLinear).
__device__ __forceinline__ void gain_coeff(int srcsize, int dstsize, int d,
                                           int& ofs, float& a1) {
  const double scale = 1.0 / ((double)dstsize / srcsize);
  float f = (float)(((double)d + 0.5) * scale - 0.5);
  int s = (int)floorf(f);
  f = __fsub_rn(f, (float)s);
  if (s < 0) {
    f = 0;
    s = 0;
  }
  if (s >= srcsize - 1) {
    f = 0;
    s = srcsize - 1;
  }
  ofs = s;
  a1 = f;
}

template <class T>
__global__ void gain_apply_kernel(T* __restrict__ img, int w, int h,
                                  size_t stride_elems,
                                  const float* __restrict__ gain, int gw,
                                  int gh) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= w || y >= h) return;
  float g;
  if (gw == w && gh == h) {
    g = gain[(size_t)y * gw + x];
  } else {
    int ox, oy;
    float al1, b1;
    gain_coeff(gw, w, x, ox, al1);
    gain_coeff(gh, h, y, oy, b1);
    const float al0 = __fsub_rn(1.f, al1), b0 = __fsub_rn(1.f, b1);
    const int ox1 = ox + 1 < gw ? ox + 1 : ox;
    const int oy1 = oy + 1 < gh ? oy + 1 : oy;
    const float* r0 = gain + (size_t)oy * gw;
    const float* r1 = gain + (size_t)oy1 * gw;
    const float h0 =
        __fadd_rn(__fmul_rn(r0[ox], al0), __fmul_rn(r0[ox1], al1));
    const float h1 =
        __fadd_rn(__fmul_rn(r1[ox], al0), __fmul_rn(r1[ox1], al1));
    g = __fadd_rn(__fmul_rn(h0, b0), __fmul_rn(h1, b1));
  }
  T* p = img + (size_t)y * stride_elems + 3 * (size_t)x;
  const int maxv = sizeof(T) == 1 ? 255 : 65535;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int q = __float2int_rn(__fmul_rn((float)p[c], g));
    p[c] = (T)(q < 0 ? 0 : (q > maxv ? maxv : q));
  }
}

class GainCompensator {
 public:
  int block_w = 32, block_h = 32, filter_iters = 2, device = 0;
  std::vector<std::vector<float>> maps;  // per image, ny x nx
  std::vector<int2> map_size;            // (nx, ny)
  std::vector<double> block_gains;       // before smoothing
  int n_pairs = 0, n_unknowns = 0, cg_iterations = 0;

  std::vector<GainImage> im;     // geometry of the last plan()
  std::vector<GainPair> pairs;   // block pairs of the last plan()
  int n_blocks = 0;

  // host: block grids and the overlapping block pairs
  void plan(int n, const int* corners, const int* widths, const int* heights);
  // host: gains and maps from the per-pair statistics
  void finish(const std::vector<GainStat>& stats);
  void feed(int n, const int* corners, const void* const* bgr,
            const int* widths, const int* heights, const size_t* strides,
            const uint8_t* const* masks, const size_t* mask_strides, int mem);
  void apply(int index, void* bgr, int w, int h, size_t stride, int depth,
             int mem) const;

 private:
  void solve(const std::vector<GainStat>& stats);
};

static int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
  return i;
}

// sepFilter2D(map, CV_32F, [.25 .5 .25], [.25 .5 .25]), BORDER_REFLECT_101:
// rows, then columns; each tap sum as centre * .5 + (left + right) * .25
// (the association OpenCV's symmetric 3-tap filters use; the float result
// is the same bits as cv2's on every case the tests hold).
static void smooth_map(std::vector<float>& m, int w, int h) {
  std::vector<float> rows(m.size());
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      const float* r = &m[(size_t)y * w];
      const float side = r[reflect101(x - 1, w)] + r[reflect101(x + 1, w)];
      const float c = r[x] * 0.5f;
      const float s = side * 0.25f;
      rows[(size_t)y * w + x] = c + s;
    }
  for (int y = 0; y < h; ++y) {
    const float* u = &rows[(size_t)reflect101(y - 1, h) * w];
    const float* d = &rows[(size_t)reflect101(y + 1, h) * w];
    const float* r = &rows[(size_t)y * w];
    for (int x = 0; x < w; ++x) {
      const float side = u[x] + d[x];
      const float c = r[x] * 0.5f;
      const float s = side * 0.25f;
      m[(size_t)y * w + x] = c + s;
    }
  }
}

// GainCompensator::singleFeed's system. A is symmetric positive definite
// (beta * sum N on the diagonal plus a weighted graph Laplacian-like term)
// with a handful of entries per row, so it is kept sparse and solved by
// Jacobi-preconditioned conjugate gradients in double down to a relative
// residual of 1e-15: the gains agree with cv::solve's LU to ~1e-13, far
// below the float32 the maps are stored in.
void GainCompensator::solve(const std::vector<GainStat>& stats) {
  const double alpha = 0.01, beta = 100.0;
  std::vector<char> skip(n_blocks, 1);
  for (size_t p = 0; p < pairs.size(); ++p)
    if (stats[p].count > 0 && pairs[p].i != pairs[p].j)
      skip[pairs[p].i] = skip[pairs[p].j] = 0;
  std::vector<int> pos(n_blocks, -1);
  int ne = 0;
  for (int i = 0; i < n_blocks; ++i)
    if (!skip[i]) pos[i] = ne++;
  n_unknowns = ne;
  block_gains.assign(n_blocks, 1.0);
  cg_iterations = 0;
  if (ne == 0) return;

  std::vector<double> diag(ne, 0.0), rhs(ne, 0.0);
  struct Off {
    int r, c;
    double v;
  };
  std::vector<Off> off;
  off.reserve(pairs.size());
  for (size_t p = 0; p < pairs.size(); ++p) {
    const int i = pairs[p].i, j = pairs[p].j;
    if (skip[i] || skip[j]) continue;
    const long long cnt = stats[p].count;
    const double N = (double)std::max<long long>(1, cnt);
    const double Iij = cnt > 0 ? stats[p].s1 / (double)cnt : 0.0;
    const double Iji = cnt > 0 ? stats[p].s2 / (double)cnt : 0.0;
    const int ki = pos[i], kj = pos[j];
    if (i == j) {
      rhs[ki] += beta * N;
      diag[ki] += beta * N;
      continue;
    }
    rhs[ki] += beta * N;
    rhs[kj] += beta * N;
    diag[ki] += beta * N + 2 * alpha * Iij * Iij * N;
    diag[kj] += beta * N + 2 * alpha * Iji * Iji * N;
    const double v = -2 * alpha * Iij * Iji * N;
    if (v != 0.0) off.push_back({ki, kj, v});
  }
  auto mul = [&](const std::vector<double>& x, std::vector<double>& y) {
    for (int k = 0; k < ne; ++k) y[k] = diag[k] * x[k];
    for (const Off& o : off) {
      y[o.r] += o.v * x[o.c];
      y[o.c] += o.v * x[o.r];
    }
  };
  std::vector<double> x(ne, 1.0), r(ne), z(ne), d(ne), q(ne);
  mul(x, q);
  double bnorm = 0;
  for (int k = 0; k < ne; ++k) {
    r[k] = rhs[k] - q[k];
    bnorm += rhs[k] * rhs[k];
  }
  double rz = 0;
  for (int k = 0; k < ne; ++k) {
    z[k] = r[k] / diag[k];
    d[k] = z[k];
    rz += r[k] * z[k];
  }
  // stop at a relative residual of 1e-15, or when rounding has taken over
  // (the residual has not halved in 25 iterations)
  const double tol2 = 1e-30 * bnorm;
  double best = 1e300;
  int since_best = 0;
  for (int it = 0; it < 20 * ne + 100; ++it) {
    double rr = 0;
    for (int k = 0; k < ne; ++k) rr += r[k] * r[k];
    if (rr <= tol2) break;
    if (rr < 0.25 * best) {
      best = rr;
      since_best = 0;
    } else if (++since_best > 25) {
      break;
    }
    mul(d, q);
    double dq = 0;
    for (int k = 0; k < ne; ++k) dq += d[k] * q[k];
    if (!(dq > 0)) break;
    const double a = rz / dq;
    for (int k = 0; k < ne; ++k) {
      x[k] += a * d[k];
      r[k] -= a * q[k];
    }
    double rz2 = 0;
    for (int k = 0; k < ne; ++k) {
      z[k] = r[k] / diag[k];
      rz2 += r[k] * z[k];
    }
    const double bt = rz2 / rz;
    rz = rz2;
    for (int k = 0; k < ne; ++k) d[k] = z[k] + bt * d[k];
    ++cg_iterations;
  }
  for (int i = 0; i < n_blocks; ++i)
    if (!skip[i]) block_gains[i] = x[pos[i]];
}

void GainCompensator::plan(int n, const int* corners, const int* widths,
                           const int* heights) {
  XP_REQUIRE(n >= 0 && (n == 0 || (corners && widths && heights)),
             "gain feed: null argument");
  XP_REQUIRE(block_w > 0 && block_h > 0, "gain feed: bad block size");
  im.assign(n, GainImage());
  n_blocks = 0;
  for (int k = 0; k < n; ++k) {
    GainImage& g = im[k];
    g.cx = corners[2 * k];
    g.cy = corners[2 * k + 1];
    g.w = widths[k];
    g.h = heights[k];
    XP_REQUIRE(g.w > 0 && g.h > 0, "gain feed: empty image");
    g.nx = div_up(g.w, block_w);
    g.ny = div_up(g.h, block_h);
    g.bw = div_up(g.w, g.nx);
    g.bh = div_up(g.h, g.ny);
    g.first = n_blocks;
    g.bgr = g.mask = nullptr;
    g.stride = g.mstride = 0;
    n_blocks += g.nx * g.ny;
  }
  // every pair i <= j of blocks whose rectangles overlap. Blocks of one
  // image are disjoint, so within an image only (i, i) occurs; across images
  // the candidate range of the second image's blocks follows from division.
  pairs.clear();
  for (int a = 0; a < n; ++a) {
    const GainImage& A = im[a];
    for (int by = 0; by < A.ny; ++by)
      for (int bx = 0; bx < A.nx; ++bx) {
        const int i = A.first + by * A.nx + bx;
        const int ax0 = A.cx + bx * A.bw, ay0 = A.cy + by * A.bh;
        const int ax1 = std::min(ax0 + A.bw, A.cx + A.w);
        const int ay1 = std::min(ay0 + A.bh, A.cy + A.h);
        pairs.push_back({a, a, i, i, ax0, ay0, ax1 - ax0, ay1 - ay0});
        for (int b = a + 1; b < n; ++b) {
          const GainImage& B = im[b];
          const int x0 = std::max(ax0, B.cx), y0 = std::max(ay0, B.cy);
          const int x1 = std::min(ax1, B.cx + B.w);
          const int y1 = std::min(ay1, B.cy + B.h);
          if (x0 >= x1 || y0 >= y1) continue;
          const int jx0 = (x0 - B.cx) / B.bw, jx1 = (x1 - 1 - B.cx) / B.bw;
          const int jy0 = (y0 - B.cy) / B.bh, jy1 = (y1 - 1 - B.cy) / B.bh;
          for (int jy = jy0; jy <= jy1; ++jy)
            for (int jx = jx0; jx <= jx1; ++jx) {
              const int bx0 = B.cx + jx * B.bw, by0 = B.cy + jy * B.bh;
              const int bx1 = std::min(bx0 + B.bw, B.cx + B.w);
              const int by1 = std::min(by0 + B.bh, B.cy + B.h);
              const int ix0 = std::max(x0, bx0), iy0 = std::max(y0, by0);
              const int ix1 = std::min(x1, bx1), iy1 = std::min(y1, by1);
              if (ix0 >= ix1 || iy0 >= iy1) continue;
              pairs.push_back({a, b, i, B.first + jy * B.nx + jx, ix0, iy0,
                               ix1 - ix0, iy1 - iy0});
            }
        }
      }
  }
  n_pairs = (int)pairs.size();
}

void GainCompensator::finish(const std::vector<GainStat>& stats) {
  XP_REQUIRE(stats.size() == pairs.size(),
             "gain solve: one statistic per planned pair expected");
  solve(stats);
  const int n = (int)im.size();
  maps.assign(n, {});
  map_size.assign(n, make_int2(0, 0));
  for (int k = 0; k < n; ++k) {
    const GainImage& g = im[k];
    std::vector<float>& m = maps[k];
    m.resize((size_t)g.nx * g.ny);
    for (int b = 0; b < g.nx * g.ny; ++b)
      m[b] = (float)block_gains[g.first + b];
    for (int it = 0; it < filter_iters; ++it) smooth_map(m, g.nx, g.ny);
    map_size[k] = make_int2(g.nx, g.ny);
  }
}

void GainCompensator::feed(int n, const int* corners, const void* const* bgr,
                           const int* widths, const int* heights,
                           const size_t* strides,
                           const uint8_t* const* masks,
                           const size_t* mask_strides, int mem) {
  XP_REQUIRE(n == 0 || (bgr && masks), "gain feed: null argument");
  plan(n, corners, widths, heights);
  select_device(device);
  cudaStream_t s = nullptr;

  std::vector<DevBuf> staging(2 * (size_t)n);
  for (int k = 0; k < n; ++k) {
    GainImage& g = im[k];
    XP_REQUIRE(bgr[k] && masks[k], "gain feed: null image or mask");
    size_t st = strides && strides[k] ? strides[k] : (size_t)g.w * 3;
    size_t ms = mask_strides && mask_strides[k] ? mask_strides[k]
                                                : (size_t)g.w;
    XP_REQUIRE(st >= (size_t)g.w * 3 && ms >= (size_t)g.w,
               "gain feed: stride smaller than a row");
    if (mem == XP_MEM_HOST) {
      DevBuf& di = staging[2 * k];
      DevBuf& dm = staging[2 * k + 1];
      di.reserve((size_t)g.w * 3 * g.h);
      dm.reserve((size_t)g.w * g.h);
      copy_in_2d(di.p, (size_t)g.w * 3, bgr[k], st, (size_t)g.w * 3,
                 (size_t)g.h, mem, s);
      copy_in_2d(dm.p, (size_t)g.w, masks[k], ms, (size_t)g.w, (size_t)g.h,
                 mem, s);
      g.bgr = di.as<uint8_t>();
      g.stride = (size_t)g.w * 3;
      g.mask = dm.as<uint8_t>();
      g.mstride = (size_t)g.w;
    } else {
      g.bgr = static_cast<const uint8_t*>(bgr[k]);
      g.stride = st;
      g.mask = masks[k];
      g.mstride = ms;
    }
    XP_REQUIRE(g.stride < (1ull << 32) && g.mstride < (1ull << 32),
               "gain feed: stride too large");
  }

  std::vector<GainStat> stats(pairs.size());
  if (!pairs.empty()) {
    std::vector<GainPairDev> hp(pairs.size());
    for (size_t p = 0; p < pairs.size(); ++p) {
      const GainPair& q = pairs[p];
      const GainImage& A = im[q.a];
      const GainImage& B = im[q.b];
      GainPairDev& d = hp[p];
      d.pa = A.bgr + (size_t)(q.y0 - A.cy) * A.stride + 3 * (q.x0 - A.cx);
      d.pb = B.bgr + (size_t)(q.y0 - B.cy) * B.stride + 3 * (q.x0 - B.cx);
      d.ma = A.mask + (size_t)(q.y0 - A.cy) * A.mstride + (q.x0 - A.cx);
      d.mb = B.mask + (size_t)(q.y0 - B.cy) * B.mstride + (q.x0 - B.cx);
      d.sa = (unsigned)A.stride;
      d.sb = (unsigned)B.stride;
      d.sma = (unsigned)A.mstride;
      d.smb = (unsigned)B.mstride;
      d.w = q.w;
      d.h = q.h;
    }
    DevBuf dp, ds;
    dp.reserve(sizeof(GainPairDev) * hp.size());
    ds.reserve(sizeof(GainStat) * hp.size());
    XP_CUDA(cudaMemcpyAsync(dp.p, hp.data(), sizeof(GainPairDev) * hp.size(),
                            cudaMemcpyHostToDevice, s));
    gain_pairs_kernel<<<(unsigned)hp.size(), 128, 0, s>>>(
        dp.as<GainPairDev>(), ds.as<GainStat>());
    XP_LAUNCHED();
    XP_CUDA(cudaMemcpyAsync(stats.data(), ds.p, sizeof(GainStat) * hp.size(),
                            cudaMemcpyDeviceToHost, s));
  }
  XP_CUDA(cudaStreamSynchronize(s));
  for (GainImage& g : im) g.bgr = g.mask = nullptr;  // the caller's buffers
  finish(stats);
}

void GainCompensator::apply(int index, void* bgr, int w, int h, size_t stride,
                            int depth, int mem) const {
  XP_REQUIRE(index >= 0 && index < (int)maps.size(),
             "gain apply: index out of range (feed first)");
  XP_REQUIRE(bgr && w > 0 && h > 0, "gain apply: empty image");
  XP_REQUIRE(depth == 8 || depth == 16, "gain apply: depth must be 8 or 16");
  const size_t bpe = depth / 8;
  if (stride == 0) stride = (size_t)w * 3 * bpe;
  XP_REQUIRE(stride >= (size_t)w * 3 * bpe && stride % bpe == 0,
             "gain apply: bad stride");
  select_device(device);
  cudaStream_t s = nullptr;
  const int gw = map_size[index].x, gh = map_size[index].y;
  DevBuf dg, staging;
  dg.reserve(sizeof(float) * (size_t)gw * gh);
  XP_CUDA(cudaMemcpyAsync(dg.p, maps[index].data(),
                          sizeof(float) * (size_t)gw * gh,
                          cudaMemcpyHostToDevice, s));
  void* d = bgr;
  size_t dstride = stride;
  if (mem == XP_MEM_HOST) {
    dstride = (size_t)w * 3 * bpe;
    staging.reserve(dstride * h);
    copy_in_2d(staging.p, dstride, bgr, stride, dstride, (size_t)h, mem, s);
    d = staging.p;
  }
  dim3 blk(32, 8), grd(div_up(w, 32), div_up(h, 8));
  if (depth == 8)
    gain_apply_kernel<uint8_t><<<grd, blk, 0, s>>>(
        static_cast<uint8_t*>(d), w, h, dstride, dg.as<float>(), gw, gh);
  else
    gain_apply_kernel<uint16_t><<<grd, blk, 0, s>>>(
        static_cast<uint16_t*>(d), w, h, dstride / 2, dg.as<float>(), gw, gh);
  XP_LAUNCHED();
  if (mem == XP_MEM_HOST)
    copy_out_2d(bgr, stride, staging.p, dstride, dstride, (size_t)h, mem, s);
  XP_CUDA(cudaStreamSynchronize(s));
}

}  // namespace xp

extern "C" {

xp_status xp_gain_compensator_create(int block_w, int block_h,
                                     int filter_iterations, int device,
                                     xp_gain_compensator** out) {
  return xp::guard([&] {
    XP_REQUIRE(out, "gain_compensator_create: null argument");
    *out = nullptr;
    XP_REQUIRE(block_w > 0 && block_h > 0 && filter_iterations >= 0,
               "gain_compensator_create: bad parameters");
    // no device call here: plan / solve are host-only; feed / apply select
    auto* g = new xp::GainCompensator();
    g->block_w = block_w;
    g->block_h = block_h;
    g->filter_iters = filter_iterations;
    g->device = device;
    *out = reinterpret_cast<xp_gain_compensator*>(g);
  });
}

void xp_gain_compensator_destroy(xp_gain_compensator* g) {
  delete reinterpret_cast<xp::GainCompensator*>(g);
}

xp_status xp_gain_compensator_feed(xp_gain_compensator* g, int n,
                                   const int* corners_xy,
                                   const void* const* bgr, const int* widths,
                                   const int* heights, const size_t* strides,
                                   const uint8_t* const* masks,
                                   const size_t* mask_strides, int mem) {
  return xp::guard([&] {
    XP_REQUIRE(g, "gain_compensator_feed: null handle");
    reinterpret_cast<xp::GainCompensator*>(g)->feed(
        n, corners_xy, bgr, widths, heights, strides, masks, mask_strides,
        mem);
  });
}

xp_status xp_gain_compensator_plan(xp_gain_compensator* g, int n,
                                   const int* corners_xy, const int* widths,
                                   const int* heights, int* n_pairs) {
  return xp::guard([&] {
    XP_REQUIRE(g && n_pairs, "gain_compensator_plan: null argument");
    auto* c = reinterpret_cast<xp::GainCompensator*>(g);
    c->plan(n, corners_xy, widths, heights);
    *n_pairs = c->n_pairs;
  });
}

xp_status xp_gain_compensator_pairs(const xp_gain_compensator* g,
                                    int* pairs6) {
  return xp::guard([&] {
    XP_REQUIRE(g && pairs6, "gain_compensator_pairs: null argument");
    auto* c = reinterpret_cast<const xp::GainCompensator*>(g);
    for (size_t p = 0; p < c->pairs.size(); ++p) {
      const xp::GainPair& q = c->pairs[p];
      int* o = pairs6 + 6 * p;
      o[0] = q.i;
      o[1] = q.j;
      o[2] = q.x0;
      o[3] = q.y0;
      o[4] = q.w;
      o[5] = q.h;
    }
  });
}

xp_status xp_gain_compensator_solve(xp_gain_compensator* g, int n_pairs,
                                    const long long* counts,
                                    const double* sums_first,
                                    const double* sums_second) {
  return xp::guard([&] {
    XP_REQUIRE(g && (n_pairs == 0 || (counts && sums_first && sums_second)),
               "gain_compensator_solve: null argument");
    auto* c = reinterpret_cast<xp::GainCompensator*>(g);
    XP_REQUIRE(n_pairs == (int)c->pairs.size(),
               "gain_compensator_solve: pair count differs from the plan");
    std::vector<xp::GainStat> st(n_pairs);
    for (int p = 0; p < n_pairs; ++p)
      st[p] = {sums_first[p], sums_second[p], counts[p]};
    c->finish(st);
  });
}

xp_status xp_gain_compensator_map_size(const xp_gain_compensator* g,
                                       int index, int wh[2]) {
  return xp::guard([&] {
    XP_REQUIRE(g && wh, "gain_compensator_map_size: null argument");
    auto* c = reinterpret_cast<const xp::GainCompensator*>(g);
    XP_REQUIRE(index >= 0 && index < (int)c->maps.size(),
               "gain_compensator_map_size: index out of range (feed first)");
    wh[0] = c->map_size[index].x;
    wh[1] = c->map_size[index].y;
  });
}

xp_status xp_gain_compensator_get_map(const xp_gain_compensator* g, int index,
                                      float* dst) {
  return xp::guard([&] {
    XP_REQUIRE(g && dst, "gain_compensator_get_map: null argument");
    auto* c = reinterpret_cast<const xp::GainCompensator*>(g);
    XP_REQUIRE(index >= 0 && index < (int)c->maps.size(),
               "gain_compensator_get_map: index out of range (feed first)");
    std::copy(c->maps[index].begin(), c->maps[index].end(), dst);
  });
}

xp_status xp_gain_compensator_set_map(xp_gain_compensator* g, int index,
                                      const float* src, int w, int h) {
  return xp::guard([&] {
    XP_REQUIRE(g && src && w > 0 && h > 0 && index >= 0,
               "gain_compensator_set_map: bad argument");
    auto* c = reinterpret_cast<xp::GainCompensator*>(g);
    if (index >= (int)c->maps.size()) {
      c->maps.resize(index + 1);
      c->map_size.resize(index + 1, make_int2(0, 0));
    }
    c->maps[index].assign(src, src + (size_t)w * h);
    c->map_size[index] = make_int2(w, h);
  });
}

xp_status xp_gain_compensator_apply(const xp_gain_compensator* g, int index,
                                    void* bgr, int w, int h, size_t stride,
                                    int depth, int mem) {
  return xp::guard([&] {
    XP_REQUIRE(g, "gain_compensator_apply: null handle");
    reinterpret_cast<const xp::GainCompensator*>(g)->apply(index, bgr, w, h,
                                                           stride, depth, mem);
  });
}

xp_status xp_gain_compensator_stats(const xp_gain_compensator* g,
                                    int out[3]) {
  return xp::guard([&] {
    XP_REQUIRE(g && out, "gain_compensator_stats: null argument");
    auto* c = reinterpret_cast<const xp::GainCompensator*>(g);
    out[0] = c->n_pairs;
    out[1] = c->n_unknowns;
    out[2] = c->cg_iterations;
  });
}

}  // extern "C"
