// Warp stage of the export hot path as sm_100a kernels: per-pixel inverse map
// + OpenCV's fixed-point bilinear gather, fused with the mask warp, exposure
// gain, seam-mask upsample and BGRA pack.
//
// Replaces: cv::detail::RotationWarper::{warpRoi,buildMaps,warp,warpPoint,
// warpPointBackward} as called from xpano/algorithm/stitcher.cc:160,377-385,
// ExposureCompensator::apply (:389-390), dilate/resize/bitwise_and (:397-401)
// and blenders.cc:99-112 (Merge). Arithmetic models: SURVEY.md section 6A.B/C
// (remap = 1/32-px fixed point, 15-bit weights for 8-bit, float weights for
// 16-bit, BORDER_REFLECT; INTER_NEAREST mask = rint in-bounds test;
// resize INTER_LINEAR_EXACT = 8.8 fixed point).
#include <climits>

#include "common.h"
#include "projector.h"
#include <mutex>

#include "warp.h"

namespace xp {

// ---------------------------------------------------------------------------
// host: ProjectorBase::setCameraParams
// ---------------------------------------------------------------------------
static void mat3_mul(const float* a, const float* b, float* out) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      float s = 0;
      for (int k = 0; k < 3; ++k) s = FADD(s, FMUL(a[i * 3 + k], b[k * 3 + j]));
      out[i * 3 + j] = s;
    }
}

static void mat3_inv(const float* m, float* out) {
  // cv::Mat::inv() for 3x3: cofactors in double, one rounding to float
  double d = (double)m[0] * ((double)m[4] * m[8] - (double)m[5] * m[7]) -
             (double)m[1] * ((double)m[3] * m[8] - (double)m[5] * m[6]) +
             (double)m[2] * ((double)m[3] * m[7] - (double)m[4] * m[6]);
  if (d == 0) {
    memset(out, 0, sizeof(float) * 9);
    return;
  }
  d = 1. / d;
  out[0] = (float)(((double)m[4] * m[8] - (double)m[5] * m[7]) * d);
  out[1] = (float)(((double)m[2] * m[7] - (double)m[1] * m[8]) * d);
  out[2] = (float)(((double)m[1] * m[5] - (double)m[2] * m[4]) * d);
  out[3] = (float)(((double)m[5] * m[6] - (double)m[3] * m[8]) * d);
  out[4] = (float)(((double)m[0] * m[8] - (double)m[2] * m[6]) * d);
  out[5] = (float)(((double)m[2] * m[3] - (double)m[0] * m[5]) * d);
  out[6] = (float)(((double)m[3] * m[7] - (double)m[4] * m[6]) * d);
  out[7] = (float)(((double)m[1] * m[6] - (double)m[0] * m[7]) * d);
  out[8] = (float)(((double)m[0] * m[4] - (double)m[1] * m[3]) * d);
}

Projector Warper::projector(const float* K, const float* R) const {
  Projector p;
  float kinv[9];
  p.type = type;
  p.portrait = portrait;
  p.scale = scale;
  p.a = a;
  p.b = b;
  memcpy(p.k, K, sizeof(float) * 9);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) p.rinv[i * 3 + j] = R[j * 3 + i];
  mat3_inv(K, kinv);
  mat3_mul(R, kinv, p.r_kinv);
  mat3_mul(K, p.rinv, p.k_rinv);
  p.t[0] = p.t[1] = p.t[2] = 0.f;
  return p;
}

// ---------------------------------------------------------------------------
// ROI detection (RotationWarperBase::detectResultRoi[ByBorder] and the
// Plane/Spherical overrides)
// ---------------------------------------------------------------------------
struct Ext {
  float tl_u = 3.402823466e+38f, tl_v = 3.402823466e+38f;
  float br_u = -3.402823466e+38f, br_v = -3.402823466e+38f;
  void add(float u, float v) {
    tl_u = u < tl_u ? u : tl_u;
    tl_v = v < tl_v ? v : tl_v;
    br_u = br_u < u ? u : br_u;
    br_v = br_v < v ? v : br_v;
  }
};

// order-preserving float <-> uint key for atomicMin/Max
__device__ __forceinline__ unsigned fkey(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
static float unkey(unsigned k) {
  unsigned u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

// generic detectResultRoi: forward-map every source pixel, min/max of (u, v)
__global__ void roi_all_pixels_kernel(Projector p, int w, int h,
                                      unsigned* out4) {
  float tl_u = 3.402823466e+38f, tl_v = tl_u, br_u = -tl_u, br_v = -tl_u;
  const size_t n = (size_t)w * h;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x) {
    float u, v;
    map_forward(p, (float)(i % w), (float)(i / w), u, v);
    tl_u = u < tl_u ? u : tl_u;
    tl_v = v < tl_v ? v : tl_v;
    br_u = br_u < u ? u : br_u;
    br_v = br_v < v ? v : br_v;
  }
  for (int o = 16; o > 0; o >>= 1) {
    tl_u = fminf(tl_u, __shfl_xor_sync(~0u, tl_u, o));
    tl_v = fminf(tl_v, __shfl_xor_sync(~0u, tl_v, o));
    br_u = fmaxf(br_u, __shfl_xor_sync(~0u, br_u, o));
    br_v = fmaxf(br_v, __shfl_xor_sync(~0u, br_v, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(out4 + 0, fkey(tl_u));
    atomicMin(out4 + 1, fkey(tl_v));
    atomicMax(out4 + 2, fkey(br_u));
    atomicMax(out4 + 3, fkey(br_v));
  }
}

// part `part` of `parts` of the border walk (pure host math, thread-safe);
// min/max merging of the parts is exact
void Warper::border_extent(const Projector& p, int w, int h, int part,
                           int parts, float ext[4]) {
  Ext e;
  float u, v;
  const int n = w + h;
  const int beg = (int)((long)n * part / parts),
            end = (int)((long)n * (part + 1) / parts);
  for (int k = beg; k < end; ++k) {
    if (k < w) {
      map_forward(p, (float)k, 0, u, v);
      e.add(u, v);
      map_forward(p, (float)k, (float)(h - 1), u, v);
      e.add(u, v);
    } else {
      const int i = k - w;
      map_forward(p, 0, (float)i, u, v);
      e.add(u, v);
      map_forward(p, (float)(w - 1), (float)i, u, v);
      e.add(u, v);
    }
  }
  ext[0] = e.tl_u, ext[1] = e.tl_v, ext[2] = e.br_u, ext[3] = e.br_v;
}

void Warper::detect_roi(const Projector& p, int w, int h, int out[4],
                        const float* border_ext) {
  Ext e;
  float u, v;
  const bool by_border =
      (p.type == XP_PROJ_CYLINDRICAL || p.type == XP_PROJ_SPHERICAL);
  if (p.type == XP_PROJ_PERSPECTIVE && !p.portrait) {
    map_forward(p, 0, 0, u, v);
    e.add(u, v);
    map_forward(p, 0, (float)(h - 1), u, v);
    e.add(u, v);
    map_forward(p, (float)(w - 1), 0, u, v);
    e.add(u, v);
    map_forward(p, (float)(w - 1), (float)(h - 1), u, v);
    e.add(u, v);
  } else if (by_border) {
    float ext[4];
    if (!border_ext) {
      border_extent(p, w, h, 0, 1, ext);
      border_ext = ext;
    }
    e.tl_u = border_ext[0], e.tl_v = border_ext[1];
    e.br_u = border_ext[2], e.br_v = border_ext[3];
  } else {
    // O(area) forward map: data-parallel, done on the device
    ensure_device();
    scratch.reserve(16);
    unsigned init[4] = {0xffffffffu, 0xffffffffu, 0u, 0u};
    XP_CUDA(cudaMemcpyAsync(scratch.p, init, 16, cudaMemcpyHostToDevice,
                            stream));
    int blocks = (int)std::min<size_t>(div_up((size_t)w * h, (size_t)256),
                                       sm_count() * 16);
    roi_all_pixels_kernel<<<blocks, 256, 0, stream>>>(
        p, w, h, scratch.as<unsigned>());
    XP_LAUNCHED();
    unsigned r[4];
    XP_CUDA(cudaMemcpyAsync(r, scratch.p, 16, cudaMemcpyDeviceToHost, stream));
    XP_CUDA(cudaStreamSynchronize(stream));
    e.tl_u = unkey(r[0]);
    e.tl_v = unkey(r[1]);
    e.br_u = unkey(r[2]);
    e.br_v = unkey(r[3]);
  }
  int tlx = (int)e.tl_u, tly = (int)e.tl_v, brx = (int)e.br_u,
      bry = (int)e.br_v;
  if (p.type == XP_PROJ_SPHERICAL) {
    // SphericalWarper / SphericalPortraitWarper::detectResultRoi pole tests
    float tl_uf = (float)tlx, tl_vf = (float)tly, br_uf = (float)brx,
          br_vf = (float)bry;
    const int c = p.portrait ? 0 : 1;
    for (int pole = 0; pole < 2; ++pole) {
      float x = p.rinv[c];
      float y = pole == 0 ? p.rinv[3 + c] : -p.rinv[3 + c];
      float z = p.rinv[6 + c];
      if (y > 0.f) {
        float x_ = FADD(FDIV(FADD(FMUL(p.k[0], x), FMUL(p.k[1], y)), z), p.k[2]);
        float y_ = FADD(FDIV(FMUL(p.k[4], y), z), p.k[5]);
        if (x_ > 0.f && x_ < w && y_ > 0.f && y_ < h) {
          float pv = pole == 0 ? (float)(XP_PI_D * p.scale) : 0.f;
          tl_uf = tl_uf < 0.f ? tl_uf : 0.f;
          tl_vf = tl_vf < pv ? tl_vf : pv;
          br_uf = br_uf > 0.f ? br_uf : 0.f;
          br_vf = br_vf > pv ? br_vf : pv;
        }
      }
    }
    tlx = (int)tl_uf;
    tly = (int)tl_vf;
    brx = (int)br_uf;
    bry = (int)br_vf;
  }
  out[0] = tlx;
  out[1] = tly;
  out[2] = brx;
  out[3] = bry;
}

// ---------------------------------------------------------------------------
// device: remap arithmetic
// ---------------------------------------------------------------------------
__device__ __forceinline__ int reflect_idx(int p, int len) {
  if ((unsigned)p < (unsigned)len) return p;
  if (len == 1) return 0;
  do {
    if (p < 0)
      p = -p - 1;
    else
      p = len - 1 - (p - len);
  } while ((unsigned)p >= (unsigned)len);
  return p;
}

__device__ __forceinline__ int sat_short(int v) {
  return v < -32768 ? -32768 : (v > 32767 ? 32767 : v);
}

// cvRound(x * 32): out-of-range and NaN give INT_MIN (cvtps_epi32)
__device__ __forceinline__ int round_q5(float x) {
  float t = __fmul_rn(x, 32.0f);
  if (!(t > -2147483648.0f && t < 2147483648.0f)) return INT_MIN;
  return __float2int_rn(t);
}
__device__ __forceinline__ int round_q0(float x) {
  if (!(x > -2147483648.0f && x < 2147483648.0f)) return INT_MIN;
  return __float2int_rn(x);
}

template <typename T>
struct Px3 {
  T c[3];
};

// INTER_LINEAR + BORDER_REFLECT for CN interleaved channels of T
template <typename T, int CN>
__device__ __forceinline__ void sample_linear_reflect(
    const T* __restrict__ src, size_t pitch_elems, int sw, int sh, float mx,
    float my, T* out) {
  const int sx = round_q5(mx), sy = round_q5(my);
  const int fx = sx & 31, fy = sy & 31;
  const int ix = sat_short(sx >> 5), iy = sat_short(sy >> 5);
  if (sizeof(T) == 1 && (unsigned)ix < (unsigned)(sw - 1) &&
      (unsigned)iy < (unsigned)(sh - 1)) {
    // interior (nearly every pixel): the four taps are two runs of 2 * CN
    // consecutive bytes, one base pointer, no border arithmetic
    const T* p0 = src + (size_t)iy * pitch_elems + ix * CN;
    const T* p1 = p0 + pitch_elems;
    const int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32,
              w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
#pragma unroll
    for (int c = 0; c < CN; ++c) {
      int p00 = __ldg(p0 + c), p01 = __ldg(p0 + CN + c), p10 = __ldg(p1 + c),
          p11 = __ldg(p1 + CN + c);
      int v = (p00 * w00 + p01 * w01 + p10 * w10 + p11 * w11 + (1 << 14)) >> 15;
      out[c] = (T)(v > 255 ? 255 : v);
    }
    return;
  }
  // border pixels: BORDER_REFLECT indices; interior 16-bit pixels take the
  // same path with x1 = x0 + 1, y1 = y0 + 1 and no reflection arithmetic
  const bool inner = (unsigned)ix < (unsigned)(sw - 1) &&
                     (unsigned)iy < (unsigned)(sh - 1);
  const int x0 = inner ? ix : reflect_idx(ix, sw),
            x1 = inner ? ix + 1 : reflect_idx(ix + 1, sw);
  const int y0 = inner ? iy : reflect_idx(iy, sh),
            y1 = inner ? iy + 1 : reflect_idx(iy + 1, sh);
  const T* r0 = src + (size_t)y0 * pitch_elems;
  const T* r1 = src + (size_t)y1 * pitch_elems;
  if (sizeof(T) == 1) {
    const int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32,
              w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
#pragma unroll
    for (int c = 0; c < CN; ++c) {
      int p00 = __ldg(r0 + x0 * CN + c), p01 = __ldg(r0 + x1 * CN + c),
          p10 = __ldg(r1 + x0 * CN + c), p11 = __ldg(r1 + x1 * CN + c);
      int v = (p00 * w00 + p01 * w01 + p10 * w10 + p11 * w11 + (1 << 14)) >> 15;
      out[c] = (T)(v > 255 ? 255 : v);
    }
  } else {
    const float ax1 = __fmul_rn((float)fx, 1.f / 32),
                ax0 = __fsub_rn(1.f, ax1);
    const float ay1 = __fmul_rn((float)fy, 1.f / 32),
                ay0 = __fsub_rn(1.f, ay1);
    const float w00 = __fmul_rn(ay0, ax0), w01 = __fmul_rn(ay0, ax1),
                w10 = __fmul_rn(ay1, ax0), w11 = __fmul_rn(ay1, ax1);
#pragma unroll
    for (int c = 0; c < CN; ++c) {
      float p00 = (float)__ldg(r0 + x0 * CN + c),
            p01 = (float)__ldg(r0 + x1 * CN + c),
            p10 = (float)__ldg(r1 + x0 * CN + c),
            p11 = (float)__ldg(r1 + x1 * CN + c);
      float v = __fadd_rn(
          __fadd_rn(__fadd_rn(__fmul_rn(p00, w00), __fmul_rn(p01, w01)),
                    __fmul_rn(p10, w10)),
          __fmul_rn(p11, w11));
      int q = __float2int_rn(v);
      out[c] = (T)(q < 0 ? 0 : (q > 65535 ? 65535 : q));
    }
  }
}

// INTER_NEAREST + BORDER_CONSTANT(0) in-bounds test
__device__ __forceinline__ bool nearest_inside(int sw, int sh, float mx,
                                               float my, int& ix, int& iy) {
  ix = sat_short(round_q0(mx));
  iy = sat_short(round_q0(my));
  return ix >= 0 && ix < sw && iy >= 0 && iy < sh;
}

// ---------------------------------------------------------------------------
// inverse map evaluation: separable tables or the generic projector
// ---------------------------------------------------------------------------
struct MapSrc {
  Projector p;
  const float2* col;  // per dst column: (S, C)   (separable only)
  const float2* row;  // per dst row: (A, B)
  int tlx, tly;
  int separable;
};

__global__ void sep_tables_kernel(Projector p, int tlx, int tly, int W, int H,
                                  float2* col, float2* row) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < W) {
    float S, C;
    sep_col(p, (float)(i + tlx), S, C);
    col[i] = make_float2(S, C);
  } else if (i < W + H) {
    int j = i - W;
    float A, B;
    sep_row(p, (float)(j + tly), A, B);
    row[j] = make_float2(A, B);
  }
}

// tables evaluated on the host (see Warper::prepare_map): pinned, mapped
// memory -> device, without a copy engine
__global__ void tables_from_host_kernel(const float2* __restrict__ src,
                                        float2* __restrict__ dst, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[i];
}

__device__ __forceinline__ void eval_map(const MapSrc& m, int dx, int dy,
                                         float& x, float& y) {
  if (m.separable) {
    float2 c = __ldg(m.col + dx), r = __ldg(m.row + dy);
    sep_combine(m.p, r.x, r.y, c.x, c.y, x, y);
  } else {
    map_backward(m.p, (float)(dx + m.tlx), (float)(dy + m.tly), x, y);
  }
}

__global__ void build_maps_kernel(MapSrc m, int W, int H, float* xmap,
                                  float* ymap, size_t stride) {
  int dx = blockIdx.x * blockDim.x + threadIdx.x;
  int dy = blockIdx.y * blockDim.y + threadIdx.y;
  if (dx >= W || dy >= H) return;
  float x, y;
  eval_map(m, dx, dy, x, y);
  xmap[(size_t)dy * stride + dx] = x;
  ymap[(size_t)dy * stride + dx] = y;
}

template <typename T, int CN>
__global__ void warp_linear_kernel(MapSrc m, const T* __restrict__ src,
                                   size_t src_pitch, int sw, int sh, int W,
                                   int H, T* __restrict__ dst,
                                   size_t dst_pitch) {
  int dx = blockIdx.x * blockDim.x + threadIdx.x;
  int dy = blockIdx.y * blockDim.y + threadIdx.y;
  if (dx >= W || dy >= H) return;
  float x, y;
  eval_map(m, dx, dy, x, y);
  T out[CN];
  sample_linear_reflect<T, CN>(src, src_pitch, sw, sh, x, y, out);
  T* d = dst + (size_t)dy * dst_pitch + (size_t)dx * CN;
#pragma unroll
  for (int c = 0; c < CN; ++c) d[c] = out[c];
}

template <typename T, int CN>
__global__ void warp_nearest_kernel(MapSrc m, const T* __restrict__ src,
                                    size_t src_pitch, int sw, int sh, int W,
                                    int H, T* __restrict__ dst,
                                    size_t dst_pitch) {
  int dx = blockIdx.x * blockDim.x + threadIdx.x;
  int dy = blockIdx.y * blockDim.y + threadIdx.y;
  if (dx >= W || dy >= H) return;
  float x, y;
  eval_map(m, dx, dy, x, y);
  int ix, iy;
  bool in = nearest_inside(sw, sh, x, y, ix, iy);
  T* d = dst + (size_t)dy * dst_pitch + (size_t)dx * CN;
#pragma unroll
  for (int c = 0; c < CN; ++c)
    d[c] = in ? __ldg(src + (size_t)iy * src_pitch + (size_t)ix * CN + c)
              : (T)0;
}

// ---------------------------------------------------------------------------
// seam-mask / gain helpers
// ---------------------------------------------------------------------------
// dilate(src, dst, Mat()): 3x3 max, outside pixels ignored (stitcher.cc:397)
__global__ void dilate3x3_kernel(const uint8_t* __restrict__ src, int w, int h,
                                 uint8_t* __restrict__ dst) {
  int x = blockIdx.x * blockDim.x + threadIdx.x;
  int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= w || y >= h) return;
  int m = 0;
  for (int dy = -1; dy <= 1; ++dy)
    for (int dx = -1; dx <= 1; ++dx) {
      int yy = y + dy, xx = x + dx;
      if (yy < 0 || yy >= h || xx < 0 || xx >= w) continue;
      m = max(m, (int)src[(size_t)yy * w + xx]);
    }
  dst[(size_t)y * w + x] = (uint8_t)m;
}

// resize_bitExact coefficient for one destination index (8.8 fixed point)
__device__ __forceinline__ void exact_coeff(int srcsize, int dstsize, int d,
                                            int& ofs, int& c0, int& c1) {
  const double scale = 1.0 / ((double)dstsize / srcsize);
  const double fval = scale * ((double)d + 0.5) - 0.5;
  const int ival = (int)floor(fval);
  if (ival >= 0 && srcsize > 1) {
    if (ival < srcsize - 1) {
      ofs = ival;
      c1 = (int)__double2ll_rn((fval - (double)ival) * 256.0);
      c0 = 256 - c1;
    } else {
      ofs = srcsize - 1;
      c0 = 256;
      c1 = 0;
    }
  } else {
    ofs = 0;
    c0 = 256;
    c1 = 0;
  }
}

__device__ __forceinline__ int resize_exact_u8(const uint8_t* __restrict__ src,
                                               int sw, int sh, int dw, int dh,
                                               int x, int y) {
  int ox, cx0, cx1, oy, cy0, cy1;
  exact_coeff(sw, dw, x, ox, cx0, cx1);
  exact_coeff(sh, dh, y, oy, cy0, cy1);
  const uint8_t* r0 = src + (size_t)oy * sw;
  const uint8_t* r1 = src + (size_t)(oy + 1 < sh ? oy + 1 : oy) * sw;
  int ox1 = ox + 1 < sw ? ox + 1 : ox;
  unsigned h0 = (unsigned)(cx0 * r0[ox] + cx1 * r0[ox1]);
  unsigned h1 = (unsigned)(cx0 * r1[ox] + cx1 * r1[ox1]);
  unsigned v = (unsigned)cy0 * h0 + (unsigned)cy1 * h1;
  v = (v + (1u << 15)) >> 16;
  return v > 255 ? 255 : (int)v;
}

// resize INTER_LINEAR float coefficient (resizeGeneric_ HResizeLinear)
__device__ __forceinline__ void linear_coeff_f32(int srcsize, int dstsize,
                                                 int d, int& ofs, float& a1) {
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

__device__ __forceinline__ float resize_linear_f32(
    const float* __restrict__ src, int sw, int sh, int dw, int dh, int x,
    int y) {
  int ox, oy;
  float al1, b1;
  linear_coeff_f32(sw, dw, x, ox, al1);
  linear_coeff_f32(sh, dh, y, oy, b1);
  const float al0 = __fsub_rn(1.f, al1), b0 = __fsub_rn(1.f, b1);
  const int ox1 = ox + 1 < sw ? ox + 1 : ox;
  const int oy1 = oy + 1 < sh ? oy + 1 : oy;
  const float* r0 = src + (size_t)oy * sw;
  const float* r1 = src + (size_t)oy1 * sw;
  float h0 = __fadd_rn(__fmul_rn(r0[ox], al0), __fmul_rn(r0[ox1], al1));
  float h1 = __fadd_rn(__fmul_rn(r1[ox], al0), __fmul_rn(r1[ox1], al1));
  return __fadd_rn(__fmul_rn(h0, b0), __fmul_rn(h1, b1));
}

// Per-column / per-row coefficients of the two resizes of the fused kernel:
// the double-precision coefficient arithmetic of exact_coeff /
// linear_coeff_f32 depends on one coordinate only, so it is evaluated once
// per column and row (W + H times) instead of four times per pixel.
struct alignas(16) AuxCoef {
  int g_ofs;   // gain map: first tap
  float g_a1;  // gain map: weight of the second tap
  int s_ofs;   // seam mask: first tap
  int s_c1;    // seam mask: 8.8 weight of the second tap (c0 = 256 - c1)
};

__global__ void aux_tables_kernel(int W, int H, int gw, int gh, int mw, int mh,
                                  AuxCoef* __restrict__ col,
                                  AuxCoef* __restrict__ row) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= W + H) return;
  const bool is_col = i < W;
  const int d = is_col ? i : i - W;
  const int dsize = is_col ? W : H;
  AuxCoef c = {0, 0.f, 0, 0};
  if (gw > 0) linear_coeff_f32(is_col ? gw : gh, dsize, d, c.g_ofs, c.g_a1);
  if (mw > 0) {
    int c0;
    exact_coeff(is_col ? mw : mh, dsize, d, c.s_ofs, c0, c.s_c1);
  }
  (is_col ? col : row)[d] = c;
}

__device__ __forceinline__ int resize_exact_u8_tab(
    const uint8_t* __restrict__ src, int sw, int sh, const AuxCoef& cx,
    const AuxCoef& cy) {
  const int ox = cx.s_ofs, oy = cy.s_ofs;
  const int cx1 = cx.s_c1, cx0 = 256 - cx1, cy1 = cy.s_c1, cy0 = 256 - cy1;
  const uint8_t* r0 = src + (size_t)oy * sw;
  const uint8_t* r1 = src + (size_t)(oy + 1 < sh ? oy + 1 : oy) * sw;
  const int ox1 = ox + 1 < sw ? ox + 1 : ox;
  const unsigned h0 = (unsigned)(cx0 * r0[ox] + cx1 * r0[ox1]);
  const unsigned h1 = (unsigned)(cx0 * r1[ox] + cx1 * r1[ox1]);
  unsigned v = (unsigned)cy0 * h0 + (unsigned)cy1 * h1;
  v = (v + (1u << 15)) >> 16;
  return v > 255 ? 255 : (int)v;
}

__device__ __forceinline__ float resize_linear_f32_tab(
    const float* __restrict__ src, int sw, int sh, const AuxCoef& cx,
    const AuxCoef& cy) {
  const int ox = cx.g_ofs, oy = cy.g_ofs;
  const float al1 = cx.g_a1, b1 = cy.g_a1;
  const float al0 = __fsub_rn(1.f, al1), b0 = __fsub_rn(1.f, b1);
  const int ox1 = ox + 1 < sw ? ox + 1 : ox;
  const int oy1 = oy + 1 < sh ? oy + 1 : oy;
  const float* r0 = src + (size_t)oy * sw;
  const float* r1 = src + (size_t)oy1 * sw;
  const float h0 = __fadd_rn(__fmul_rn(r0[ox], al0), __fmul_rn(r0[ox1], al1));
  const float h1 = __fadd_rn(__fmul_rn(r1[ox], al0), __fmul_rn(r1[ox1], al1));
  return __fadd_rn(__fmul_rn(h0, b0), __fmul_rn(h1, b1));
}

// Fused stitcher.cc:377-405 + blenders.cc:99-112 for one image.
// kSkipTransparent: pixels that come out with alpha 0 get colour 0 instead of
// their (unused) reflected sample -- the compose driver's form; the public
// xp_warper_warp_bgra keeps the reference's colours everywhere.
// Optional by-product of the composer's warps (null = off): the opacity of
// every pixel as a bit plane, one 32-bit word per 32 consecutive x (bit =
// x & 31). The blender's trim (image.cpp:396-512) then reads one bit per pixel
// instead of a 4- or 8-byte pixel, and so does its cover map.
struct AlphaOut {
  uint32_t* bits;   // H rows of bits_pitch words
  int bits_pitch;
};

// every thread of a warp calls this once, from one place, inactive ones with
// opaque = false; threadIdx.x is the lane (blocks are 32 wide)
__device__ __forceinline__ void alpha_out(const AlphaOut& ao, bool opaque,
                                          int dx, int dy, int W, int H) {
  if (!ao.bits) return;
  const unsigned bal = __ballot_sync(0xffffffffu, opaque);
  if (threadIdx.x == 0 && dy < H && dx < W)
    ao.bits[(size_t)dy * ao.bits_pitch + (dx >> 5)] = bal;
}

template <typename T, bool kSkipTransparent>
__global__ void warp_bgra_kernel(MapSrc m, const T* __restrict__ src,
                                 size_t src_pitch, int sw, int sh, int W, int H,
                                 const float* __restrict__ gain, int gw, int gh,
                                 const uint8_t* __restrict__ seam_dilated,
                                 int mw, int mh,
                                 const AuxCoef* __restrict__ aux_col,
                                 const AuxCoef* __restrict__ aux_row,
                                 T* __restrict__ dst, size_t dst_pitch,
                                 AlphaOut ao) {
  const int dx = blockIdx.x * blockDim.x + threadIdx.x;
  const int dy = blockIdx.y * blockDim.y + threadIdx.y;
  const bool in = dx < W && dy < H;
  float x = 0.f, y = 0.f;
  bool opaque = false;
  AuxCoef acx = {0, 0.f, 0, 0}, acy = {0, 0.f, 0, 0};
  if (in) {
    eval_map(m, dx, dy, x, y);
    int ix, iy;
    opaque = nearest_inside(sw, sh, x, y, ix, iy);
    if (gain || seam_dilated) {
      acx = aux_col[dx];
      acy = aux_row[dy];
    }
    if (seam_dilated && opaque)
      opaque = resize_exact_u8_tab(seam_dilated, mw, mh, acx, acy) != 0;
  }
  // one call site for the whole block (ballot + barrier inside)
  alpha_out(ao, opaque, dx, dy, W, H);
  if (!in) return;
  T* d = dst + (size_t)dy * dst_pitch + (size_t)dx * 4;
  if (kSkipTransparent && !opaque) {
    // The colour of a pixel that is fed with alpha 0 never reaches the
    // panorama: multiblend trims it away or overwrites it with the inpaint
    // fill (image.cpp:396-812) and sums only opaque pixels. With graph-cut
    // seam masks that is 30-45 % of a warp ROI, whose reflected samples the
    // reference computes for nothing.
    if (sizeof(T) == 1)
      *reinterpret_cast<uchar4*>(d) = make_uchar4(0, 0, 0, 0);
    else
      *reinterpret_cast<ushort4*>(d) = make_ushort4(0, 0, 0, 0);
    return;
  }
  T px[3];
  sample_linear_reflect<T, 3>(src, src_pitch, sw, sh, x, y, px);
  if (gain) {
    // multiply(CV_8UC3|CV_16UC3, CV_32F gain): saturate_cast(rint(a * g))
    const float g = resize_linear_f32_tab(gain, gw, gh, acx, acy);
    const int maxv = sizeof(T) == 1 ? 255 : 65535;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      int q = __float2int_rn(__fmul_rn((float)px[c], g));
      px[c] = (T)(q < 0 ? 0 : (q > maxv ? maxv : q));
    }
  }
  const T amax = (T)(sizeof(T) == 1 ? 255 : 65535);
  if (sizeof(T) == 1) {
    *reinterpret_cast<uchar4*>(d) =
        make_uchar4(px[0], px[1], px[2], opaque ? 255 : 0);
  } else {
    *reinterpret_cast<ushort4*>(d) =
        make_ushort4(px[0], px[1], px[2], opaque ? amax : 0);
  }
}

// The alpha channel of warp_bgra_kernel alone (colour samples zero): what a
// rank of a canvas-tiled run needs of the images outside its strip -- their
// opacity enters the cover map and the seams, their colours never reach its
// pixels. No source read, no gain.
template <typename T>
__global__ void warp_alpha_kernel(MapSrc m, int sw, int sh, int W, int H,
                                  const uint8_t* __restrict__ seam_dilated,
                                  int mw, int mh,
                                  const AuxCoef* __restrict__ aux_col,
                                  const AuxCoef* __restrict__ aux_row,
                                  T* __restrict__ dst, size_t dst_pitch,
                                  AlphaOut ao) {
  const int dx = blockIdx.x * blockDim.x + threadIdx.x;
  const int dy = blockIdx.y * blockDim.y + threadIdx.y;
  const bool in = dx < W && dy < H;
  bool opaque = false;
  if (in) {
    float x, y;
    eval_map(m, dx, dy, x, y);
    int ix, iy;
    opaque = nearest_inside(sw, sh, x, y, ix, iy);
    if (seam_dilated && opaque)
      opaque = resize_exact_u8_tab(seam_dilated, mw, mh, aux_col[dx],
                                   aux_row[dy]) != 0;
  }
  alpha_out(ao, opaque, dx, dy, W, H);
  if (!in) return;
  const T amax = (T)(sizeof(T) == 1 ? 255 : 65535);
  T* d = dst + (size_t)dy * dst_pitch + (size_t)dx * 4;
  if (sizeof(T) == 1) {
    *reinterpret_cast<uchar4*>(d) = make_uchar4(0, 0, 0, opaque ? 255 : 0);
  } else {
    *reinterpret_cast<ushort4*>(d) = make_ushort4(0, 0, 0, opaque ? amax : 0);
  }
}

// ---------------------------------------------------------------------------
// Warper methods
// ---------------------------------------------------------------------------
void Warper::ensure_device() {
  select_device(device);
  device_ready = true;
  if (!stream && !external_stream) {
    XP_CUDA(cudaStreamCreateWithFlags(&own_stream, cudaStreamNonBlocking));
    stream = own_stream;
  }
}

Warper::~Warper() {
  // scratch buffers return to the pool: the stream (possibly the caller's
  // legacy default stream, a null handle) must be idle, also after an error
  if (device_ready && !(external_stream && detached)) {
    int cur = 0;
    cudaGetDevice(&cur);
    if (cur != device) cudaSetDevice(device);
    cudaStreamSynchronize(stream);
    if (cur != device) cudaSetDevice(cur);
  }
  if (own_stream) cudaStreamDestroy(own_stream);
}

// Scratch that queued kernels may still be reading is only replaced after the
// stream drained (the old block returns to the pool, which other objects
// allocate from); growth is rare -- the ROIs of a panorama are alike.
static void grow_scratch(DevBuf& b, size_t n, cudaStream_t s) {
  if (n <= b.bytes) return;
  if (b.p) XP_CUDA(cudaStreamSynchronize(s));
  b.reserve(n);
}

static bool host_trig_tables() {
  static const bool on = [] {
    const char* e = getenv("XP_HOST_TABLES");
    return e ? atoi(e) != 0 : true;
  }();
  return on;
}

void Warper::tab_reset(size_t bytes) {
  tab_box.reset(std::max(tab_box.cap, std::max(bytes, (size_t)1 << 20)));
}

MapSrc Warper::prepare_map(const Projector& p, const int roi[4]) {
  MapSrc m;
  m.p = p;
  m.tlx = roi[0];
  m.tly = roi[1];
  m.col = m.row = nullptr;
  m.separable = is_separable(p.type) ? 1 : 0;
  const int W = roi[2] - roi[0] + 1, H = roi[3] - roi[1] + 1;
  if (m.separable) {
    grow_scratch(tables, sizeof(float2) * (size_t)(W + H), stream);
    float2* col = tables.as<float2>();
    float2* row = col + W;
    if (host_trig_tables()) {
      // The W + H sines and cosines of a separable projection come from the
      // host's libm -- the one OpenCV's projectors call (warpers_inl.hpp:
      // sinf / cosf) -- so the inverse map is the reference's bit for bit;
      // CUDA's sinf / cosf differ from glibc's in the last bit for ~7 % of
      // the arguments, which moves a handful of mask pixels per panorama.
      // Everything per pixel (sep_combine: products, the K R^-1 transform,
      // one division) is IEEE arithmetic on both sides.
      const size_t bytes = sizeof(float2) * (size_t)(W + H);
      if (tab_box.cap - tab_box.used < ((bytes + 15) & ~(size_t)15)) {
        XP_CUDA(cudaStreamSynchronize(stream));
        tab_box.reset(std::max(bytes * 4, (size_t)1 << 20));
      }
      float2* h = static_cast<float2*>(tab_box.take(bytes));
      for (int i = 0; i < W; ++i) {
        float S, C;
        sep_col(p, (float)(i + roi[0]), S, C);
        h[i] = make_float2(S, C);
      }
      for (int j = 0; j < H; ++j) {
        float A, B;
        sep_row(p, (float)(j + roi[1]), A, B);
        h[W + j] = make_float2(A, B);
      }
      void* dh = nullptr;
      XP_CUDA(cudaHostGetDevicePointer(&dh, h, 0));
      tables_from_host_kernel<<<div_up(W + H, 256), 256, 0, stream>>>(
          static_cast<const float2*>(dh), col, W + H);
    } else {
      sep_tables_kernel<<<div_up(W + H, 256), 256, 0, stream>>>(
          p, roi[0], roi[1], W, H, col, row);
    }
    XP_LAUNCHED();
    m.col = col;
    m.row = row;
  }
  return m;
}

static dim3 grid2d(int W, int H, dim3 b) {
  return dim3(div_up(W, (int)b.x), div_up(H, (int)b.y));
}

void Warper::warp_roi(int sw, int sh, const float* K, const float* R,
                      int out_xywh[4]) {
  XP_REQUIRE(sw > 0 && sh > 0, "warpRoi: empty source size");
  Projector p = projector(K, R);
  int roi[4];
  detect_roi(p, sw, sh, roi);
  out_xywh[0] = roi[0];
  out_xywh[1] = roi[1];
  out_xywh[2] = roi[2] - roi[0] + 1;
  out_xywh[3] = roi[3] - roi[1] + 1;
}

static void check_roi(const int roi[4]) {
  const long W = (long)roi[2] - roi[0] + 1, H = (long)roi[3] - roi[1] + 1;
  if (W <= 0 || H <= 0 || W > (1 << 20) || H > (1 << 20) ||
      W * H > (1L << 33))
    fail(XP_ERR_INVALID, "warp: degenerate result ROI %ldx%ld", W, H);
}

void Warper::build_maps(int sw, int sh, const float* K, const float* R,
                        float* xmap, float* ymap, size_t stride, int mem,
                        int out_xywh[4]) {
  ensure_device();
  Projector p = projector(K, R);
  int roi[4];
  detect_roi(p, sw, sh, roi);
  check_roi(roi);
  const int W = roi[2] - roi[0] + 1, H = roi[3] - roi[1] + 1;
  out_xywh[0] = roi[0];
  out_xywh[1] = roi[1];
  out_xywh[2] = W;
  out_xywh[3] = H;
  if (!xmap || !ymap) return;
  if (stride == 0) stride = W;
  MapSrc m = prepare_map(p, roi);
  float *dx = xmap, *dy = ymap;
  size_t dstride = stride;
  if (mem == XP_MEM_HOST) {
    staging_dst.reserve(sizeof(float) * 2 * (size_t)W * H);
    dx = staging_dst.as<float>();
    dy = dx + (size_t)W * H;
    dstride = W;
  }
  dim3 b(32, 8);
  build_maps_kernel<<<grid2d(W, H, b), b, 0, stream>>>(m, W, H, dx, dy,
                                                        dstride);
  XP_LAUNCHED();
  if (mem == XP_MEM_HOST) {
    copy_out_2d(xmap, stride * 4, dx, (size_t)W * 4, (size_t)W * 4, H, mem,
                stream);
    copy_out_2d(ymap, stride * 4, dy, (size_t)W * 4, (size_t)W * 4, H, mem,
                stream);
  }
  XP_CUDA(cudaStreamSynchronize(stream));
}

const void* Warper::stage_src(const void* src, int sw, int sh, size_t stride,
                              size_t row_bytes, int mem, size_t* dev_pitch) {
  if (mem == XP_MEM_DEVICE) {
    *dev_pitch = stride;
    return src;
  }
  staging_src.reserve(row_bytes * (size_t)sh);
  copy_in_2d(staging_src.p, row_bytes, src, stride, row_bytes, sh, mem, stream);
  *dev_pitch = row_bytes;
  return staging_src.p;
}

void Warper::warp(const void* src, int sw, int sh, size_t src_stride, int depth,
                  int channels, const float* K, const float* R, int interp,
                  int border, void* dst, size_t dst_stride, int mem,
                  int out_corner[2]) {
  XP_REQUIRE(src && dst, "warp: null buffer");
  XP_REQUIRE(depth == 8 || depth == 16, "warp: depth must be 8 or 16");
  XP_REQUIRE(channels == 1 || channels == 3, "warp: channels must be 1 or 3");
  const bool linear = interp == XP_INTER_LINEAR && border == XP_BORDER_REFLECT;
  const bool nearest =
      interp == XP_INTER_NEAREST && border == XP_BORDER_CONSTANT;
  if (!linear && !nearest)
    fail(XP_ERR_UNSUPPORTED,
         "warp: only INTER_LINEAR+BORDER_REFLECT and "
         "INTER_NEAREST+BORDER_CONSTANT are on the export path");
  ensure_device();
  Projector p = projector(K, R);
  int roi[4];
  detect_roi(p, sw, sh, roi);
  check_roi(roi);
  const int W = roi[2] - roi[0] + 1, H = roi[3] - roi[1] + 1;
  const size_t bs = depth / 8;
  const size_t src_row = (size_t)sw * channels * bs;
  const size_t dst_row = (size_t)W * channels * bs;
  if (src_stride == 0) src_stride = src_row;
  if (dst_stride == 0) dst_stride = dst_row;
  XP_REQUIRE(src_stride % bs == 0 && dst_stride % bs == 0,
             "warp: stride not a multiple of the sample size");
  size_t spitch;
  const void* dsrc = stage_src(src, sw, sh, src_stride, src_row, mem, &spitch);
  void* ddst = dst;
  size_t dpitch = dst_stride;
  if (mem == XP_MEM_HOST) {
    staging_dst.reserve(dst_row * (size_t)H);
    ddst = staging_dst.p;
    dpitch = dst_row;
  }
  MapSrc m = prepare_map(p, roi);
  dim3 b(32, 8);
  dim3 g = grid2d(W, H, b);
#define XP_WARP_CASE(T, CN)                                                  \
  if (linear)                                                                \
    warp_linear_kernel<T, CN><<<g, b, 0, stream>>>(                          \
        m, (const T*)dsrc, spitch / sizeof(T), sw, sh, W, H, (T*)ddst,       \
        dpitch / sizeof(T));                                                 \
  else                                                                       \
    warp_nearest_kernel<T, CN><<<g, b, 0, stream>>>(                         \
        m, (const T*)dsrc, spitch / sizeof(T), sw, sh, W, H, (T*)ddst,       \
        dpitch / sizeof(T));
  if (depth == 8 && channels == 1) {
    XP_WARP_CASE(uint8_t, 1)
  } else if (depth == 8) {
    XP_WARP_CASE(uint8_t, 3)
  } else if (channels == 1) {
    XP_WARP_CASE(uint16_t, 1)
  } else {
    XP_WARP_CASE(uint16_t, 3)
  }
#undef XP_WARP_CASE
  XP_LAUNCHED();
  if (mem == XP_MEM_HOST)
    copy_out_2d(dst, dst_stride, ddst, dpitch, dst_row, H, mem, stream);
  XP_CUDA(cudaStreamSynchronize(stream));
  out_corner[0] = roi[0];
  out_corner[1] = roi[1];
}

void Warper::warp_bgra_device(const void* dsrc, int sw, int sh, size_t spitch,
                              int depth, const Projector& p, const int roi[4],
                              const float* d_gain, int gw, int gh,
                              const uint8_t* d_seam, int mw, int mh,
                              void* ddst, size_t dpitch,
                              bool skip_transparent, uint32_t* abits,
                              int abits_pitch) {
  const int W = roi[2] - roi[0] + 1, H = roi[3] - roi[1] + 1;
  const AlphaOut ao = {abits, abits_pitch};
  const uint8_t* d_dil = nullptr;
  if (d_seam) {
    grow_scratch(dilated, (size_t)mw * mh, stream);
    dim3 b(32, 8);
    dilate3x3_kernel<<<grid2d(mw, mh, b), b, 0, stream>>>(d_seam, mw, mh,
                                                          dilated.as<uint8_t>());
    XP_LAUNCHED();
    d_dil = dilated.as<uint8_t>();
  }
  if (!dsrc) d_gain = nullptr;  // alpha only: colours are not produced
  const AuxCoef *aux_col = nullptr, *aux_row = nullptr;
  if (d_gain || d_dil) {
    grow_scratch(aux_tables, sizeof(AuxCoef) * (size_t)(W + H), stream);
    AuxCoef* t = aux_tables.as<AuxCoef>();
    aux_tables_kernel<<<div_up(W + H, 256), 256, 0, stream>>>(
        W, H, d_gain ? gw : 0, d_gain ? gh : 0, d_dil ? mw : 0,
        d_dil ? mh : 0, t, t + W);
    XP_LAUNCHED();
    aux_col = t;
    aux_row = t + W;
  }
  MapSrc m = prepare_map(p, roi);
  dim3 b(32, 8);
  dim3 g = grid2d(W, H, b);
  if (!dsrc) {
    if (depth == 8)
      warp_alpha_kernel<uint8_t><<<g, b, 0, stream>>>(
          m, sw, sh, W, H, d_dil, mw, mh, aux_col, aux_row, (uint8_t*)ddst,
          dpitch, ao);
    else
      warp_alpha_kernel<uint16_t><<<g, b, 0, stream>>>(
          m, sw, sh, W, H, d_dil, mw, mh, aux_col, aux_row, (uint16_t*)ddst,
          dpitch / 2, ao);
    XP_LAUNCHED();
    return;
  }
#define XP_WARP_BGRA(T, SKIP, DIV)                                          \
  warp_bgra_kernel<T, SKIP><<<g, b, 0, stream>>>(                           \
      m, (const T*)dsrc, spitch / DIV, sw, sh, W, H, d_gain, gw, gh, d_dil, \
      mw, mh, aux_col, aux_row, (T*)ddst, dpitch / DIV, ao)
  if (depth == 8) {
    if (skip_transparent)
      XP_WARP_BGRA(uint8_t, true, 1);
    else
      XP_WARP_BGRA(uint8_t, false, 1);
  } else {
    if (skip_transparent)
      XP_WARP_BGRA(uint16_t, true, 2);
    else
      XP_WARP_BGRA(uint16_t, false, 2);
  }
#undef XP_WARP_BGRA
  XP_LAUNCHED();
}

void Warper::warp_bgra(const void* src, int sw, int sh, size_t src_stride,
                       int depth, const float* K, const float* R,
                       const xp_compose_aux* aux, int aux_mem, void* dst,
                       size_t dst_stride, int mem, int out_xywh[4]) {
  XP_REQUIRE(src, "warp_bgra: null source");
  XP_REQUIRE(depth == 8 || depth == 16, "warp_bgra: depth must be 8 or 16");
  ensure_device();
  Projector p = projector(K, R);
  int roi[4];
  detect_roi(p, sw, sh, roi);
  check_roi(roi);
  const int W = roi[2] - roi[0] + 1, H = roi[3] - roi[1] + 1;
  out_xywh[0] = roi[0];
  out_xywh[1] = roi[1];
  out_xywh[2] = W;
  out_xywh[3] = H;
  if (!dst) return;  // size query
  const size_t bs = depth / 8;
  const size_t src_row = (size_t)sw * 3 * bs, dst_row = (size_t)W * 4 * bs;
  if (src_stride == 0) src_stride = src_row;
  if (dst_stride == 0) dst_stride = dst_row;
  size_t spitch;
  const void* dsrc = stage_src(src, sw, sh, src_stride, src_row, mem, &spitch);
  void* ddst = dst;
  size_t dpitch = dst_stride;
  if (mem == XP_MEM_HOST) {
    staging_dst.reserve(dst_row * (size_t)H);
    ddst = staging_dst.p;
    dpitch = dst_row;
  }
  XP_REQUIRE(dpitch % (4 * bs) == 0, "warp_bgra: dst stride must keep pixels aligned");
  const float* d_gain = nullptr;
  const uint8_t* d_seam = nullptr;
  int gw = 0, gh = 0, mw = 0, mh = 0;
  if (aux && aux->gain) {
    gw = aux->gain_w;
    gh = aux->gain_h;
    XP_REQUIRE(gw > 0 && gh > 0, "warp_bgra: empty gain map");
    if (aux_mem == XP_MEM_DEVICE) {
      d_gain = aux->gain;
    } else {
      aux_gain.reserve(sizeof(float) * (size_t)gw * gh);
      XP_CUDA(cudaMemcpyAsync(aux_gain.p, aux->gain,
                              sizeof(float) * (size_t)gw * gh,
                              cudaMemcpyHostToDevice, stream));
      d_gain = aux_gain.as<float>();
    }
  }
  if (aux && aux->seam_mask) {
    mw = aux->seam_w;
    mh = aux->seam_h;
    XP_REQUIRE(mw > 0 && mh > 0, "warp_bgra: empty seam mask");
    if (aux_mem == XP_MEM_DEVICE) {
      d_seam = aux->seam_mask;
    } else {
      aux_seam.reserve((size_t)mw * mh);
      XP_CUDA(cudaMemcpyAsync(aux_seam.p, aux->seam_mask, (size_t)mw * mh,
                              cudaMemcpyHostToDevice, stream));
      d_seam = aux_seam.as<uint8_t>();
    }
  }
  warp_bgra_device(dsrc, sw, sh, spitch, depth, p, roi, d_gain, gw, gh, d_seam,
                   mw, mh, ddst, dpitch);
  if (mem == XP_MEM_HOST)
    copy_out_2d(dst, dst_stride, ddst, dpitch, dst_row, H, mem, stream);
  XP_CUDA(cudaStreamSynchronize(stream));
}

// ---------------------------------------------------------------------------
// Allocation-free warp (xp_warper_warp_noalloc). The only scratch of a warp
// with a separable projection is the per-column / per-row trig table
// (sep_tables_kernel); here it lives in a file-scope device array that every
// GPU loading the library holds until exit, so the call needs no cudaMalloc.
// ---------------------------------------------------------------------------
constexpr int kNoAllocTableEntries = XP_NOALLOC_MAX_ROI_PERIMETER;  // float2 each
static __device__ __align__(256) float2 g_noalloc_tables[kNoAllocTableEntries];

// In host code `g_noalloc_tables` is only a handle for the runtime, not a
// device address: the address is looked up once per device, with that device
// current, and cached.
static float2* noalloc_tables(int device) {
  static std::mutex mu;
  static float2* ptr[64] = {};
  XP_REQUIRE(device >= 0 && device < 64, "device %d out of range", device);
  std::lock_guard<std::mutex> lk(mu);
  if (!ptr[device]) {
    void* p = nullptr;
    XP_CUDA(cudaGetSymbolAddress(&p, g_noalloc_tables));
    XP_REQUIRE(p, "cudaGetSymbolAddress returned a null address");
    ptr[device] = static_cast<float2*>(p);
  }
  return ptr[device];
}

void warp_noalloc(const Warper& wr, const void* d_src, int sw, int sh,
                  size_t src_stride, int depth, int channels, const float* K,
                  const float* R, int interp, int border, void* d_dst,
                  size_t dst_stride, cudaStream_t stream, int out_corner[2]) {
  XP_REQUIRE(d_src && d_dst, "warp_noalloc: null buffer");
  XP_REQUIRE(sw > 0 && sh > 0, "warp_noalloc: empty source size");
  XP_REQUIRE(depth == 8 || depth == 16, "warp_noalloc: depth must be 8 or 16");
  XP_REQUIRE(channels == 1 || channels == 3,
             "warp_noalloc: channels must be 1 or 3");
  const bool linear = interp == XP_INTER_LINEAR && border == XP_BORDER_REFLECT;
  const bool nearest =
      interp == XP_INTER_NEAREST && border == XP_BORDER_CONSTANT;
  if (!linear && !nearest)
    fail(XP_ERR_UNSUPPORTED,
         "warp_noalloc: only INTER_LINEAR+BORDER_REFLECT and "
         "INTER_NEAREST+BORDER_CONSTANT are on the export path");
  if (wr.type != XP_PROJ_SPHERICAL && wr.type != XP_PROJ_CYLINDRICAL)
    fail(XP_ERR_UNSUPPORTED,
         "warp_noalloc: only the spherical and cylindrical projections (host "
         "ROI walk, separable maps) run without device allocations");
  select_device(wr.device);
  Projector p = wr.projector(K, R);
  int roi[4];
  {  // border-walk ROI: host arithmetic only (see Warper::detect_roi)
    Warper host_only;
    host_only.type = wr.type;
    host_only.detect_roi(p, sw, sh, roi);
  }
  check_roi(roi);
  const int W = roi[2] - roi[0] + 1, H = roi[3] - roi[1] + 1;
  if ((long)W + H > kNoAllocTableEntries)
    fail(XP_ERR_INVALID,
         "warp_noalloc: result ROI %dx%d needs %ld table entries, the static "
         "scratch holds %d (width + height <= XP_NOALLOC_MAX_ROI_PERIMETER)",
         W, H, (long)W + H, kNoAllocTableEntries);
  const size_t bs = depth / 8;
  if (src_stride == 0) src_stride = (size_t)sw * channels * bs;
  if (dst_stride == 0) dst_stride = (size_t)W * channels * bs;
  XP_REQUIRE(src_stride % bs == 0 && dst_stride % bs == 0,
             "warp_noalloc: stride not a multiple of the sample size");
  MapSrc m;
  m.p = p;
  m.tlx = roi[0];
  m.tly = roi[1];
  m.separable = 1;
  float2* col = noalloc_tables(wr.device);
  float2* row = col + W;
  // every entry that is read (W + H of them) is written here first: nothing
  // carried over from the previous call is used
  if (host_trig_tables()) {
    // as in Warper::prepare_map, the sines and cosines come from the host's
    // libm; they are formed in a static host array (plain memory, no
    // allocation) and handed to the driver as a pageable copy, which has
    // taken the bytes when the call returns
    static std::mutex mu;
    static float2 h_tab[kNoAllocTableEntries];
    std::lock_guard<std::mutex> lk(mu);
    for (int i = 0; i < W; ++i) {
      float S, C;
      sep_col(p, (float)(i + roi[0]), S, C);
      h_tab[i] = make_float2(S, C);
    }
    for (int j = 0; j < H; ++j) {
      float A, B;
      sep_row(p, (float)(j + roi[1]), A, B);
      h_tab[W + j] = make_float2(A, B);
    }
    XP_CUDA(cudaMemcpyAsync(col, h_tab, sizeof(float2) * (size_t)(W + H),
                            cudaMemcpyHostToDevice, stream));
  } else {
    sep_tables_kernel<<<div_up(W + H, 256), 256, 0, stream>>>(
        p, roi[0], roi[1], W, H, col, row);
    XP_LAUNCHED();
  }
  m.col = col;
  m.row = row;
  dim3 b(32, 8);
  dim3 g = grid2d(W, H, b);
#define XP_WARP_CASE(T, CN)                                                  \
  if (linear)                                                                \
    warp_linear_kernel<T, CN><<<g, b, 0, stream>>>(                          \
        m, (const T*)d_src, src_stride / sizeof(T), sw, sh, W, H, (T*)d_dst, \
        dst_stride / sizeof(T));                                             \
  else                                                                       \
    warp_nearest_kernel<T, CN><<<g, b, 0, stream>>>(                         \
        m, (const T*)d_src, src_stride / sizeof(T), sw, sh, W, H, (T*)d_dst, \
        dst_stride / sizeof(T));
  if (depth == 8 && channels == 1) {
    XP_WARP_CASE(uint8_t, 1)
  } else if (depth == 8) {
    XP_WARP_CASE(uint8_t, 3)
  } else if (channels == 1) {
    XP_WARP_CASE(uint16_t, 1)
  } else {
    XP_WARP_CASE(uint16_t, 3)
  }
#undef XP_WARP_CASE
  XP_LAUNCHED();
  out_corner[0] = roi[0];
  out_corner[1] = roi[1];
}

}  // namespace xp

// ---------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------
using xp::guard;

struct xp_warper : xp::Warper {};

extern "C" {

xp_status xp_warper_create(int projection, int portrait, float scale, float a,
                           float b, int device, xp_warper** out) {
  return guard([&] {
    XP_REQUIRE(out, "xp_warper_create: null out");
    XP_REQUIRE(projection >= 0 && projection <= 8,
               "unknown projection %d", projection);
    auto* w = new xp_warper();
    w->type = projection;
    w->portrait = portrait ? 1 : 0;
    w->scale = scale;
    w->a = a;
    w->b = b;
    w->device = device;
    *out = w;
  });
}

void xp_warper_destroy(xp_warper* w) { delete w; }

float xp_warper_get_scale(const xp_warper* w) { return w ? w->scale : 0.f; }

xp_status xp_warper_set_scale(xp_warper* w, float scale) {
  return guard([&] {
    XP_REQUIRE(w, "null warper");
    w->scale = scale;
  });
}

xp_status xp_warper_set_stream(xp_warper* w, void* cuda_stream) {
  return guard([&] {
    XP_REQUIRE(w, "null warper");
    w->stream = (cudaStream_t)cuda_stream;
    w->external_stream = true;
  });
}

xp_status xp_warper_warp_roi(xp_warper* w, int src_w, int src_h,
                             const float K[9], const float R[9],
                             int out_xywh[4]) {
  return guard([&] {
    XP_REQUIRE(w && K && R && out_xywh, "warp_roi: null argument");
    w->warp_roi(src_w, src_h, K, R, out_xywh);
  });
}

xp_status xp_warper_warp_point(xp_warper* w, const float xy[2],
                               const float K[9], const float R[9],
                               float uv_out[2]) {
  return guard([&] {
    XP_REQUIRE(w && xy && K && R && uv_out, "warp_point: null argument");
    xp::Projector p = w->projector(K, R);
    xp::map_forward(p, xy[0], xy[1], uv_out[0], uv_out[1]);
  });
}

xp_status xp_warper_warp_point_backward(xp_warper* w, const float uv[2],
                                        const float K[9], const float R[9],
                                        float xy_out[2]) {
  return guard([&] {
    XP_REQUIRE(w && uv && K && R && xy_out, "warp_point: null argument");
    xp::Projector p = w->projector(K, R);
    xp::map_backward(p, uv[0], uv[1], xy_out[0], xy_out[1]);
  });
}

xp_status xp_warper_build_maps(xp_warper* w, int src_w, int src_h,
                               const float K[9], const float R[9], float* xmap,
                               float* ymap, size_t stride, int mem,
                               int out_xywh[4]) {
  return guard([&] {
    XP_REQUIRE(w && K && R && out_xywh, "build_maps: null argument");
    w->build_maps(src_w, src_h, K, R, xmap, ymap, stride, mem, out_xywh);
  });
}

xp_status xp_warper_warp(xp_warper* w, const void* src, int src_w, int src_h,
                         size_t src_stride, int depth, int channels,
                         const float K[9], const float R[9], int interp,
                         int border, void* dst, size_t dst_stride, int mem,
                         int out_corner[2]) {
  return guard([&] {
    XP_REQUIRE(w && K && R && out_corner, "warp: null argument");
    w->warp(src, src_w, src_h, src_stride, depth, channels, K, R, interp,
            border, dst, dst_stride, mem, out_corner);
  });
}

xp_status xp_warper_warp_noalloc(xp_warper* w, const void* d_src, int src_w,
                                 int src_h, size_t src_stride, int depth,
                                 int channels, const float K[9],
                                 const float R[9], int interp, int border,
                                 void* d_dst, size_t dst_stride,
                                 void* cuda_stream, int out_corner[2]) {
  return guard([&] {
    XP_REQUIRE(w && K && R && out_corner, "warp_noalloc: null argument");
    xp::warp_noalloc(*w, d_src, src_w, src_h, src_stride, depth, channels, K,
                     R, interp, border, d_dst, dst_stride,
                     (cudaStream_t)cuda_stream, out_corner);
  });
}

xp_status xp_warper_warp_bgra(xp_warper* w, const void* src, int src_w,
                              int src_h, size_t src_stride, int depth,
                              const float K[9], const float R[9],
                              const xp_compose_aux* aux, int aux_mem, void* dst,
                              size_t dst_stride, int mem, int out_xywh[4]) {
  return guard([&] {
    XP_REQUIRE(w && K && R && out_xywh, "warp_bgra: null argument");
    w->warp_bgra(src, src_w, src_h, src_stride, depth, K, R, aux, aux_mem, dst,
                 dst_stride, mem, out_xywh);
  });
}

}  // extern "C"
