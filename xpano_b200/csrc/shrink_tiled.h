// Tiled Shrink (pyramid.cpp:584-823): a block produces 64 x 8 lo samples.
// The hi rows and columns they reference are staged in shared memory once
// (coalesced loads), the vertical form is applied per column into a second
// shared tile, then the five horizontal taps are combined. All border rules
// (row forms, column clamps, parity shifts) come from pyramid_math.h index
// functions. Blocks whose window touches no border (every row the five-tap
// form on consecutive hi rows, no column clamped) take a path with fixed
// loop shapes that converts each hi sample once; the same arithmetic per
// sample, so both paths agree bit for bit.
//
// `Src` is the hi level: raw(row, col) loads one stored element, value(raw, c)
// converts it to the float sample of channel c (kChannels channels per
// element: 3 for interleaved source pixels, 1 for a float plane).
#pragma once

#include "pyramid_math.h"

namespace xp {

constexpr int kShrinkTileW = 64, kShrinkTileH = 8;
constexpr int kShrinkMaxCols = 2 * kShrinkTileW + 6;  // 2*(W-1) + 4 + 1 (+xs)
constexpr int kShrinkMaxRows = 2 * kShrinkTileH + 4;

// ---- TMA / mbarrier primitives (sm_90+: cp.async.bulk.tensor) -------------
// One elected thread arms an mbarrier with the byte count of the box and
// issues the bulk tensor copy; the copy engine walks the rows of the box,
// writes them densely into shared memory and completes the transaction on
// the barrier, which every thread of the block then waits on. No per-element
// address arithmetic, no per-thread loads.
__device__ __forceinline__ unsigned smem_addr(const void* p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)),
               "r"(count)
               : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar,
                                               unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(
                   smem_addr(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap,
                                            int x, int y,
                                            unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::"
      "bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_addr(smem_dst)),
      "l"(tmap), "r"(x), "r"(y), "r"(smem_addr(bar))
      : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar,
                                          unsigned parity) {
  unsigned done;
  do {
    asm volatile(
        "{\n.reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n}"
        : "=r"(done)
        : "r"(smem_addr(bar)), "r"(parity)
        : "memory");
  } while (!done);
}

// `tmap`: a CUtensorMap over the source pixels (2-D, 32-bit elements, box
// kFastPitch x kFastRows) in global memory, or null.
template <class Src>
__device__ __forceinline__ void shrink_tile(const Src& hi, int hh, int hp,
                                            int xs, int ys,
                                            float* __restrict__ lo, int lo_h,
                                            int lo_pitch, size_t chan_stride,
                                            int bx, int by, int z,
                                            const void* tmap = nullptr) {
  typedef typename Src::Raw Raw;
  constexpr int NC = Src::kChannels;
  constexpr int kFastRows = 2 * kShrinkTileH + 3, kFastCols = 2 * kShrinkTileW + 3;
  constexpr int kFastPitch = kFastCols + 1;
  constexpr size_t kRawBytes = sizeof(Raw) * kShrinkMaxRows * kShrinkMaxCols;
  constexpr size_t kFastBytes = sizeof(float) * NC * kFastRows * kFastPitch;
  __shared__ __align__(128) unsigned char
      stage[kRawBytes > kFastBytes ? kRawBytes : kFastBytes];
  __shared__ __align__(8) unsigned long long tma_bar;
  Raw* tile = reinterpret_cast<Raw*>(stage);
  __shared__ float vert[NC][kShrinkTileH][kShrinkMaxCols];
  __shared__ pm::VForm forms[kShrinkTileH];
  __shared__ int s_rmin, s_rmax;
  const int tid = threadIdx.x;
  const int j0 = bx * kShrinkTileW, y0 = by * kShrinkTileH;
  const int ny = min(kShrinkTileH, lo_h - y0);
  // z: plane for single-channel sources
  if (tid == 0) s_rmin = 1 << 30, s_rmax = -1;
  __syncthreads();
  if (tid < ny) {
    const pm::VForm f = pm::shrink_vform(hh, ys, lo_h, y0 + tid);
    forms[tid] = f;
    const int taps = f.kind == 3 ? 5 : (f.kind == 1 ? 3 : (f.kind == 0 ? 1 : 2));
    int lo_r = f.r[0], hi_r = f.r[0];
    for (int k = 1; k < taps; ++k) lo_r = min(lo_r, f.r[k]), hi_r = max(hi_r, f.r[k]);
    atomicMin(&s_rmin, lo_r);
    atomicMax(&s_rmax, hi_r);
  }
  // interior block: full tile, five consecutive unclamped rows per lo row,
  // no clamped column
  bool mine_ok = true;
  if (tid < kShrinkTileH) {
    mine_ok = false;
    if (tid < ny) {
      const pm::VForm& f = forms[tid];
      mine_ok = f.kind == 3 && f.r[0] == f.r[2] - 2 && f.r[1] == f.r[2] - 1 &&
                f.r[3] == f.r[2] + 1 && f.r[4] == f.r[2] + 2;
    }
  }
  const bool cols_ok = 2 * j0 - 4 - xs >= 0 &&
                       2 * (j0 + kShrinkTileW - 1) <= hp - 1 &&
                       j0 + kShrinkTileW <= lo_pitch;
  const bool fast = __syncthreads_and(mine_ok && cols_ok && ny == kShrinkTileH);
  const int rmin = s_rmin, nrows = s_rmax - rmin + 1;
  if constexpr (Src::kPackedBytes) {
    // 8-bit source pixels, no bit-depth conversion: every intermediate of the
    // five-tap forms is an integer below 2^16 (255 * 16 * 16 = 65280), so the
    // sums are formed on two channels at a time in 16-bit halves of a word
    // and converted once; the float arithmetic of the general path has
    // nothing to round on these values, the results are the same bits.
    if (fast && hi.mul == 1.0f) {
      // staged source pixels: kFastRows rows of kTmaPitch words; the window
      // starts `off` words into a row (TMA boxes start on 16-byte boundaries
      // of global memory: the innermost coordinate is rounded down to a
      // multiple of four pixels and the box is four pixels wider)
      constexpr int kTmaPitch = (kFastCols + 3 + 3) & ~3;  // 136
      unsigned* R = reinterpret_cast<unsigned*>(stage);
      unsigned(*VE)[kShrinkMaxCols] =
          reinterpret_cast<unsigned(*)[kShrinkMaxCols]>(&vert[0][0][0]);
      unsigned(*VO)[kShrinkMaxCols] =
          reinterpret_cast<unsigned(*)[kShrinkMaxCols]>(&vert[1][0][0]);
      const int c0 = 2 * j0 - 4 - xs;
      static_assert(kTmaPitch * 4 % 16 == 0, "TMA box rows are 16-byte units");
      static_assert(kTmaPitch == 2 * kShrinkTileW + 8, "box of the tensor map");
      static_assert(sizeof(stage) >= sizeof(unsigned) * kFastRows * kTmaPitch,
                    "the box fits the staging tile");
      int off = 0;
      // the box must not reach the columns right of the image, which the
      // plain path replicates (pyramid.cpp:268-330)
      if (tmap != nullptr && c0 + kFastCols - 1 <= hi.w - 1) {
        const int gx = hi.left + c0;
        off = gx & 3;
        if (tid == 0) mbar_init(&tma_bar, 1);
        __syncthreads();
        if (tid == 0) {
          mbar_expect_tx(&tma_bar, sizeof(unsigned) * kFastRows * kTmaPitch);
          tma_load_2d(stage, tmap, gx - off, hi.top + rmin, &tma_bar);
        }
        mbar_wait(&tma_bar, 0);
      } else {
        constexpr int kPer = (kFastRows * kFastCols + 255) / 256;
        unsigned v[kPer];
#pragma unroll
        for (int k = 0; k < kPer; ++k) {
          const int i = tid + 256 * k;
          const int r = i / kFastCols, c = i - r * kFastCols;
          if (i < kFastRows * kFastCols)
            v[k] = (unsigned)hi.raw(z, rmin + r, c0 + c);
        }
#pragma unroll
        for (int k = 0; k < kPer; ++k) {
          const int i = tid + 256 * k;
          const int r = i / kFastCols, c = i - r * kFastCols;
          if (i < kFastRows * kFastCols) R[r * kTmaPitch + c] = v[k];
        }
        __syncthreads();
      }
      for (int i = tid; i < kShrinkTileH * kFastCols; i += 256) {
        const int yy = i / kFastCols, c = i - yy * kFastCols;
        const unsigned* col = R + (2 * yy) * kTmaPitch + c + off;
        const unsigned r0 = col[0], r1 = col[kTmaPitch],
                       r2 = col[2 * kTmaPitch], r3 = col[3 * kTmaPitch],
                       r4 = col[4 * kTmaPitch];
        constexpr unsigned kE = 0x00ff00ffu;  // channels 0 and 2
        const unsigned e = ((r0 & kE) + (r4 & kE)) +
                           (((r1 & kE) + (r3 & kE)) << 2) + (r2 & kE) * 6u;
        const unsigned o = (((r0 >> 8) & 0xffu) + ((r4 >> 8) & 0xffu)) +
                           ((((r1 >> 8) & 0xffu) + ((r3 >> 8) & 0xffu)) << 2) +
                           ((r2 >> 8) & 0xffu) * 6u;
        VE[yy][c] = e;
        VO[yy][c] = o;
      }
      __syncthreads();
      const int jj = tid & (kShrinkTileW - 1);
#pragma unroll
      for (int it = 0; it < kShrinkTileH * kShrinkTileW / 256; ++it) {
        const int yy = (tid / kShrinkTileW) + it * (256 / kShrinkTileW);
        const unsigned* pe = &VE[yy][2 * jj];
        const unsigned* po = &VO[yy][2 * jj];
        const unsigned he = (pe[0] + pe[4]) + (((pe[1] + pe[3]) << 2) + pe[2] * 6u);
        const unsigned ho = (po[0] + po[4]) + (((po[1] + po[3]) << 2) + po[2] * 6u);
        const size_t o = (size_t)(y0 + yy) * lo_pitch + j0 + jj;
        lo[o] = (float)(he & 0xffffu) * (1.0f / 256);
        lo[chan_stride + o] = (float)ho * (1.0f / 256);
        lo[2 * chan_stride + o] = (float)(he >> 16) * (1.0f / 256);
      }
      return;
    }
  }
  if (fast) {
    float(*S)[kFastRows][kFastPitch] =
        reinterpret_cast<float(*)[kFastRows][kFastPitch]>(stage);
    const int c0 = 2 * j0 - 4 - xs;
    // all loads of a thread are issued before the first conversion
    constexpr int kPer = (kFastRows * kFastCols + 255) / 256;
    Raw v[kPer];
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      const int i = tid + 256 * k;
      const int r = i / kFastCols, c = i - r * kFastCols;
      if (i < kFastRows * kFastCols) v[k] = hi.raw(z, rmin + r, c0 + c);
    }
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      const int i = tid + 256 * k;
      const int r = i / kFastCols, c = i - r * kFastCols;
      if (i < kFastRows * kFastCols) {
#pragma unroll
        for (int ch = 0; ch < NC; ++ch) S[ch][r][c] = hi.value(v[k], ch);
      }
    }
    __syncthreads();
    for (int i = tid; i < kShrinkTileH * kFastCols; i += 256) {
      const int yy = i / kFastCols, c = i - yy * kFastCols;
#pragma unroll
      for (int ch = 0; ch < NC; ++ch)
        vert[ch][yy][c] = pm::shrink_vcombine(
            3, S[ch][2 * yy][c], S[ch][2 * yy + 1][c], S[ch][2 * yy + 2][c],
            S[ch][2 * yy + 3][c], S[ch][2 * yy + 4][c]);
    }
    __syncthreads();
    const int jj = tid & (kShrinkTileW - 1);
#pragma unroll
    for (int it = 0; it < kShrinkTileH * kShrinkTileW / 256; ++it) {
      const int yy = (tid / kShrinkTileW) + it * (256 / kShrinkTileW);
#pragma unroll
      for (int ch = 0; ch < NC; ++ch) {
        float t[5];
#pragma unroll
        for (int m = 0; m < 5; ++m) t[m] = vert[ch][yy][2 * jj + m];
        lo[(size_t)(NC == 1 ? z : ch) * chan_stride +
           (size_t)(y0 + yy) * lo_pitch + j0 + jj] =
            pm::squeeze_combine(t, 1.0f / 256);
      }
    }
    return;
  }
  const int jl = min(j0 + kShrinkTileW - 1, lo_pitch - 1);
  const int cmin = pm::squeeze_col(j0, 0, hp, xs);
  const int ncols = pm::squeeze_col(jl, 4, hp, xs) - cmin + 1;
  // ---- stage the hi window ----
  for (int i = tid; i < nrows * ncols; i += 256) {
    const int r = i / ncols, c = i - r * ncols;
    tile[r * kShrinkMaxCols + c] = hi.raw(z, rmin + r, cmin + c);
  }
  __syncthreads();
  // ---- vertical forms, one (lo row, hi column) per item ----
  for (int i = tid; i < ny * ncols; i += 256) {
    const int yy = i / ncols, c = i - yy * ncols;
    const pm::VForm& f = forms[yy];
    Raw r[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      // rows of unused taps alias a staged row; their value is ignored
      int rr = f.r[k] - rmin;
      rr = rr < 0 ? 0 : (rr >= nrows ? nrows - 1 : rr);
      r[k] = tile[rr * kShrinkMaxCols + c];
    }
#pragma unroll
    for (int ch = 0; ch < NC; ++ch)
      vert[ch][yy][c] = pm::shrink_vcombine(
          f.kind, hi.value(r[0], ch), hi.value(r[1], ch), hi.value(r[2], ch),
          hi.value(r[3], ch), hi.value(r[4], ch));
  }
  __syncthreads();
  // ---- horizontal taps ----
  for (int i = tid; i < ny * kShrinkTileW; i += 256) {
    const int yy = i / kShrinkTileW, jj = i - yy * kShrinkTileW;
    const int j = j0 + jj;
    if (j >= lo_pitch) continue;
    int col[5];
#pragma unroll
    for (int m = 0; m < 5; ++m) col[m] = pm::squeeze_col(j, m, hp, xs) - cmin;
    const float mul = forms[yy].mul;
#pragma unroll
    for (int ch = 0; ch < NC; ++ch) {
      float t[5];
#pragma unroll
      for (int m = 0; m < 5; ++m) t[m] = vert[ch][yy][col[m]];
      lo[(size_t)(NC == 1 ? z : ch) * chan_stride +
         (size_t)(y0 + yy) * lo_pitch + j] = pm::squeeze_combine(t, mul);
    }
  }
}

template <class Src>
__global__ void __launch_bounds__(256)
    shrink_tiled_kernel(Src hi, int hh, int hp, int xs, int ys,
                        float* __restrict__ lo, int lo_h, int lo_pitch,
                        size_t chan_stride) {
  shrink_tile(hi, hh, hp, xs, ys, lo, lo_h, lo_pitch, chan_stride,
              (int)blockIdx.x, (int)blockIdx.y, (int)blockIdx.z);
}

// The same level of several pyramids in one launch (the upper levels of an
// image pyramid are a few hundred blocks each): blockIdx.z = job * planes +
// plane, blocks beyond a job's own extent leave at once.
template <class Src>
struct ShrinkJob {
  Src hi;
  int hh, hp, xs, ys;
  float* lo;
  int lo_h, lo_pitch;
  size_t chan_stride;
};
constexpr int kShrinkBatch = 16;
struct alignas(64) TensorMapBytes {
  unsigned char b[128];  // a CUtensorMap
};
template <class Src>
struct ShrinkBatch {
  ShrinkJob<Src> job[kShrinkBatch];
  // TMA tensor maps of the jobs' sources, in global memory (tmaps[k] belongs
  // to job k; bit k of `tma`: job k has one)
  const TensorMapBytes* tmaps;
  unsigned tma;
};

template <class Src>
__global__ void __launch_bounds__(256)
    shrink_batch_kernel(const __grid_constant__ ShrinkBatch<Src> batch,
                        int planes) {
  const ShrinkJob<Src>& j = batch.job[blockIdx.z / planes];
  if ((int)blockIdx.x * kShrinkTileW >= j.lo_pitch ||
      (int)blockIdx.y * kShrinkTileH >= j.lo_h)
    return;
  const int k = blockIdx.z / planes;
  shrink_tile(j.hi, j.hh, j.hp, j.xs, j.ys, j.lo, j.lo_h, j.lo_pitch,
              j.chan_stride, (int)blockIdx.x, (int)blockIdx.y,
              (int)(blockIdx.z % planes),
              (batch.tma >> k) & 1u ? (const void*)(batch.tmaps + k) : nullptr);
}

// float plane(s) as the hi level
struct PlaneSrc {
  typedef float Raw;
  static constexpr int kChannels = 1;
  static constexpr bool kPackedBytes = false;
  const float* p;  // channel 0
  int pitch;
  size_t cs;
  __device__ __forceinline__ float raw(int z, int row, int col) const {
    return __ldg(p + z * cs + (size_t)row * pitch + col);
  }
  __device__ __forceinline__ float value(float v, int) const { return v; }
};

}  // namespace xp
