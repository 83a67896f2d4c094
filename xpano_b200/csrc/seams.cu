// Mask / seam stage on the device: trim bbox, cover map, the two inpaint
// passes and the two seam passes (wavefront kernels), canvas statistics.
//
// Replaces external/multiblend/src/image.cpp:396-812 (trim, inpaint) and
// src/multiblend.cpp:229-823 (seams, full/xor masks, channel sums).
#include <climits>

#ifdef XP_WF_PROFILE
#include <cstdio>
static __host__ __device__ inline long long wf_now() {
#ifdef __CUDA_ARCH__
  return clock64();
#else
  return 0;
#endif
}
static __host__ __device__ inline unsigned long long wf_ns() {
#ifdef __CUDA_ARCH__
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
#else
  return 0;
#endif
}
#define XP_WF_CLOCK_BEGIN()                                             \
  long long wf_prof[8] = {0, 0, 0, 0, 0, 0, 0, 0}, wf_prev = wf_now(); \
  const unsigned long long wf_t0 = wf_ns()
#define XP_WF_CLOCK(i)                \
  {                                   \
    const long long now = wf_now();   \
    wf_prof[i] += now - wf_prev;      \
    wf_prev = now;                    \
  }
#define XP_WF_CLOCK_END(band, nb, fast, slow)                                 \
  if (lane == 0 && (nb) < 200 && (band % 4 == 0 || band == (nb)-1))           \
  printf("band %d/%d fast %d slow %d | t0 %llu t1 %llu dep %lld cpwait %lld " \
         "steps %lld loop %lld\n",                                            \
         band, nb, fast, slow, wf_t0 % 100000000ull, wf_ns() % 100000000ull,  \
         wf_prof[1], wf_prof[2], wf_prof[4], wf_prof[7])
#endif

#include "blend_impl.h"
#include "common.h"
#include "scan_problems.h"
#include "wavefront.h"

namespace xp {

// ---------------------------------------------------------------------------
// device warp context for the wavefront driver
// ---------------------------------------------------------------------------
struct DevCtx {
  __device__ __forceinline__ int lane() const { return threadIdx.x & 31; }
  __device__ __forceinline__ int role() const { return threadIdx.x >> 5; }
  // whole-block operations (all three warps call them together)
  __device__ __forceinline__ int block_bcast(int v) {
    __shared__ int slot;
    if (threadIdx.x == 0) slot = v;
    __syncthreads();
    const int r = slot;
    __syncthreads();
    return r;
  }
  __device__ __forceinline__ void block_sync() { __syncthreads(); }
  // hand-over signals between two warps: named barriers 1..kSignals, 64
  // threads;
  // the signalling warp arrives without waiting, the waiting warp syncs
  __device__ __forceinline__ void signal(int id) {
    __syncwarp();
    asm volatile("bar.arrive %0, 64;" ::"r"(id + 1) : "memory");
  }
  __device__ __forceinline__ void wait(int id) {
    asm volatile("bar.sync %0, 64;" ::"r"(id + 1) : "memory");
  }
  __device__ __forceinline__ int atomic_add(int* p, int v) {
    return atomicAdd(p, v);
  }
  __device__ __forceinline__ int bcast(int v) {
    return __shfl_sync(0xffffffffu, v, 0);
  }
  template <class T>
  __device__ __forceinline__ T shfl_up1(T v) {
    return __shfl_up_sync(0xffffffffu, v, 1);
  }
  template <class T>
  __device__ __forceinline__ T shfl_idx(T v, int src) {
    return __shfl_sync(0xffffffffu, v, src);
  }
  __device__ __forceinline__ void sync() { __syncwarp(); }
  __device__ __forceinline__ bool all(bool v) {
    return __all_sync(0xffffffffu, v);
  }
  __device__ __forceinline__ unsigned ballot(bool v) {
    return __ballot_sync(0xffffffffu, v);
  }
  // polling loads go to L2 every time (values are published by plain stores
  // of whole words; a word that is not the sentinel is final)
  __device__ __forceinline__ u32 ld_poll(const u32* p) {
    u32 v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];"
                 : "=r"(v)
                 : "l"(p)
                 : "memory");
    return v;
  }
  __device__ __forceinline__ u64 ld_poll(const u64* p) {
    u64 v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];"
                 : "=l"(v)
                 : "l"(p)
                 : "memory");
    return v;
  }
  __device__ __forceinline__ void backoff() { __nanosleep(200); }
  __device__ __forceinline__ void cp_async(void* smem_dst, const void* gsrc,
                                           int bytes) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    if (bytes == 4)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d),
                   "l"(gsrc)
                   : "memory");
    else
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d),
                   "l"(gsrc)
                   : "memory");
  }
  __device__ __forceinline__ void cp_commit() {
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  template <int kPending>
  __device__ __forceinline__ void cp_wait_pending() {
    asm volatile("cp.async.wait_group %0;" ::"n"(kPending) : "memory");
  }
};

// Scratch of one wavefront problem: the tile flags of the classification
// pass, the first item of every band (+ the item count) and the item list.
struct WfScratch {
  uint8_t* flags;
  int* first;       // n_bands + 1
  wf::Item* items;  // wf::max_items(n_bands, flag_pitch)
};
static size_t wf_flag_bytes(int n_bands, int fp) {
  return ((size_t)n_bands * fp + 15) & ~(size_t)15;
}
static size_t wf_first_bytes(int n_bands) {
  return (sizeof(int) * ((size_t)n_bands + 1) + 15) & ~(size_t)15;
}
static size_t wf_scratch_bytes(int n_bands, int fp) {
  return wf_flag_bytes(n_bands, fp) + wf_first_bytes(n_bands) +
         ((sizeof(wf::Item) * wf::max_items(n_bands, fp) + 15) & ~(size_t)15);
}
static WfScratch wf_scratch(uint8_t* base, int n_bands, int fp) {
  WfScratch w;
  w.flags = base;
  w.first = reinterpret_cast<int*>(base + wf_flag_bytes(n_bands, fp));
  w.items = reinterpret_cast<wf::Item*>(base + wf_flag_bytes(n_bands, fp) +
                                        wf_first_bytes(n_bands));
  return w;
}

template <class P>
__global__ void __launch_bounds__(32 * wf::kWarpsPerBand)
    wavefront_kernel(P prob, int* counter, const wf::Item* items,
                     const int* n_items, const uint8_t* flags,
                     int flag_pitch) {
  extern __shared__ __align__(16) unsigned char smem[];
  DevCtx ctx;
  wf::run(ctx, prob, counter, items, n_items, flags, flag_pitch, smem);
}

template <class P>
__global__ void __launch_bounds__(128)
    classify_kernel(P prob, uint8_t* flags, int flag_pitch) {
  DevCtx ctx;
  const size_t tile = blockIdx.x * (size_t)(blockDim.x / 32) + threadIdx.x / 32;
  const int band = (int)(tile / flag_pitch);
  if (band >= prob.n_bands) return;
  wf::classify_tile(ctx, prob, band, (int)(tile % flag_pitch), flags,
                    flag_pitch);
}

// Work items of a classified problem (wavefront.h), band-major: a warp per
// band counts the runs of unflagged chunks, one warp turns the counts into
// first-item indices, then a warp per band writes its runs.
template <class P, bool kEmit>
__global__ void __launch_bounds__(256)
    segment_kernel(P prob, const uint8_t* flags, int flag_pitch, int* first,
                   wf::Item* items) {
  DevCtx ctx;
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= prob.n_bands) return;
  const int n = wf::segment_band(ctx, flags + (size_t)b * flag_pitch,
                                 wf::chunks_of(prob.band(b).W), b,
                                 kEmit ? items + first[b] : (wf::Item*)nullptr);
  if (!kEmit && (threadIdx.x & 31) == 0) first[b] = n;
}

__global__ void segment_scan_kernel(int* first, int nb) {
  const int lane = threadIdx.x;
  int carry = 0;
  for (int base = 0; base < nb; base += 32) {
    const int b = base + lane;
    const int v = b < nb ? first[b] : 0;
    int incl = v;
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (b < nb) first[b] = carry + incl - v;
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) first[nb] = carry;
}

// `scratch` needs wavefront_flag_bytes(n_bands, widest band) bytes
template <class P>
static void launch_classify(const P& prob, uint8_t* scratch, int flag_pitch,
                            cudaStream_t s) {
  if (prob.n_bands <= 0) return;
  const WfScratch w = wf_scratch(scratch, prob.n_bands, flag_pitch);
  const size_t tiles = (size_t)prob.n_bands * flag_pitch;
  classify_kernel<P><<<(unsigned)div_up(tiles, (size_t)4), 128, 0, s>>>(
      prob, w.flags, flag_pitch);
  XP_LAUNCHED();
  const int sb = div_up(prob.n_bands, 8);
  segment_kernel<P, false><<<sb, 256, 0, s>>>(prob, w.flags, flag_pitch,
                                              w.first, w.items);
  XP_LAUNCHED();
  segment_scan_kernel<<<1, 32, 0, s>>>(w.first, prob.n_bands);
  XP_LAUNCHED();
  segment_kernel<P, true><<<sb, 256, 0, s>>>(prob, w.flags, flag_pitch,
                                             w.first, w.items);
  XP_LAUNCHED();
}

// Upper bound of the blocks a wavefront launch uses: the SMs of the partition
// its stream runs on (one block per SM, see below). 0: the whole device.
static thread_local int t_wavefront_blocks = 0;
void set_wavefront_blocks(int n) { t_wavefront_blocks = n; }

template <class P>
static void launch_wavefront(const P& prob, int* counter, uint8_t* scratch,
                             int flag_pitch, cudaStream_t s,
                             bool classified = false) {
  if (prob.n_bands <= 0) return;
  XP_CUDA(cudaMemsetAsync(counter, 0, sizeof(int), s));
  if (!classified) launch_classify(prob, scratch, flag_pitch, s);
  const WfScratch w = wf_scratch(scratch, prob.n_bands, flag_pitch);
  const size_t smem = wf::tile_bytes<P>();
  static_assert(wf::kSignals <= 15, "named barriers 1..15");
  // more than 48 KB of dynamic shared memory needs the opt-in (per device)
  XP_CUDA(cudaFuncSetAttribute(wavefront_kernel<P>,
                               cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)smem));
  // The item count lives on the device; blocks without an item leave at
  // once. One block per SM: its three warps then sit on three different
  // schedulers, whereas a second block's loader or storer warp would share a
  // scheduler with the first block's compute warp and slow every step of it.
  // Items are short, so one per SM (148) in flight cover the active front.
  const int blocks = (int)std::min<size_t>(
      wf::max_items(prob.n_bands, flag_pitch),
      (size_t)(t_wavefront_blocks > 0 ? t_wavefront_blocks : sm_count()));
  wavefront_kernel<P><<<blocks, 32 * wf::kWarpsPerBand, smem, s>>>(
      prob, counter, w.items, w.first + prob.n_bands, w.flags, flag_pitch);
  XP_LAUNCHED();
}

// ---------------------------------------------------------------------------
// trim: bbox of alpha == max (image.cpp:396-447; the widening loops make the
// result the exact bounding box of the opaque pixels)
// ---------------------------------------------------------------------------
// all images of a blend in one launch: blockIdx.y = image
struct TrimBatch {
  const void* data[kMaxImages + 1];
  int w[kMaxImages + 1], h[kMaxImages + 1], pitch[kMaxImages + 1];
};

template <int BPP>
__global__ void trim_kernel(const __grid_constant__ TrimBatch batch,
                            int* bbox_all,
                            unsigned long long* opaque_count_all) {
  const void* data = batch.data[blockIdx.y];
  const int w = batch.w[blockIdx.y], h = batch.h[blockIdx.y];
  const int pitch = batch.pitch[blockIdx.y];
  int* bbox = bbox_all + 4 * blockIdx.y;
  unsigned long long* opaque_count = opaque_count_all + blockIdx.y;
  int minx = INT_MAX, miny = INT_MAX, maxx = -1, maxy = -1, cnt = 0;
  // a warp scans 32 consecutive x; rows are strided over the grid
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  for (int y = warp; y < h; y += warps) {
    const size_t row = (size_t)y * pitch;
    bool any = false;
    // eight independent loads per lane in flight
    for (int x0 = lane; x0 < w; x0 += 32 * 8) {
      bool op[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int x = x0 + 32 * k;
        op[k] = x < w && px_opaque<BPP>(data, row + x);
      }
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (op[k]) {
          const int x = x0 + 32 * k;
          minx = min(minx, x);
          maxx = max(maxx, x);
          ++cnt;
          any = true;
        }
    }
    if (any) {
      miny = min(miny, y);
      maxy = max(maxy, y);
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    cnt += __shfl_xor_sync(~0u, cnt, o);
    minx = min(minx, __shfl_xor_sync(~0u, minx, o));
    miny = min(miny, __shfl_xor_sync(~0u, miny, o));
    maxx = max(maxx, __shfl_xor_sync(~0u, maxx, o));
    maxy = max(maxy, __shfl_xor_sync(~0u, maxy, o));
  }
  if (lane == 0 && maxx >= 0) {
    atomicMin(bbox + 0, minx);
    atomicMin(bbox + 1, miny);
    atomicMax(bbox + 2, maxx);
    atomicMax(bbox + 3, maxy);
    atomicAdd(opaque_count, (unsigned long long)cnt);
  }
}

__global__ void trim_init_kernel(int* bbox, unsigned long long* opaque, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  bbox[4 * i] = bbox[4 * i + 1] = INT_MAX;
  bbox[4 * i + 2] = bbox[4 * i + 3] = -1;
  opaque[i] = 0;
}

void launch_trim_init(int* d_bbox, unsigned long long* d_opaque, int n,
                      cudaStream_t s) {
  trim_init_kernel<<<div_up(n, 64), 64, 0, s>>>(d_bbox, d_opaque, n);
  XP_LAUNCHED();
}

// Trim from the opacity bit planes the composer's warps left (warp.cu
// AlphaOut): one bit per pixel. blockIdx.y = image.
struct TrimBits {
  const uint32_t* bits[kMaxImages + 1];
  int pitch[kMaxImages + 1], h[kMaxImages + 1];
};

__global__ void __launch_bounds__(256)
    trim_bits_kernel(const __grid_constant__ TrimBits tb, int* bbox_all,
                     unsigned long long* opaque_count_all) {
  const uint32_t* bits = tb.bits[blockIdx.y];
  const int pitch = tb.pitch[blockIdx.y];
  const int nw = pitch * tb.h[blockIdx.y];
  int minx = INT_MAX, miny = INT_MAX, maxx = -1, maxy = -1;
  unsigned long long cnt = 0;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < nw;
       k += gridDim.x * blockDim.x) {
    const uint32_t v = __ldg(bits + k);
    if (v) {
      const int y = k / pitch, x0 = (k - y * pitch) * 32;
      minx = min(minx, x0 + __ffs((int)v) - 1);
      maxx = max(maxx, x0 + 31 - __clz((int)v));
      miny = min(miny, y);
      maxy = max(maxy, y);
      cnt += (unsigned long long)__popc(v);
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    cnt += __shfl_xor_sync(~0u, cnt, o);
    minx = min(minx, __shfl_xor_sync(~0u, minx, o));
    miny = min(miny, __shfl_xor_sync(~0u, miny, o));
    maxx = max(maxx, __shfl_xor_sync(~0u, maxx, o));
    maxy = max(maxy, __shfl_xor_sync(~0u, maxy, o));
  }
  if ((threadIdx.x & 31) == 0 && maxx >= 0) {
    int* bbox = bbox_all + 4 * blockIdx.y;
    atomicMin(bbox + 0, minx);
    atomicMin(bbox + 1, miny);
    atomicMax(bbox + 2, maxx);
    atomicMax(bbox + 3, maxy);
    atomicAdd(opaque_count_all + blockIdx.y, cnt);
  }
}

// d_bbox / d_opaque_count initialised by launch_trim_init
void launch_trim_bits(const uint32_t* const* bits, const int* pitch,
                      const int* h, int n, int* d_bbox,
                      unsigned long long* d_opaque_count, cudaStream_t s) {
  XP_REQUIRE(n <= kMaxImages, "too many images");
  TrimBits tb;
  int most = 0;
  for (int i = 0; i < n; ++i) {
    tb.bits[i] = bits[i], tb.pitch[i] = pitch[i], tb.h[i] = h[i];
    most = std::max(most, pitch[i] * h[i]);
  }
  dim3 g(std::max(1, std::min(div_up(most, 256 * 8), 64)), n);
  trim_bits_kernel<<<g, 256, 0, s>>>(tb, d_bbox, d_opaque_count);
  XP_LAUNCHED();
}

// trim boxes and opaque counts of n images (d_bbox: 4 ints, d_opaque_count:
// one word per image; both initialised by launch_trim_init)
void launch_trim(const void* const* data, const int* w, const int* h,
                 const int* pitch, int n, int bpp, int* d_bbox,
                 unsigned long long* d_opaque_count, cudaStream_t s) {
  TrimBatch batch;
  int max_h = 0;
  for (int i = 0; i < n; ++i) {
    batch.data[i] = data[i];
    batch.w[i] = w[i];
    batch.h[i] = h[i];
    batch.pitch[i] = pitch[i];
    max_h = std::max(max_h, h[i]);
  }
  // rows are strided over gridDim.x blocks of eight warps
  dim3 g(std::min(div_up(max_h, 8), sm_count() * 4), n);
  if (bpp == 8)
    trim_kernel<8><<<g, 256, 0, s>>>(batch, d_bbox, d_opaque_count);
  else
    trim_kernel<16><<<g, 256, 0, s>>>(batch, d_bbox, d_opaque_count);
  XP_LAUNCHED();
}

// ---------------------------------------------------------------------------
// cover map: bit i = image i opaque here; top bit = a run of the reference's
// RLE walk starts here (bbox edges and opacity changes of any image,
// multiblend.cpp:509-534).
// ---------------------------------------------------------------------------
constexpr int kCoverTileW = 128, kCoverTileH = 8;

template <class C, int F, int BPP>
__global__ void __launch_bounds__(256)
    cover_kernel(C* __restrict__ cover, int W, int H,
                 const ImgDev* __restrict__ imgs, int n, int reverse) {
  // One pass over the canvas, a 128 x 8 tile per block. Every pixel looks at
  // the images whose bounding box (plus the column right of it, where the
  // reference's walk starts a new run) meets the tile; warp 0 compacts that
  // list, in image order. A thread owns the four pixels x0 + 32 j of its row,
  // so every load and store of a warp is one contiguous run, and the pixel
  // left of it -- whose opacity decides whether a run starts here -- is the
  // neighbour lane's.
  extern __shared__ ImgDev s_img[];
  __shared__ int s_count;
  const int lane = threadIdx.x;
  const int bx0 = blockIdx.x * kCoverTileW, by0 = blockIdx.y * kCoverTileH;
  if (threadIdx.y == 0) {
    int m = 0;
    for (int base = 0; base < n; base += 32) {
      const int i = base + lane;
      ImgDev im;
      bool hit = false;
      if (i < n) {
        im = imgs[i];
        hit = bx0 + kCoverTileW - 1 >= im.xpos && bx0 <= im.xpos + im.w &&
              by0 + kCoverTileH - 1 >= im.ypos && by0 < im.ypos + im.h;
      }
      const unsigned hits = __ballot_sync(~0u, hit);
      if (hit) {
        im.tiff_h = i;  // the image's bit position
        s_img[m + __popc(hits & ((1u << lane) - 1u))] = im;
      }
      m += __popc(hits);
    }
    if (lane == 0) s_count = m;
  }
  __syncthreads();
  const int y = by0 + threadIdx.y;
  const int x0 = bx0 + lane;
  C bits[4] = {0, 0, 0, 0};
  bool rs[4] = {false, false, false, false};
  const int m = s_count;
  for (int k = 0; k < m; ++k) {
    const ImgDev& im = s_img[k];
    const int ly = y - im.ypos;
    const bool row_in = y < H && (unsigned)ly < (unsigned)im.h;
    const size_t row = (size_t)(im.top + (row_in ? ly : 0)) * im.tiff_w + im.left;
    const int lx0 = x0 - im.xpos;
    // with a bit plane (composer warps) one word holds 32 pixels' opacity
    const uint32_t* brow =
        im.abits ? im.abits + (size_t)(im.top + (row_in ? ly : 0)) *
                                  im.abits_pitch
                 : nullptr;
    auto opaque_at = [&](int lx) {
      if (brow) {
        const int X = im.left + lx;
        return ((__ldg(brow + (X >> 5)) >> (X & 31)) & 1u) != 0;
      }
      return px_opaque<BPP>(im.data, row + lx);
    };
    bool op[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int lx = lx0 + 32 * j;
      op[j] = row_in && (unsigned)lx < (unsigned)im.w && opaque_at(lx);
    }
    // opacity of the pixel left of the tile's first column
    bool carry = row_in && lane == 0 && lx0 >= 1 && lx0 - 1 < im.w &&
                 opaque_at(lx0 - 1);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int lx = lx0 + 32 * j;
      bool prev = __shfl_up_sync(~0u, op[j], 1);
      if (lane == 0) prev = carry;
      carry = __shfl_sync(~0u, op[j], 31);
      if (row_in && (unsigned)lx <= (unsigned)im.w) {
        if (op[j]) bits[j] |= (C)1 << im.tiff_h;
        if (lx == 0 || lx == im.w || op[j] != prev) rs[j] = true;
      }
    }
  }
  if (y >= H) return;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int x = x0 + 32 * j;
    if (x >= W) break;
    C out;
    if (F == kCoverDigest)
      out = (C)make_digest((u32)bits[j], rs[j], reverse);
    else
      out = bits[j] | (rs[j] ? CoverTraits<C>::rs_bit : (C)0);
    cover[(size_t)y * W + x] = out;
  }
}

// The cover map when every image brought an opacity bit plane (the composer's
// warps, warp.cu AlphaOut): a thread owns four consecutive pixels and takes
// their opacity in image i -- and that of the pixel left of them, which
// decides whether a run starts -- from one funnel shift over two words of the
// plane, instead of five pixel loads. Same tiles, same image list, same
// output as cover_kernel; about a third of its instructions.
template <int F>
__global__ void __launch_bounds__(256)
    cover_bits_kernel(u32* __restrict__ cover, int W, int H,
                      const ImgDev* __restrict__ imgs, int n, int reverse) {
  extern __shared__ ImgDev s_img[];
  __shared__ int s_count;
  const int lane = threadIdx.x;
  const int bx0 = blockIdx.x * kCoverTileW, by0 = blockIdx.y * kCoverTileH;
  if (threadIdx.y == 0) {
    int m = 0;
    for (int base = 0; base < n; base += 32) {
      const int i = base + lane;
      ImgDev im;
      bool hit = false;
      if (i < n) {
        im = imgs[i];
        hit = bx0 + kCoverTileW - 1 >= im.xpos && bx0 <= im.xpos + im.w &&
              by0 + kCoverTileH - 1 >= im.ypos && by0 < im.ypos + im.h;
      }
      const unsigned hits = __ballot_sync(~0u, hit);
      if (hit) {
        im.tiff_h = i;  // the image's bit position
        s_img[m + __popc(hits & ((1u << lane) - 1u))] = im;
      }
      m += __popc(hits);
    }
    if (lane == 0) s_count = m;
  }
  __syncthreads();
  const int y = by0 + threadIdx.y;
  const int x0 = bx0 + 4 * lane;
  if (y >= H || x0 >= W) return;
  u32 bits[4] = {0, 0, 0, 0};
  unsigned rs = 0;  // bit j: a run starts at pixel x0 + j
  const int m = s_count;
  for (int k = 0; k < m; ++k) {
    const ImgDev& im = s_img[k];
    const int ly = y - im.ypos;
    if ((unsigned)ly >= (unsigned)im.h) continue;
    const int lx0 = x0 - im.xpos;      // local x of the first of the four
    if (lx0 + 3 < 0 || lx0 > im.w) continue;
    // five bits, for local x = lx0 - 1 .. lx0 + 3, zero outside [0, w)
    unsigned valid = 0;
#pragma unroll
    for (int t = 0; t < 5; ++t)
      if ((unsigned)(lx0 - 1 + t) < (unsigned)im.w) valid |= 1u << t;
    unsigned five = 0;
    if (valid) {
      const int Xs = im.left + lx0 - 1;  // plane coordinate of bit 0
      const int X0 = max(Xs, 0), sh = X0 - Xs;
      const uint32_t* brow = im.abits + (size_t)(im.top + ly) * im.abits_pitch;
      const int wi = X0 >> 5;
      const u32 lo = __ldg(brow + wi);
      const u32 hi = wi + 1 < im.abits_pitch ? __ldg(brow + wi + 1) : 0u;
      five = ((__funnelshift_r(lo, hi, X0 & 31) << sh) & 31u) & valid;
    }
    const u32 bit = 1u << im.tiff_h;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int lx = lx0 + j;
      if ((unsigned)lx > (unsigned)im.w) continue;  // [0, w]: w = right edge
      const bool op = (five >> (j + 1)) & 1u, prev = (five >> j) & 1u;
      if (op) bits[j] |= bit;
      if (lx == 0 || lx == im.w || op != prev) rs |= 1u << j;
    }
  }
  u32 out[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const bool r = (rs >> j) & 1u;
    out[j] = F == kCoverDigest
                 ? make_digest(bits[j], r, reverse)
                 : bits[j] | (r ? CoverTraits<u32>::rs_bit : 0u);
  }
  u32* dst = cover + (size_t)y * W + x0;
  if (x0 + 3 < W && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
    *reinterpret_cast<uint4*>(dst) = make_uint4(out[0], out[1], out[2], out[3]);
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (x0 + j < W) dst[j] = out[j];
  }
}

int cover_format(int n) {
  return n <= kDigestMaxImages ? kCoverDigest : kCoverBits;
}

void launch_cover(void* cover, int cover_bytes, int W, int H,
                  const ImgDev* d_imgs, int n, int bpp, int reverse,
                  cudaStream_t s, bool with_bits) {
  dim3 b(32, kCoverTileH);
  dim3 g(div_up(W, kCoverTileW), div_up(H, kCoverTileH));
  const size_t sm = sizeof(ImgDev) * (size_t)n;
  if (with_bits && cover_bytes == 4) {
    if (cover_format(n) == kCoverDigest)
      cover_bits_kernel<kCoverDigest><<<g, b, sm, s>>>((u32*)cover, W, H,
                                                       d_imgs, n, reverse);
    else
      cover_bits_kernel<kCoverBits><<<g, b, sm, s>>>((u32*)cover, W, H, d_imgs,
                                                     n, reverse);
    XP_LAUNCHED();
    return;
  }
#define XP_COVER(C, F, BPP)                                        \
  cover_kernel<C, F, BPP><<<g, b, sm, s>>>((C*)cover, W, H, d_imgs, n, \
                                           reverse)
  if (cover_bytes == 8) {
    if (bpp == 8)
      XP_COVER(u64, kCoverBits, 8);
    else
      XP_COVER(u64, kCoverBits, 16);
  } else if (cover_format(n) == kCoverDigest) {
    if (bpp == 8)
      XP_COVER(u32, kCoverDigest, 8);
    else
      XP_COVER(u32, kCoverDigest, 16);
  } else {
    if (bpp == 8)
      XP_COVER(u32, kCoverBits, 8);
    else
      XP_COVER(u32, kCoverBits, 16);
  }
#undef XP_COVER
  XP_LAUNCHED();
}

// ---------------------------------------------------------------------------
// seams
// ---------------------------------------------------------------------------
size_t wavefront_flag_bytes(int n_bands, int max_w) {
  return wf_scratch_bytes(n_bands, wf::chunks_of(max_w));
}

__global__ void seam_unpack_kernel(const u32* __restrict__ in, size_t n,
                                   u64* __restrict__ out) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    out[i] = seam_unpack(in[i]);
}

// debug view: the reference's u64 encoding of the packed seam values
void launch_seam_unpack(const u32* in, size_t n, u64* out, cudaStream_t s) {
  const int blocks = (int)std::min<size_t>(div_up(n, (size_t)256), sm_count() * 16);
  seam_unpack_kernel<<<blocks, 256, 0, s>>>(in, n, out);
  XP_LAUNCHED();
}

template <class C, int F>
static void seams_t(const void* cover, int W, int H, int reverse, u32* bwd,
                    u32* fwd, uint8_t* labels, int* counter, uint8_t* flags,
                    cudaStream_t s, cudaStream_t beside, cudaEvent_t ev) {
  const int n_bands = div_up(H, wf::kLanes);
  const int fp = wf::chunks_of(W);
  // Both classification passes depend on the cover map only (the forward
  // one loads the backward value of each pixel but never looks at it). The
  // forward one leaves the critical chain: it runs on `beside`, which the
  // caller has made wait for the cover map, under the backward wavefront.
  uint8_t* flags_fwd = flags + wf_scratch_bytes(n_bands, fp);
  SeamBwd<C, F> pb{n_bands, W, H, (const C*)cover, bwd};
  SeamFwd<C, F> pf{n_bands, W, H, reverse, (const C*)cover, bwd, fwd, labels};
  launch_classify(pb, flags, fp, s);
  if (beside) {
    launch_classify(pf, flags_fwd, fp, beside);
    XP_CUDA(cudaEventRecord(ev, beside));
  } else {
    launch_classify(pf, flags_fwd, fp, s);
  }
  launch_wavefront(pb, counter, flags, fp, s, true);
  if (beside) XP_CUDA(cudaStreamWaitEvent(s, ev, 0));
  launch_wavefront(pf, counter, flags_fwd, fp, s, true);
}

void launch_seams(const void* cover, int cover_bytes, int n_images, int W,
                  int H, int reverse, u32* bwd, u32* fwd, uint8_t* labels,
                  int* counter, uint8_t* flags, cudaStream_t s,
                  cudaStream_t beside, cudaEvent_t ev) {
  XP_REQUIRE((long)W + H < kSeamMaxExtent,
             "canvas %d x %d exceeds the seam distance range", W, H);
  if (cover_bytes == 8)
    seams_t<u64, kCoverBits>(cover, W, H, reverse, bwd, fwd, labels, counter,
                             flags, s, beside, ev);
  else if (cover_format(n_images) == kCoverDigest)
    seams_t<u32, kCoverDigest>(cover, W, H, reverse, bwd, fwd, labels, counter,
                               flags, s, beside, ev);
  else
    seams_t<u32, kCoverBits>(cover, W, H, reverse, bwd, fwd, labels, counter,
                             flags, s, beside, ev);
}

// ---------------------------------------------------------------------------
// inpaint
// ---------------------------------------------------------------------------
void launch_inpaint(const InpaintJob* d_jobs, const InpaintBandRef* d_bands,
                    int n_bands, int max_w, int bpp, int* counter,
                    uint8_t* flags, cudaStream_t s) {
  const int fp = wf::chunks_of(max_w);
  if (bpp == 8) {
    InpaintFwd<8> pf{n_bands, d_jobs, d_bands};
    launch_wavefront(pf, counter, flags, fp, s);
  } else {
    InpaintFwd<16> pf{n_bands, d_jobs, d_bands};
    launch_wavefront(pf, counter, flags, fp, s);
  }
  InpaintBwd pb{n_bands, d_jobs, d_bands};
  launch_wavefront(pb, counter, flags, fp, s);
}

// Inpaint gather, in place, over the tiles the backward pass did not flag
// (a flagged tile consists of opaque pixels: nothing to fill): every
// non-opaque bbox pixel takes the colour of its source -- always an opaque
// pixel, never written here; alpha is kept, so the cover map and the
// exclusive-pixel sums, which may be read concurrently, are unaffected.
// One warp per tile of the InpaintBwd problem.
template <int BPP>
__global__ void __launch_bounds__(128)
    inpaint_gather_tiles_kernel(InpaintBwd prob, const uint8_t* flags,
                                int flag_pitch) {
  typedef typename PixelOf<BPP>::type Px;
  const size_t tile = blockIdx.x * (size_t)(blockDim.x / 32) + threadIdx.x / 32;
  const int band = (int)(tile / flag_pitch), c = (int)(tile % flag_pitch);
  if (band >= prob.n_bands) return;
  if (flags[(size_t)band * flag_pitch + c] == wf::kTileTrivial) return;
  const wf::Band B = prob.band(band);
  if (c >= wf::chunks_of(B.W)) return;
  const InpaintJob& j = prob.jobs[B.job];
  Px* data = static_cast<Px*>(const_cast<void*>(j.data));
  const int lane = threadIdx.x & 31;
  const int ab = BPP == 8 ? 24 : 48;
  const Px amask = (Px)(BPP == 8 ? 0xffull : 0xffffull) << ab;
  for (int r = 0; r < wf::kLanes; ++r) {
    const int rho = B.rho0 + r, q = c * wf::kChunk - wf::kSkew * r + lane;
    if (rho >= B.H || q < 0 || q >= B.W) continue;
    const int x = B.W - 1 - q, y = B.H - 1 - rho;
    const size_t me = (size_t)(j.top + y) * j.tiff_w + j.left + x;
    const Px mine = data[me];
    if (pixel_opaque(mine)) continue;
    const u32 from = (u32)j.bwd[(size_t)y * B.W + x];
    if (from == (u32)me) continue;
    data[me] = (data[from] & ~amask) | (mine & amask);
  }
}

void launch_inpaint_gather_tiles(const InpaintJob* d_jobs,
                                 const InpaintBandRef* d_bands, int n_bands,
                                 int max_w, int bpp, const uint8_t* scratch,
                                 cudaStream_t s) {
  if (n_bands <= 0) return;
  const int fp = wf::chunks_of(max_w);
  InpaintBwd pb{n_bands, d_jobs, d_bands};
  const WfScratch w = wf_scratch(const_cast<uint8_t*>(scratch), n_bands, fp);
  const size_t tiles = (size_t)n_bands * fp;
  const unsigned g = (unsigned)div_up(tiles, (size_t)4);
  if (bpp == 8)
    inpaint_gather_tiles_kernel<8><<<g, 128, 0, s>>>(pb, w.flags, fp);
  else
    inpaint_gather_tiles_kernel<16><<<g, 128, 0, s>>>(pb, w.flags, fp);
  XP_LAUNCHED();
}

// ---------------------------------------------------------------------------
// canvas statistics (multiblend.cpp:536-616, 821-823): full / xor masks,
// per-channel sums of the inputs over exclusively covered pixels, whether any
// pixel is uncovered, and the bounding box of every image's owner region.
// ---------------------------------------------------------------------------
// kMode bit 0: masks, exclusive-pixel sums, uncovered flag (needs the cover
// map only); bit 1: owner-region boxes (needs the labels).
template <class C, int F, int BPP, int kMode>
__global__ void canvas_stats_kernel(const C* __restrict__ cover,
                                    const uint8_t* __restrict__ labels, int W,
                                    int H, const ImgDev* __restrict__ imgs,
                                    int n, uint8_t* __restrict__ full_mask,
                                    uint8_t* __restrict__ xor_mask,
                                    CanvasStats* stats) {
  __shared__ int s_box[64][4];
  __shared__ unsigned long long s_sum[4];
  __shared__ int s_uncovered;
  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  for (int i = tid; i < 64; i += blockDim.x * blockDim.y) {
    s_box[i][0] = INT_MAX;
    s_box[i][1] = INT_MAX;
    s_box[i][2] = -1;
    s_box[i][3] = -1;
  }
  if (tid < 4) s_sum[tid] = 0;
  if (tid == 0) s_uncovered = 0;
  __syncthreads();
  // persistent blocks: tiles of 32 x 8 pixels are strided over the grid and
  // the block's partial results leave through a handful of global atomics at
  // the end (one block per tile would serialise ~10^6 atomics on 36 words)
  const int tiles_x = (W + 31) / 32, tiles_y = (H + 7) / 8;
  for (int tile = blockIdx.x; tile < tiles_x * tiles_y; tile += gridDim.x) {
    const int x = (tile % tiles_x) * 32 + threadIdx.x;
    const int y = (tile / tiles_x) * 8 + threadIdx.y;
    unsigned sum[4] = {0, 0, 0, 0};
    int label = -1;
    if (x < W && y < H && (kMode & 1)) {
      const size_t p = (size_t)y * W + x;
      typedef Cov<C, F> Cv;
      const C cov = cover[p];
      const bool any = Cv::any(cov), one = Cv::seed(cov);
      full_mask[p] = any ? 255 : 0;
      xor_mask[p] = one ? 1 : 0;
      if (!any) s_uncovered = 1;
      if (one) {
        const ImgDev im = imgs[Cv::low(cov)];
        const size_t q =
            (size_t)(im.top + y - im.ypos) * im.tiff_w + im.left + x - im.xpos;
        if (BPP == 8) {
          const u32 v = static_cast<const u32*>(im.data)[q];
          sum[0] = v & 0xff;
          sum[1] = (v >> 8) & 0xff;
          sum[2] = (v >> 16) & 0xff;
        } else {
          const u64 v = static_cast<const u64*>(im.data)[q];
          sum[0] = v & 0xffff;
          sum[1] = (v >> 16) & 0xffff;
          sum[2] = (v >> 32) & 0xffff;
        }
        sum[3] = 1;
      }
    }
    if (x < W && y < H && (kMode & 2) && n > 1)
      label = labels[(size_t)y * W + x];
    // a warp is 32 consecutive x of one row: lanes with the same label elect
    // one of them to extend that label's box
    if (kMode & 2) {
      const unsigned peers = __match_any_sync(~0u, label);
      const int lane = tid & 31;
      if (label >= 0 && lane == __ffs(peers) - 1) {
        const int xb = x - lane;
        atomicMin(&s_box[label][0], xb + __ffs(peers) - 1);
        atomicMin(&s_box[label][1], y);
        atomicMax(&s_box[label][2], xb + 31 - __clz(peers));
        atomicMax(&s_box[label][3], y);
      }
    }
    for (int k = 0; k < ((kMode & 1) ? 4 : 0); ++k) {
      unsigned v = sum[k];  // <= 32 * 65535
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(~0u, v, o);
      if ((tid & 31) == 0 && v) atomicAdd(&s_sum[k], (unsigned long long)v);
    }
  }
  __syncthreads();
  if (tid < 4 && s_sum[tid]) atomicAdd(&stats->totals[tid], s_sum[tid]);
  if (tid == 0 && s_uncovered) stats->uncovered = 1;
  for (int i = tid; i < n; i += blockDim.x * blockDim.y) {
    if (s_box[i][2] >= 0) {
      atomicMin(&stats->box[i][0], s_box[i][0]);
      atomicMin(&stats->box[i][1], s_box[i][1]);
      atomicMax(&stats->box[i][2], s_box[i][2]);
      atomicMax(&stats->box[i][3], s_box[i][3]);
    }
  }
}

// The cover-only part (kMode 1) as a streaming pass: four consecutive pixels
// of the linear canvas per thread (one 16/32-byte load of cover words, one
// 4-byte store to each mask), sums kept in registers over the whole grid
// stride and reduced once per block.
template <class C, int F, int BPP>
__global__ void __launch_bounds__(256)
    canvas_masks_kernel(const C* __restrict__ cover, int W, size_t n_px,
                        const ImgDev* __restrict__ imgs, int n,
                        uint8_t* __restrict__ full_mask,
                        uint8_t* __restrict__ xor_mask, CanvasStats* stats,
                        int sum_x0, int sum_x1) {
  // sums are taken over the columns [sum_x0, sum_x1) only: the whole canvas,
  // or the strip a rank of a canvas-tiled run owns
  typedef Cov<C, F> Cv;
  __shared__ ImgDev s_img[64];
  __shared__ unsigned long long s_sum[4];
  __shared__ int s_uncovered;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s_img[i] = imgs[i];
  if (threadIdx.x < 4) s_sum[threadIdx.x] = 0;
  if (threadIdx.x == 0) s_uncovered = 0;
  __syncthreads();
  unsigned long long sum[4] = {0, 0, 0, 0};
  bool uncovered = false;
  const size_t groups = (n_px + 3) / 4;
  for (size_t g = blockIdx.x * (size_t)blockDim.x + threadIdx.x; g < groups;
       g += (size_t)gridDim.x * blockDim.x) {
    const size_t p0 = 4 * g;
    C cv[4];
    if (p0 + 3 < n_px) {
      if (sizeof(C) == 4) {
        const uint4 v = *reinterpret_cast<const uint4*>(cover + p0);
        cv[0] = v.x, cv[1] = v.y, cv[2] = v.z, cv[3] = v.w;
      } else {
        const ulonglong2 a = *reinterpret_cast<const ulonglong2*>(cover + p0);
        const ulonglong2 b =
            *reinterpret_cast<const ulonglong2*>(cover + p0 + 2);
        cv[0] = (C)a.x, cv[1] = (C)a.y, cv[2] = (C)b.x, cv[3] = (C)b.y;
      }
    } else {
      for (int k = 0; k < 4; ++k) cv[k] = p0 + k < n_px ? cover[p0 + k] : (C)0;
    }
    unsigned fm = 0, xm = 0;
    int y = (int)(p0 / W), x = (int)(p0 - (size_t)y * W);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const bool in = p0 + k < n_px;
      const bool any = Cv::any(cv[k]), one = Cv::seed(cv[k]);
      fm |= any ? 255u << (8 * k) : 0u;
      xm |= one ? 1u << (8 * k) : 0u;
      uncovered = uncovered || (in && !any);
      if (in && one && x >= sum_x0 && x < sum_x1) {
        const ImgDev& im = s_img[Cv::low(cv[k])];
        const size_t q =
            (size_t)(im.top + y - im.ypos) * im.tiff_w + im.left + x - im.xpos;
        if (BPP == 8) {
          const u32 v = static_cast<const u32*>(im.data)[q];
          sum[0] += v & 0xff;
          sum[1] += (v >> 8) & 0xff;
          sum[2] += (v >> 16) & 0xff;
        } else {
          const u64 v = static_cast<const u64*>(im.data)[q];
          sum[0] += v & 0xffff;
          sum[1] += (v >> 16) & 0xffff;
          sum[2] += (v >> 32) & 0xffff;
        }
        sum[3] += 1;
      }
      if (++x == W) x = 0, ++y;
    }
    if (p0 + 3 < n_px) {
      *reinterpret_cast<unsigned*>(full_mask + p0) = fm;
      *reinterpret_cast<unsigned*>(xor_mask + p0) = xm;
    } else {
      for (int k = 0; p0 + k < n_px; ++k) {
        full_mask[p0 + k] = (uint8_t)(fm >> (8 * k));
        xor_mask[p0 + k] = (uint8_t)(xm >> (8 * k));
      }
    }
  }
  for (int k = 0; k < 4; ++k) {
    unsigned long long v = sum[k];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(~0u, v, o);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(&s_sum[k], v);
  }
  if (uncovered) s_uncovered = 1;
  __syncthreads();
  if (threadIdx.x < 4 && s_sum[threadIdx.x])
    atomicAdd(&stats->totals[threadIdx.x], s_sum[threadIdx.x]);
  if (threadIdx.x == 0 && s_uncovered) stats->uncovered = 1;
}

// Owner-region boxes (kMode 2) as a streaming pass over the label map: four
// consecutive pixels of the linear canvas per thread. Nearly every thread
// sees one label; the lanes of a warp that share (row, label) elect their
// lowest lane, which extends that label's box by the span of all of them.
__global__ void __launch_bounds__(256)
    label_boxes_kernel(const uint8_t* __restrict__ labels, int W, size_t n_px,
                       int n, CanvasStats* stats) {
  __shared__ int s_box[64][4];
  for (int i = threadIdx.x; i < 64; i += blockDim.x) {
    s_box[i][0] = s_box[i][1] = INT_MAX;
    s_box[i][2] = s_box[i][3] = -1;
  }
  __syncthreads();
  auto extend = [&](int label, int xa, int xb, int y) {
    if (label >= 64) return;
    atomicMin(&s_box[label][0], xa);
    atomicMin(&s_box[label][1], y);
    atomicMax(&s_box[label][2], xb);
    atomicMax(&s_box[label][3], y);
  };
  const int lane = threadIdx.x & 31;
  const size_t groups = (n_px + 3) / 4;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  // whole warps iterate together (the vote below needs every lane)
  const size_t rounds = (groups + stride - 1) / stride;
  for (size_t r = 0; r < rounds; ++r) {
    const size_t g = r * stride + blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const size_t p0 = 4 * g;
    unsigned key = 0x80000000u | (unsigned)lane;  // no peers
    int x = 0, y = 0;
    if (g < groups) {
      y = (int)(p0 / W);
      x = (int)(p0 - (size_t)y * W);
      if (p0 + 3 < n_px && x + 3 < W) {
        const unsigned v = *reinterpret_cast<const unsigned*>(labels + p0);
        if (v == (v & 0xffu) * 0x01010101u) {
          key = ((unsigned)y << 6) | (v & 63u);
          if ((v & 0xffu) >= 64u) key = 0x80000000u | (unsigned)lane;
        } else {
          for (int k = 0; k < 4; ++k)
            extend((int)((v >> (8 * k)) & 0xffu), x + k, x + k, y);
        }
      } else {
        int xx = x, yy = y;
        for (int k = 0; k < 4 && p0 + k < n_px; ++k) {
          extend(labels[p0 + k], xx, xx, yy);
          if (++xx == W) xx = 0, ++yy;
        }
      }
    }
    const unsigned peers = __match_any_sync(~0u, key);
    const int hi_lane = 31 - __clz(peers);
    const int x_hi = __shfl_sync(~0u, x + 3, hi_lane);
    if (!(key & 0x80000000u) && lane == __ffs(peers) - 1)
      extend((int)(key & 63u), x, x_hi, y);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (s_box[i][2] >= 0) {
      atomicMin(&stats->box[i][0], s_box[i][0]);
      atomicMin(&stats->box[i][1], s_box[i][1]);
      atomicMax(&stats->box[i][2], s_box[i][2]);
      atomicMax(&stats->box[i][3], s_box[i][3]);
    }
  }
}

__global__ void window_copy_kernel(const uint8_t* __restrict__ labels,
                                   const uint8_t* __restrict__ xor_mask,
                                   const uint8_t* __restrict__ full_mask,
                                   int W, int H, int wx0, int ww, int ox0,
                                   int ox1,
                                   const uint8_t* __restrict__ relabel,
                                   uint8_t* __restrict__ wlabels,
                                   uint8_t* __restrict__ wxor,
                                   uint8_t* __restrict__ wfull) {
  __shared__ uint8_t map[64];
  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  if (tid < 64) map[tid] = relabel[tid];
  __syncthreads();
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= ww || y >= H) return;
  const size_t src = (size_t)y * W + wx0 + x, dst = (size_t)y * ww + x;
  const uint8_t l = labels[src];
  wlabels[dst] = l < 64 ? map[l] : (uint8_t)255;
  const int gx = wx0 + x;
  wxor[dst] = gx >= ox0 && gx < ox1 ? xor_mask[src] : (uint8_t)0;
  wfull[dst] = full_mask[src];
}

void launch_window_copy(const uint8_t* labels, const uint8_t* xor_mask,
                        const uint8_t* full_mask, int W, int H, int wx0,
                        int ww, int ox0, int ox1, const uint8_t* d_relabel,
                        uint8_t* wlabels, uint8_t* wxor, uint8_t* wfull,
                        cudaStream_t s) {
  dim3 b(64, 4);
  dim3 g(div_up(ww, 64), div_up(H, 4));
  window_copy_kernel<<<g, b, 0, s>>>(labels, xor_mask, full_mask, W, H, wx0,
                                     ww, ox0, ox1, d_relabel, wlabels, wxor,
                                     wfull);
  XP_LAUNCHED();
}

__global__ void stats_init_kernel(CanvasStats* st) {
  const int i = threadIdx.x;
  if (i < 4) st->totals[i] = 0;
  if (i == 0) st->uncovered = 0, st->pad = 0;
  st->box[i][0] = st->box[i][1] = INT_MAX;
  st->box[i][2] = st->box[i][3] = -1;
}

void launch_stats_init(CanvasStats* d_stats, cudaStream_t s) {
  stats_init_kernel<<<1, 64, 0, s>>>(d_stats);
  XP_LAUNCHED();
}

void launch_canvas_stats(const void* cover, int cover_bytes,
                         const uint8_t* labels, int W, int H,
                         const ImgDev* d_imgs, int n, int bpp,
                         uint8_t* full_mask, uint8_t* xor_mask,
                         CanvasStats* d_stats, int mode, cudaStream_t s,
                         int sum_x0, int sum_x1) {
  if (mode == 2) {  // needs the labels only
    if (n > 1) {
      label_boxes_kernel<<<sm_count() * 8, 256, 0, s>>>(labels, W, (size_t)W * H, n,
                                                 d_stats);
      XP_LAUNCHED();
    }
    return;
  }
  dim3 b(32, 8);
  const int g = std::min(div_up(W, 32) * div_up(H, 8), sm_count() * 8);
#define XP_STATS3(C, F, BPP, M)                                               \
  canvas_stats_kernel<C, F, BPP, M><<<g, b, 0, s>>>((const C*)cover, labels, \
                                                    W, H, d_imgs, n,         \
                                                    full_mask, xor_mask,     \
                                                    d_stats)
#define XP_STATS(C, F, BPP)                                                 \
  if (mode == 1)                                                            \
    canvas_masks_kernel<C, F, BPP><<<sm_count() * 8, 256, 0, s>>>(                 \
        (const C*)cover, W, (size_t)W * H, d_imgs, n, full_mask, xor_mask, \
        d_stats, sum_x0, sum_x1);                                           \
  else if (mode == 2)          \
    XP_STATS3(C, F, BPP, 2);   \
  else                         \
    XP_STATS3(C, F, BPP, 3)
  if (cover_bytes == 8) {
    if (bpp == 8) {
      XP_STATS(u64, kCoverBits, 8);
    } else {
      XP_STATS(u64, kCoverBits, 16);
    }
  } else if (cover_format(n) == kCoverDigest) {
    if (bpp == 8) {
      XP_STATS(u32, kCoverDigest, 8);
    } else {
      XP_STATS(u32, kCoverDigest, 16);
    }
  } else {
    if (bpp == 8) {
      XP_STATS(u32, kCoverBits, 8);
    } else {
      XP_STATS(u32, kCoverBits, 16);
    }
  }
#undef XP_STATS
#undef XP_STATS3
  XP_LAUNCHED();
}

}  // namespace xp
